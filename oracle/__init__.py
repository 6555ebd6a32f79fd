"""CPU oracle for the XLB per-timestep LBM update  --  TEST INFRASTRUCTURE ONLY.

This package is a NumPy restatement of the *JAX* code path of the reference
(hsalehipour/XLB): ``IncompressibleNavierStokesStepper.jax_implementation_pull``
(xlb/operator/stepper/nse_stepper.py:238-282) and every operator it composes.
Each function cites the reference file:line it follows.

Rules (enforced by tests/test_no_oracle_in_product.py):
  * only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
    ``--impl reference`` legs may import anything from here;
  * the product package ``xlb_b200`` never imports it and has no CPU fallback.

Pinning status
--------------
The reference cannot be imported in the build container (``jax`` / ``warp`` are not
installed and every ``xlb`` module imports them at top level, xlb/operator/operator.py:17-18),
and its own tests hold **no stored golden vectors** -- only analytic identities
(tests/kernels/*, tests/boundary_conditions/*).  Those identities are restated against
this oracle in ``tests/test_oracle_*.py``.  They pin: QuadraticEquilibrium, Macroscopic,
BGK, Stream, EquilibriumBC, FullwayBounceBackBC and the ``bc_mask`` output of
IndicesBoundaryMasker.  For KBC, HalfwayBounceBackBC, ZouHeBC, ExtrapolationOutflowBC,
the ``missing_mask`` contents, the composed stepper, FP64 / FP16 policies and
``distribute`` the reference has no tests at all: for those rows **parity is unpinned**
(the restatement is the only pin).

Third-party arithmetic outside /root/reference: ``jax>=0.8.2`` (setup.py:95, unpinned,
not vendored) supplies jnp.roll / jnp.sum / jnp.tensordot / jnp.where / lax.ppermute;
their semantics are exact data movement or IEEE ops, but XLA's reduction order and FMA
contraction are unspecified, hence the FP tolerances (1e-5 rel FP32, 1e-12 FP64).
"""

from .lattice import Lattice, D2Q9, D3Q19, D3Q27, make_lattice  # noqa: F401
