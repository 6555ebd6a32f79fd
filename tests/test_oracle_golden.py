"""Pins the NumPy oracle to the reference itself: tests/golden/*.npz are outputs of the UNMODIFIED reference source
(hsalehipour/XLB, JAX code path) executed under oracle/refshim by ``python -m oracle.gen_golden``.
Masks must be bit-exact; populations agree to far below the parity tolerance (they are in fact bit-identical in
this container -- both sides are NumPy -- the slack only absorbs BLAS / NumPy version differences)."""

import numpy as np
import pytest

from oracle import cases
from tests.golden_util import golden_cases, index, load
from tests.util import rel_err

CASES = golden_cases()


def test_index_lists_every_case():
    idx = index()["cases"]
    assert sorted(idx) == sorted(c["name"] for c, _ in CASES)


@pytest.mark.parametrize("case,n_steps", CASES, ids=lambda v: v["name"] if isinstance(v, dict) else str(v))
def test_oracle_matches_reference_run(case, n_steps):
    g = load(case)
    assert g["n_steps"] == n_steps
    ost, _ = cases.build_oracle(case)
    f0, f1, bc, mm = ost.prepare_fields()
    assert bc.dtype == g["bc_mask"].dtype and np.array_equal(bc, g["bc_mask"])
    assert np.array_equal(mm, g["missing_mask"])
    assert np.array_equal(f0.astype(g["f_init"].dtype), g["f_init"])
    tol = 1e-13 if case["policy"].startswith("FP64") else 1e-7
    _, one = ost(cases.random_state(case, seed=0), None, bc, mm, case["omega"], 0)
    assert one.dtype == g["f_one"].dtype
    assert rel_err(one, g["f_one"]) <= (2.0**-11 if case["policy"].endswith("FP16") else tol)
    fn, _ = ost.run(f0.astype(ost.sdt), f1, bc, mm, case["omega"], n_steps)
    assert rel_err(fn, g["f_n"]) <= (2.0**-9 if case["policy"].endswith("FP16") else 50 * tol)


@pytest.mark.parametrize("case,n_steps", [c for c in CASES if any(b["kind"] == "HalfwayBounceBackBC" for b in c[0]["bcs"])], ids=lambda v: v["name"] if isinstance(v, dict) else str(v))
def test_oracle_momentum_transfer_matches_reference_run(case, n_steps):
    """oracle/force.py against MomentumTransfer of the unmodified reference (momentum_transfer.py:174-211) on the
    reference's own populations after n steps."""
    from oracle import bc as obc
    from oracle.force import momentum_transfer

    g = load(case)
    ost, _ = cases.build_oracle(case)
    solid = [b for b in ost.boundary_conditions if isinstance(b, obc.HalfwayBounceBackBC)][-1]
    force = momentum_transfer(ost.lat, solid, g["f_n"], g["bc_mask"], g["missing_mask"])
    assert force.shape == g["force_n"].shape == (ost.lat.d,)
    assert force.dtype == g["force_n"].dtype
    assert np.array_equal(force, g["force_n"])
