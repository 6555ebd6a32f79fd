"""Multi-threaded CPU transcription of the oracle step for the benchmark cavity (TEST / BASELINE INFRASTRUCTURE).

The reference's CPU path is its JAX backend on XLA:CPU, which runs whole-array ops on all host cores; NumPy (the
oracle proper) is single-threaded for these ops.  This module re-states the SAME array program as
``oracle/stepper.py`` for the BCs of ``examples/performance/mlups_3d.py`` (EquilibriumBC lid + FullwayBounceBackBC
walls, BGK, any lattice / policy) with ``torch`` CPU tensors so that ``bench.py``'s reference arm can use every host
thread.  ``tests/test_oracle_identities.py`` checks it against the NumPy oracle.  It is never imported by ``xlb_b200``.
"""

import numpy as np
import torch

from . import ops
from .bc import COLLISION, STREAMING, EquilibriumBC, FullwayBounceBackBC


class TorchCavityStepper:
    """Same operation order as ``Stepper.__call__`` (nse_stepper.py:238-282) on torch CPU tensors."""

    def __init__(self, oracle_stepper):
        st = oracle_stepper
        if st.collision_type != "BGK":
            raise NotImplementedError("torch CPU transcription: BGK only (the mlups_3d.py benchmark)")
        for bc in st.boundary_conditions:
            if not isinstance(bc, (EquilibriumBC, FullwayBounceBackBC)):
                raise NotImplementedError("torch CPU transcription: EquilibriumBC and FullwayBounceBackBC only")
        self.st = st
        self.lat = st.lat
        tdt = {np.dtype(np.float32): torch.float32, np.dtype(np.float64): torch.float64, np.dtype(np.float16): torch.float16}
        self.cdt, self.sdt = tdt[st.cdt], tdt[st.sdt]
        d = self.lat.d
        self.c = torch.tensor(self.lat.c.astype(np.float64), dtype=self.cdt)  # (d, q)
        self.w = torch.tensor(self.lat.w, dtype=self.cdt).reshape((-1,) + (1,) * d)
        self.opp = torch.tensor(self.lat.opp_indices, dtype=torch.long)
        self.bc_consts = {}
        for bc in st.boundary_conditions:
            if isinstance(bc, EquilibriumBC):
                self.bc_consts[bc.id] = torch.tensor(bc.feq_constants().astype(np.float64), dtype=self.cdt).reshape((-1,) + (1,) * d)

    def __call__(self, f_0, f_1, bc_mask, missing_mask, omega, timestep=0):
        lat = self.lat
        d = lat.d
        f0 = f_0.to(self.cdt)
        dims = tuple(range(d))
        fps = torch.empty_like(f0)
        for l in range(lat.q):  # stream.py:29-62
            fps[l] = torch.roll(f0[l], tuple(int(v) for v in lat.c[:, l]), dims)
        for bc in self.st.boundary_conditions:  # nse_stepper.py:250-257
            if bc.implementation_step == STREAMING:
                fps = torch.where(bc_mask == bc.id, self.bc_consts[bc.id], fps)
        rho = fps.sum(dim=0, keepdim=True)  # zero_moment.py:14-17
        u = torch.tensordot(self.c, fps, dims=([1], [0])) / rho  # first_moment.py:14-18
        cu = 3.0 * torch.tensordot(self.c, u, dims=([0], [0]))  # quadratic_equilibrium.py:23-30
        usqr = 1.5 * (u * u).sum(dim=0, keepdim=True)
        feq = rho * self.w * (1.0 + cu * (1.0 + 0.5 * cu) - usqr)
        fpc = fps - omega * (fps - feq)  # bgk.py:27-32
        for bc in self.st.boundary_conditions:  # nse_stepper.py:269-277
            if bc.implementation_step == COLLISION:
                fpc = torch.where(bc_mask == bc.id, fps[self.opp], fpc)
        return f_0, fpc.to(self.sdt)


def to_torch(a):
    return torch.from_numpy(np.ascontiguousarray(a))
