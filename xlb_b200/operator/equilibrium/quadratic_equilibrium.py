"""Second-order Hermite equilibrium (xlb/operator/equilibrium/quadratic_equilibrium.py:23-30)."""

from ...compute_backend import ComputeBackend
from ...default_config import DefaultConfig
from ..operator import Operator, byref, check, lib, stream_ptr
from .equilibrium import Equilibrium


class QuadraticEquilibrium(Equilibrium):
    @Operator.register_backend(ComputeBackend.CUDA)
    def cuda_implementation(self, rho, u, f):
        """Out-parameter form (quadratic_equilibrium.py:89-99): fills and returns ``f``."""
        vs = self.velocity_set
        self._check_field(rho, "rho", 1)
        self._check_field(u, "u", vs.d)
        self._check_field(f, "f", vs.q)
        if rho.dtype != u.dtype:
            raise ValueError("QuadraticEquilibrium: rho and u must share a dtype")
        dom = self._domain(f)
        check(lib().xlb_equilibrium(byref(dom), self._code(rho), rho.data_ptr(), u.data_ptr(), self._code(f), f.data_ptr(), stream_ptr()), "QuadraticEquilibrium")
        return f

    @Operator.register_backend(ComputeBackend.CUDA)
    def cuda_functional(self, rho, u):
        """Functional form (quadratic_equilibrium.py:23-30): returns feq in the compute dtype."""
        import torch

        f = torch.empty((self.velocity_set.q,) + tuple(rho.shape[1:]), dtype=self.compute_dtype, device=rho.device)
        return self.cuda_implementation(rho, u, f)
