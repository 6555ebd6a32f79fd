"""Exact-difference body force (xlb/operator/force/exact_difference_force.py:14-76; Kupershtokh 2004).

Inside ``IncompressibleNavierStokesStepper`` the force is applied by the fused step kernel (``xlb_step_ex`` with
``XlbCollisionParams.has_force``).  This stand-alone operator keeps the reference's call signature
``force(f_postcollision, feq, rho, u) -> f_postcollision`` for scripts that compose operators by hand; it runs the
equilibrium kernel once and two element-wise device ops, and is not on the timed path.
"""

import torch

from ...compute_backend import ComputeBackend
from ..equilibrium import QuadraticEquilibrium
from ..operator import Operator


class ExactDifference(Operator):
    def __init__(self, force_vector, velocity_set=None, precision_policy=None, compute_backend=None):
        # like the reference: one constant force vector, not a spatially varying field (exact_difference_force.py:34)
        self.force_vector = force_vector
        self.equilibrium = QuadraticEquilibrium(compute_backend=compute_backend)
        super().__init__(velocity_set, precision_policy, compute_backend)

    @Operator.register_backend(ComputeBackend.CUDA)
    def cuda_implementation(self, f_postcollision, feq, rho, u):
        d = self.velocity_set.d
        force = torch.as_tensor(self.force_vector).to(device=u.device, dtype=u.dtype).reshape((d,) + (1,) * (u.dim() - 1))
        feq_force = self.equilibrium(rho, (u + force).contiguous())  # exact_difference_force.py:72-73
        f_postcollision += feq_force.to(f_postcollision.dtype) - feq  # :74
        return f_postcollision
