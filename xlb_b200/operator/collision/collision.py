"""Collision operators (xlb/operator/collision/{collision,bgk,kbc,smagorinsky_les_bgk,forced_collision}.py)."""

import torch

from ... import _lib
from ...compute_backend import ComputeBackend
from ...velocity_set import D2Q9, D3Q27
from ..operator import Operator, byref, check, lib, stream_ptr


class Collision(Operator):
    """Base class: ``collision(f, feq, omega) -> fout`` or ``collision(f, feq, fout, omega)``."""

    abi_model = None

    def abi_params(self):
        """``XlbCollisionParams`` of this collision for ``xlb_step_ex`` / ``xlb_collide_ex`` (None = defaults)."""
        return None

    def _run(self, f, feq, fout, omega):
        q = self.velocity_set.q
        for name, t in (("f", f), ("feq", feq), ("fout", fout)):
            self._check_field(t, name, q)
        if not (f.dtype == feq.dtype == fout.dtype):
            raise ValueError("collision: f, feq and fout must share a dtype")
        dom = self._domain(f)
        cp = self.abi_params()
        check(
            lib().xlb_collide_ex(
                byref(dom), self.abi_model, byref(cp) if cp is not None else None, self._code(f), f.data_ptr(), feq.data_ptr(), fout.data_ptr(),
                float(omega), stream_ptr(),
            ),
            self.__class__.__name__,
        )
        return fout


class BGK(Collision):
    """fout = f - omega (f - feq)  (bgk.py:27-32)."""

    abi_model = _lib.BGK

    @Operator.register_backend(ComputeBackend.CUDA)
    def cuda_implementation(self, f, feq, fout, omega):
        return self._run(f, feq, fout, omega)

    @Operator.register_backend(ComputeBackend.CUDA)
    def cuda_functional(self, f, feq, omega):
        return self._run(f, feq, torch.empty_like(f), omega)


class KBC(Collision):
    """Entropic two-relaxation-time KBC (kbc.py:40-174); D2Q9 and D3Q27 only, like the reference."""

    abi_model = _lib.KBC

    def __init__(self, velocity_set=None, precision_policy=None, compute_backend=None):
        self.epsilon = 1e-32
        super().__init__(velocity_set=velocity_set, precision_policy=precision_policy, compute_backend=compute_backend)

    def _construct_cuda(self):
        if not isinstance(self.velocity_set, (D2Q9, D3Q27)):
            raise NotImplementedError("Velocity set not supported for KBC: {}".format(type(self.velocity_set)))  # kbc.py:65-66, 178-179

    @Operator.register_backend(ComputeBackend.CUDA)
    def cuda_implementation(self, f, feq, fout, omega):
        return self._run(f, feq, fout, omega)

    @Operator.register_backend(ComputeBackend.CUDA)
    def cuda_functional(self, f, feq, omega):
        return self._run(f, feq, torch.empty_like(f), omega)


class SmagorinskyLESBGK(Collision):
    """BGK with the Smagorinsky sub-grid model (smagorinsky_les_bgk.py:18-66): the relaxation time is raised per
    cell from the non-equilibrium stress, ``tau = (tau0 + sqrt(tau0^2 + 36 cs^2 sqrt(strain))) / 2``."""

    abi_model = _lib.SMAGORINSKY_LES_BGK

    def __init__(self, velocity_set=None, precision_policy=None, compute_backend=None, smagorinsky_coef: float = 0.17):
        self.smagorinsky_coef = smagorinsky_coef
        super().__init__(velocity_set=velocity_set, precision_policy=precision_policy, compute_backend=compute_backend)

    def abi_params(self):
        cp = _lib.XlbCollisionParams()
        cp.smagorinsky_coef = float(self.smagorinsky_coef)
        return cp

    @Operator.register_backend(ComputeBackend.CUDA)
    def cuda_implementation(self, f, feq, fout, omega):
        return self._run(f, feq, fout, omega)

    @Operator.register_backend(ComputeBackend.CUDA)
    def cuda_functional(self, f, feq, omega):
        return self._run(f, feq, torch.empty_like(f), omega)


class ForcedCollision(Collision):
    """Inner collision followed by a body-force correction (forced_collision.py:18-57).  Inside the fused stepper
    the whole chain runs in the step kernel (``XlbCollisionParams.has_force``); called stand-alone it composes the
    inner collision, ``Macroscopic`` and ``ExactDifference`` like the reference does."""

    def __init__(self, collision_operator, forcing_scheme="exact_difference", force_vector=None):
        from ..force import ExactDifference
        from ..macroscopic import Macroscopic

        assert collision_operator is not None
        self.collision_operator = collision_operator
        self.macroscopic = Macroscopic()
        super().__init__()
        assert forcing_scheme == "exact_difference", NotImplementedError(f"Force model {forcing_scheme} not implemented!")
        assert force_vector.shape[0] == self.velocity_set.d, "Check the dimensions of the input force!"
        self.force_vector = force_vector
        self.forcing_operator = ExactDifference(force_vector)

    @property
    def abi_model(self):
        return self.collision_operator.abi_model

    def abi_params(self):
        cp = self.collision_operator.abi_params() or _lib.XlbCollisionParams()
        force = [float(v) for v in torch.as_tensor(self.force_vector).reshape(-1).tolist()]
        cp.has_force = 1
        kernel_force = [0.0] * (3 - len(force)) + force  # kernel axes: a 2-D lattice lives on (y, z)
        for a in range(3):
            cp.force[a] = kernel_force[a]
        return cp

    @Operator.register_backend(ComputeBackend.CUDA)
    def cuda_functional(self, f, feq, omega):
        fout = self.collision_operator(f, feq, omega)
        rho, u = self.macroscopic(fout)
        return self.forcing_operator(fout, feq, rho, u)

    @Operator.register_backend(ComputeBackend.CUDA)
    def cuda_implementation(self, f, feq, fout, omega):
        fout.copy_(self.cuda_functional(f, feq, omega))
        return fout
