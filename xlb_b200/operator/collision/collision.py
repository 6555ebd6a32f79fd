"""Collision operators (xlb/operator/collision/{collision,bgk,kbc}.py)."""

import torch

from ... import _lib
from ...compute_backend import ComputeBackend
from ...velocity_set import D2Q9, D3Q27
from ..operator import Operator, byref, check, lib, stream_ptr


class Collision(Operator):
    """Base class: ``collision(f, feq, omega) -> fout`` or ``collision(f, feq, fout, omega)``."""

    abi_model = None

    def _run(self, f, feq, fout, omega):
        q = self.velocity_set.q
        for name, t in (("f", f), ("feq", feq), ("fout", fout)):
            self._check_field(t, name, q)
        if not (f.dtype == feq.dtype == fout.dtype):
            raise ValueError("collision: f, feq and fout must share a dtype")
        dom = self._domain(f)
        check(
            lib().xlb_collide(byref(dom), self.abi_model, self._code(f), f.data_ptr(), feq.data_ptr(), fout.data_ptr(), float(omega), stream_ptr()),
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
