"""Moments of the populations (xlb/operator/macroscopic/{zero_moment,first_moment,second_moment,macroscopic}.py)."""

import torch

from ...compute_backend import ComputeBackend
from ..operator import Operator, byref, check, lib, stream_ptr


class _MomentBase(Operator):
    def _run(self, f, rho, u):
        vs = self.velocity_set
        self._check_field(f, "f", vs.q)
        out = rho if rho is not None else u
        dom = self._domain(f)
        check(
            lib().xlb_macroscopic(byref(dom), self._code(f), f.data_ptr(), self._code(out), self._ptr(rho), self._ptr(u), stream_ptr()),
            self.__class__.__name__,
        )

    def _new(self, f, card):
        return torch.empty((card,) + tuple(f.shape[1:]), dtype=self._out_dtype(f), device=f.device)

    def _out_dtype(self, f):
        # jnp.sum / tensordot keep the dtype of f (int32 lattice tables promote to it)
        return f.dtype


class ZeroMoment(_MomentBase):
    """rho = sum_l f_l (zero_moment.py:14-17)."""

    @Operator.register_backend(ComputeBackend.CUDA)
    def cuda_implementation(self, f, rho):
        self._check_field(rho, "rho", 1)
        self._run(f, rho, None)
        return rho

    @Operator.register_backend(ComputeBackend.CUDA)
    def cuda_functional(self, f):
        return self.cuda_implementation(f, self._new(f, 1))


class FirstMoment(_MomentBase):
    """u = (sum_l c_l f_l) / rho (first_moment.py:14-18).  The kernel recomputes rho = sum_l f_l; the ``rho``
    argument is accepted for signature parity and must be that zeroth moment."""

    @Operator.register_backend(ComputeBackend.CUDA)
    def cuda_implementation(self, f, rho, u):
        self._check_field(u, "u", self.velocity_set.d)
        self._run(f, None, u)
        return u

    @Operator.register_backend(ComputeBackend.CUDA)
    def cuda_functional(self, f, rho):
        return self.cuda_implementation(f, rho, self._new(f, self.velocity_set.d))


class Macroscopic(_MomentBase):
    """(rho, u) in one pass (macroscopic.py:21-26 JAX; :37-53 Warp kernel)."""

    def __init__(self, *args, **kwargs):
        self.zero_moment = ZeroMoment(*args, **kwargs)
        self.first_moment = FirstMoment(*args, **kwargs)
        super().__init__(*args, **kwargs)

    @Operator.register_backend(ComputeBackend.CUDA)
    def cuda_implementation(self, f, rho, u):
        self._check_field(rho, "rho", 1)
        self._check_field(u, "u", self.velocity_set.d)
        if rho.dtype != u.dtype:
            raise ValueError("Macroscopic: rho and u must share a dtype")
        self._run(f, rho, u)
        return rho, u

    @Operator.register_backend(ComputeBackend.CUDA)
    def cuda_functional(self, f):
        return self.cuda_implementation(f, self._new(f, 1), self._new(f, self.velocity_set.d))


class SecondMoment(Operator):
    """Pi = sum_l c_l c_l f_l, packed (xx, xy, xz, yy, yz, zz) / (xx, xy, yy) (second_moment.py:35-55)."""

    @Operator.register_backend(ComputeBackend.CUDA)
    def cuda_implementation(self, f, pi):
        vs = self.velocity_set
        self._check_field(f, "f", vs.q)
        self._check_field(pi, "pi", vs.d * (vs.d + 1) // 2)
        dom = self._domain(f)
        check(lib().xlb_second_moment(byref(dom), self._code(f), f.data_ptr(), self._code(pi), pi.data_ptr(), stream_ptr()), "SecondMoment")
        return pi

    @Operator.register_backend(ComputeBackend.CUDA)
    def cuda_functional(self, fneq):
        vs = self.velocity_set
        pi = torch.empty((vs.d * (vs.d + 1) // 2,) + tuple(fneq.shape[1:]), dtype=fneq.dtype, device=fneq.device)
        return self.cuda_implementation(fneq, pi)
