"""Pull streaming (xlb/operator/stream/stream.py:29-62): f_out[l, x] = f[l, x - c_l], periodic."""

import torch

from ...compute_backend import ComputeBackend
from ..operator import Operator, byref, check, lib, stream_ptr


class Stream(Operator):
    @Operator.register_backend(ComputeBackend.CUDA)
    def cuda_implementation(self, f_0, f_1):
        """Out-parameter form (Warp convention, stream.py:113-120): writes and returns ``f_1``."""
        q = self.velocity_set.q
        self._check_field(f_0, "f_0", q)
        self._check_field(f_1, "f_1", q)
        if f_0.dtype != f_1.dtype or f_0.shape != f_1.shape:
            raise ValueError("Stream: f_0 and f_1 must have the same shape and dtype")
        dom = self._domain(f_0)
        check(lib().xlb_stream(byref(dom), self._code(f_0), f_0.data_ptr(), f_1.data_ptr(), stream_ptr()), "Stream")
        return f_1

    @Operator.register_backend(ComputeBackend.CUDA)
    def cuda_functional(self, f):
        """Functional form (JAX convention, stream.py:29-62): returns a new field.  Also streams bool masks."""
        return self.cuda_implementation(f, torch.empty_like(f))
