"""Lattice velocity sets (xlb/velocity_set/velocity_set.py:20-253).

Host-side constants only: the kernels carry their own compile-time copies of the same tables
(csrc/xlb_common.cuh); ``tests/test_host_logic.py::test_velocity_set_matches_oracle`` checks these against the oracle's tables and
the stream / step parity tests (``tests/test_gpu_operators.py``, ``tests/test_gpu_step.py``) check the kernels' copies through their results.
Attributes keep the reference's names: ``c`` (d, q) int, ``w`` (q,), ``opp_indices``, ``cc``, ``qi``,
``main_indices``, ``right_indices``, ``left_indices``, ``center_index``, ``cs``, ``cs2``, ``inv_cs2``.
"""

import math

import numpy as np

from ..default_config import DefaultConfig
from ..compute_backend import ComputeBackend, require_cuda_backend


class VelocitySet:
    def __init__(self, d, q, c, w, precision_policy=None, compute_backend=None):
        self.d = d
        self.q = q
        self.precision_policy = precision_policy or DefaultConfig.default_precision_policy
        self.compute_backend = compute_backend or DefaultConfig.default_backend or ComputeBackend.CUDA
        require_cuda_backend(self.compute_backend)

        self._c = np.array(c)
        self._w = np.array(w, dtype=np.float64)
        dirs = self._c.T.tolist()
        self._opp_indices = np.array([dirs.index([-v for v in ci]) for ci in dirs])
        pairs = [(a, b) for a in range(d) for b in range(a, d)]
        self._cc = np.stack([self._c[a] * self._c[b] for a, b in pairs], axis=1).astype(np.float64)
        self._c_float = self._c.astype(np.float64)
        self._qi = self._cc.copy()
        for k, (a, b) in enumerate(pairs):
            if a == b:
                self._qi[:, k] -= 1.0 / 3.0  # Q = cc - cs^2 I
            else:
                self._qi[:, k] *= 2.0  # symmetric tensor stored once
        speed = np.abs(self._c).sum(axis=0)
        self.main_indices = np.nonzero(speed == 1)[0]
        self.right_indices = np.nonzero(self._c[0] == 1)[0]
        self.left_indices = np.nonzero(self._c[0] == -1)[0]
        self.center_index = int(np.nonzero(speed == 0)[0][0])
        self.cs = math.sqrt(3) / 3.0
        self.cs2 = 1.0 / 3.0
        self.inv_cs2 = 3.0

        # public views (numpy; the reference exposes backend arrays here)
        self.c = self._c.astype(np.int32)
        self.w = self._w
        self.opp_indices = self._opp_indices.astype(np.int32)
        self.cc = self._cc
        self.c_float = self._c_float
        self.qi = self._qi

    def __repr__(self):
        return "D{}Q{}".format(self.d, self.q)

    __str__ = __repr__
