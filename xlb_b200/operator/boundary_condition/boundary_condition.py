"""Boundary-condition base class (xlb/operator/boundary_condition/boundary_condition.py:26-180).

A BC object is a host-side descriptor: id, implementation step, index list, prescribed data.  The fused
step kernel reads one ``XlbBcEntry`` row per id (``bc_table_row``); the same row drives the stand-alone
``bc(f_pre, f_post, bc_mask, missing_mask)`` call, which runs the generic per-BC kernel like the
reference's Warp path (:146-180) with the JAX semantics of each ``bc_*.py``.
"""

from enum import Enum, auto

import numpy as np
import torch

from ... import _lib
from ...compute_backend import ComputeBackend
from ...default_config import DefaultConfig, current_device
from ..operator import Operator, byref, check, lib, stream_ptr
from .boundary_condition_registry import boundary_condition_registry


class ImplementationStep(Enum):
    COLLISION = auto()
    STREAMING = auto()


def kernel_axes(values, d, fill=0):
    """Map per-lattice-dimension values onto the three kernel axes (2-D fields run as nx=1, ny=NX, nz=NY)."""
    values = [int(v) for v in values]
    return [fill] * (3 - d) + values


class BoundaryCondition(Operator):
    abi_kind = _lib.BC_NONE

    def __init__(
        self,
        implementation_step,
        velocity_set=None,
        precision_policy=None,
        compute_backend=None,
        indices=None,
        mesh_vertices=None,
        voxelization_method=None,
    ):
        self.id = boundary_condition_registry.register_boundary_condition(self.__class__.__name__ + "_" + str(hash(self)))
        velocity_set = velocity_set or DefaultConfig.velocity_set
        precision_policy = precision_policy or DefaultConfig.default_precision_policy
        compute_backend = compute_backend or DefaultConfig.default_backend
        super().__init__(velocity_set, precision_policy, compute_backend)
        self.indices = indices
        self.mesh_vertices = mesh_vertices  # (N, 3) triangle soup in lattice units, voxelised by MeshMaskerRay (boundary_condition.py:76)
        self.voxelization_method = voxelization_method
        self.implementation_step = implementation_step
        # flags of the reference API (boundary_condition.py:84-105)
        self.needs_padding = False
        self.needs_mesh_distance = False
        self.needs_aux_init = False
        self.is_initialized_with_aux_data = False
        self.num_of_aux_data = 0
        self.needs_aux_recovery = False
        self._row_cache = None

    def pad_indices(self):
        """indices + c_l for every l, unique (boundary_condition.py:123-136); identity unless ``needs_padding``."""
        idx = np.array(self.indices)
        if not self.needs_padding:
            return idx
        padded = idx[:, :, None] + self.velocity_set._c[:, None, :]
        return np.unique(padded.reshape(self.velocity_set.d, -1), axis=1)

    def assemble_auxiliary_data(self, f_pre, f_post, bc_mask, missing_mask):
        """Post-collision hook (boundary_condition.py:138-144): identity unless overridden."""
        return f_post

    # ---- table row consumed by the kernels
    def bc_table_row(self, row, field_shape=None):
        """Fill ``row`` (one element of a ``_lib.BC_ENTRY_DTYPE`` array).  Subclasses extend."""
        row["kind"] = self.abi_kind

    def _single_row_table(self, field_shape):
        key = tuple(field_shape)
        if self._row_cache is None or self._row_cache[0] != key:
            table = np.zeros(256, dtype=_lib.BC_ENTRY_DTYPE)
            self.bc_table_row(table[self.id], field_shape)
            dev = torch.from_numpy(table.view(np.uint8).copy()).to(current_device())
            self._row_cache = (key, dev)
        return self._row_cache[1]

    def _apply(self, f_pre, f_post, bc_mask, missing_mask, assemble=0):
        q = self.velocity_set.q
        self._check_field(f_pre, "f_pre", q)
        self._check_field(f_post, "f_post", q)
        self._check_field(bc_mask, "bc_mask", 1)
        self._check_field(missing_mask, "missing_mask", q)
        if f_pre.dtype != f_post.dtype:
            raise ValueError("boundary condition: f_pre and f_post must share a dtype")
        if bc_mask.dtype != torch.uint8 or missing_mask.dtype not in (torch.uint8, torch.bool):
            raise ValueError("bc_mask must be uint8 and missing_mask bool or uint8")
        table = self._single_row_table(f_post.shape[1:])
        dom = self._domain(f_post)
        check(
            lib().xlb_apply_bc(
                byref(dom), self._code(f_post), f_pre.data_ptr(), f_post.data_ptr(), f_post.data_ptr(), bc_mask.data_ptr(), missing_mask.data_ptr(),
                table.data_ptr(), self.id, assemble, stream_ptr(),
            ),
            self.__class__.__name__,
        )
        return f_post

    @Operator.register_backend(ComputeBackend.CUDA)
    def cuda_implementation(self, f_pre, f_post, bc_mask, missing_mask):
        """``bc(f_pre, f_post, bc_mask, missing_mask)``: updates ``f_post`` in place on cells tagged with this id."""
        return self._apply(f_pre, f_post, bc_mask, missing_mask)

    def __call__(self, *args, callback=None, **kwargs):
        # every subclass shares the registered implementation of the base class
        result = BoundaryCondition.cuda_implementation(self, *args, **kwargs)
        if callback and callable(callback):
            callback(result)
        return result
