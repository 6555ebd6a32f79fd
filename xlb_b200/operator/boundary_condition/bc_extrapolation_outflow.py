"""Extrapolation outflow (xlb/operator/boundary_condition/bc_extrapolation_outflow.py:48-145).

STREAMING step: missing populations pick up the values stored by the previous step in the outgoing
slots; after collision ``assemble_auxiliary_data`` stores (1 - cs) f(x) + cs f(x - n) of this step's
post-stream populations into those slots again.  In the fused kernel the neighbour's post-stream value is
gathered straight from ``f_0`` (``f_0[l, x - n - c_l]``, like the Warp kernel :203-210); it equals the JAX
whole-array result unless the neighbour cell x - n is itself a STREAMING-step boundary cell.
"""

from collections import Counter

import numpy as np

from ... import _lib
from .boundary_condition import BoundaryCondition, ImplementationStep, kernel_axes


class ExtrapolationOutflowBC(BoundaryCondition):
    abi_kind = _lib.BC_EXTRAPOLATION

    def __init__(self, velocity_set=None, precision_policy=None, compute_backend=None, indices=None, mesh_vertices=None, voxelization_method=None):
        super().__init__(ImplementationStep.STREAMING, velocity_set, precision_policy, compute_backend, indices, mesh_vertices, voxelization_method)
        if indices is None:
            raise ValueError("ExtrapolationOutflowBC derives its normal from the index list; pass indices")
        self._get_normal_vectors(indices)

    def _get_normal_vectors(self, indices):
        """Integer outward normal from the most frequent coordinate value per axis (:78-92)."""
        top = [Counter(np.asarray(coord).tolist()).most_common(1)[0] for coord in indices]
        counts = np.array([n for _, n in top])
        values = np.array([v for v, _ in top])
        self.normal = counts // counts.max()
        if values[np.argmax(counts)] == 0:
            self.normal *= -1

    def bc_table_row(self, row, field_shape=None):
        super().bc_table_row(row, field_shape)
        row["normal"][:] = kernel_axes(self.normal, self.velocity_set.d)

    def assemble_auxiliary_data(self, f_pre, f_post, bc_mask, missing_mask):
        """Stand-alone form: ``f_pre`` = post-stream, ``f_post`` = post-collision (updated in place)."""
        return self._apply(f_pre, f_post, bc_mask, missing_mask, assemble=1)
