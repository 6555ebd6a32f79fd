"""Zou-He velocity / pressure boundary condition (xlb/operator/boundary_condition/bc_zouhe.py:48-304).

Prescribed data follow the JAX convention of the reference: ``prescribed_value`` (a velocity vector with
one non-zero normal component, or a scalar density for ``"pressure"``) or a ``profile`` callable returning
the full array, e.g. ``(d, ny, nz)`` for an x-face inlet (examples/cfd/flow_past_sphere_3d.py:58-81).  The
array is kept in a side buffer on the device and addressed with broadcast strides -- nothing is encoded
in ``f_1`` (that encoding is an artefact of the Warp backend, helper_functions_bc.py:371-562).
"""

import numpy as np
import torch

from ... import _lib
from ...default_config import current_device
from ...precision_policy import abi_code_of
from .boundary_condition import BoundaryCondition, ImplementationStep, kernel_axes


class ZouHeBC(BoundaryCondition):
    def __init__(
        self, bc_type, profile=None, prescribed_value=None, velocity_set=None, precision_policy=None, compute_backend=None, indices=None,
        mesh_vertices=None, voxelization_method=None,
    ):
        assert bc_type in ["velocity", "pressure"], f"type = {bc_type} not supported! Use 'pressure' or 'velocity'."
        self.bc_type = bc_type
        super().__init__(ImplementationStep.STREAMING, velocity_set, precision_policy, compute_backend, indices, mesh_vertices, voxelization_method)
        self.abi_kind = _lib.BC_ZOUHE_VELOCITY if bc_type == "velocity" else _lib.BC_ZOUHE_PRESSURE
        self.needs_padding = True
        self.profile = profile
        if prescribed_value is not None:
            if profile is not None:
                raise ValueError("Cannot specify both profile and prescribed_value")
            if bc_type == "velocity":
                if not isinstance(prescribed_value, (tuple, list, np.ndarray)):
                    raise ValueError("Velocity prescribed_value must be a tuple, list, or array-like")
                prescribed_value = np.asarray(prescribed_value, dtype=np.float64)
            else:
                if not isinstance(prescribed_value, (int, float)):
                    raise ValueError("Pressure prescribed_value must be a scalar (int or float)")
                prescribed_value = float(prescribed_value)
            self.prescribed_value = prescribed_value
            sdt = self.precision_policy.store_precision.np_dtype
            self.profile = lambda: np.array(prescribed_value, dtype=sdt).reshape(-1, 1)  # bc_zouhe.py:156-157
        if self.profile is None:
            raise ValueError("ZouHeBC needs either prescribed_value or profile")
        values = self.profile()  # evaluated eagerly like the JAX path (bc_zouhe.py:119-120)
        if isinstance(values, torch.Tensor):
            values = values.detach().cpu().numpy()
        values = np.asarray(values)
        if values.dtype not in (np.float16, np.float32, np.float64):
            values = values.astype(self.precision_policy.store_precision.np_dtype)
        self.prescribed_values = values
        self._pv_device = None

    def _broadcast_shape(self, field_shape):
        """Shape rule of ``_broadcast_prescribed_values`` (bc_zouhe.py:178-215): singleton axes are inserted
        after the component axis; every axis must then be 1 or match the field."""
        d = self.velocity_set.d
        lead = d if self.bc_type == "velocity" else 1
        target = (lead,) + tuple(field_shape)
        shp = tuple(self.prescribed_values.shape)
        if len(shp) > len(target):
            raise ValueError("prescribed_values has more dimensions than target_shape")
        if len(shp) == 0:
            shp = (1,) * len(target)
        elif len(shp) < len(target):
            shp = (shp[0],) + (1,) * (len(target) - len(shp)) + shp[1:]
        for have, want in zip(shp, target):
            if have != 1 and have != want:
                raise ValueError(f"Cannot broadcast dimension {have} to {want}")
        return shp

    def bc_table_row(self, row, field_shape=None):
        super().bc_table_row(row, field_shape)
        if field_shape is None:
            raise ValueError("ZouHeBC needs the field shape to lay out its prescribed values")
        shp = self._broadcast_shape(field_shape)
        if self._pv_device is None or tuple(self._pv_device.shape) != shp:
            self._pv_device = torch.from_numpy(np.ascontiguousarray(self.prescribed_values.reshape(shp))).to(current_device())
        pv = self._pv_device
        strides = [0 if n == 1 else s for n, s in zip(shp, pv.stride())]
        row["pv"] = pv.data_ptr()
        row["pv_dtype"] = abi_code_of(pv.dtype)
        row["pv_stride"][0] = strides[0]
        row["pv_stride"][1:] = kernel_axes(strides[1:], self.velocity_set.d)
