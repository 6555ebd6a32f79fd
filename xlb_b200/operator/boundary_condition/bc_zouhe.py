"""Zou-He velocity / pressure boundary condition (xlb/operator/boundary_condition/bc_zouhe.py:48-304).

Prescribed data: ``prescribed_value`` (a velocity vector with one non-zero normal component, or a scalar density for
``"pressure"``) or a ``profile`` callable.  Both profile conventions of the reference are accepted, told apart by the
callable's arity:

* ``profile()`` -> the full array, e.g. ``(d, ny, nz)`` for an x-face inlet (the JAX convention,
  examples/cfd/flow_past_sphere_3d.py:62-79, bc_zouhe.py:119-120);
* ``profile(index)`` -> one value per boundary cell, the inflow speed along the inward normal (velocity) or the density
  (pressure), ``index`` being the cell's integer coordinates (the Warp convention: a ``wp.func`` over ``wp.vec3i``,
  flow_past_sphere_3d.py:89-102, bc_zouhe.py:332-337).  It is evaluated once, at construction, over the BC's index list
  -- vectorised (``index[a]`` are NumPy arrays) where the function's body allows it, cell by cell otherwise.

Either way the values live in a side buffer on the device and are addressed with broadcast strides -- nothing is
encoded in ``f_1`` (that encoding is an artefact of the Warp backend, helper_functions_bc.py:371-562).
"""

import inspect

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
        self._pv_device = None
        self._pv_box = None  # per-index profile: (lower corner of the index set's bounding box, dense value box)
        fn = getattr(self.profile, "func", self.profile)  # a warp.Function keeps the Python function in .func
        required = [a for a in inspect.signature(fn).parameters.values()
                    if a.default is inspect.Parameter.empty and a.kind in (a.POSITIONAL_ONLY, a.POSITIONAL_OR_KEYWORD)]
        if required:  # profile(index): the Warp convention; a callable without required arguments is the JAX convention
            self.prescribed_values = None
            self._evaluate_index_profile(fn, indices)
            return
        values = self.profile()  # evaluated eagerly like the JAX path (bc_zouhe.py:119-120)
        if isinstance(values, torch.Tensor):
            values = values.detach().cpu().numpy()
        values = np.asarray(values)
        if values.dtype not in (np.float16, np.float32, np.float64):
            values = values.astype(self.precision_policy.store_precision.np_dtype)
        self.prescribed_values = values

    def _evaluate_index_profile(self, fn, indices):
        """``profile(index) -> value`` over the BC's own cells, stored as a dense box around the index set."""
        if indices is None:
            raise NotImplementedError("a per-index profile needs the boundary condition's indices (mesh-based BCs: pass an array profile)")
        idx = np.asarray(indices.detach().cpu() if isinstance(indices, torch.Tensor) else indices, dtype=np.int64).reshape(self.velocity_set.d, -1)
        if idx.shape[1] == 0:
            raise ValueError("a per-index profile needs a non-empty index list")
        sdt = self.precision_policy.store_precision.np_dtype
        if sdt == np.float16:
            sdt = np.float32
        full = tuple(idx) + ((np.zeros_like(idx[0]),) if idx.shape[0] == 2 else ())  # the reference always passes a vec3i
        try:
            values = np.asarray(fn(full), dtype=np.float64).reshape(-1)  # vectorised: index[a] are arrays
            if values.size == 1:
                values = np.full(idx.shape[1], float(values[0]))
            if values.size != idx.shape[1]:
                raise ValueError("profile did not broadcast over the index arrays")
        except Exception:  # the body only works on scalars (e.g. real warp builtins): cell by cell
            values = np.array([float(np.asarray(fn(tuple(int(v) for v in col))).reshape(-1)[0]) for col in np.stack(full, axis=1)])
        lo = idx.min(axis=1)
        box = np.zeros(tuple(idx.max(axis=1) - lo + 1), dtype=sdt)
        box[tuple(idx - lo[:, None])] = values.astype(sdt)
        self._pv_box = (lo, box)

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
        if self._pv_box is not None:
            # one value per cell (XlbBcEntry.pv_mode = 1), addressed by the kernel's local cell coordinates: axes along which the
            # index set is one cell thick are broadcast, the others are shifted by the box corner minus the block's origin
            lo, box = self._pv_box
            if self._pv_device is None:
                self._pv_device = torch.from_numpy(np.ascontiguousarray(box)).to(current_device())
            pv = self._pv_device
            origin = getattr(self, "_block_origin", None) or (0,) * self.velocity_set.d
            strides = [0 if n == 1 else s for n, s in zip(box.shape, pv.stride())]
            shift = sum(s * (int(origin[a]) - int(lo[a])) for a, s in enumerate(strides))
            row["pv"] = pv.data_ptr() + shift * pv.element_size()
            row["pv_dtype"] = abi_code_of(pv.dtype)
            row["pv_mode"] = 1
            row["pv_stride"][0] = 0
            row["pv_stride"][1:] = kernel_axes(strides, self.velocity_set.d)
            return
        shp = self._broadcast_shape(field_shape)
        if self._pv_device is None or tuple(self._pv_device.shape) != shp:
            self._pv_device = torch.from_numpy(np.ascontiguousarray(self.prescribed_values.reshape(shp))).to(current_device())
        pv = self._pv_device
        strides = [0 if n == 1 else s for n, s in zip(shp, pv.stride())]
        row["pv"] = pv.data_ptr()
        row["pv_dtype"] = abi_code_of(pv.dtype)
        row["pv_stride"][0] = strides[0]
        row["pv_stride"][1:] = kernel_axes(strides[1:], self.velocity_set.d)
