"""Ray-cast mesh boundary masker on the device (xlb/operator/boundary_masker/ray.py:17-121 with the shared base
mesh_boundary_masker.py:181-245).

Same call as the reference's Warp operator: ``masker(bc, distances, bc_mask, missing_mask) -> (distances, bc_mask,
missing_mask)``.  ``bc.mesh_vertices`` is an ``(N, 3)`` triangle soup in lattice units (three consecutive rows per
triangle, as ``trimesh.load_mesh(..., process=False).vertices`` gives for an STL, windtunnel_3d.py:80-81).  Cells whose
unit-length lattice rays hit the surface get ``bc.id`` in ``bc_mask`` and ``missing_mask[opp l] = True``; with
``bc.needs_mesh_distance`` the fractional distance to the surface is written to ``distances[l]`` (the reference passes
``f_1`` there, nse_stepper.py:203).  One ``xlb_mask_mesh_ray`` call: a triangle-parallel kernel plus the out-of-bound
post-pass.  The reference's BVH query (``warp.mesh_query_ray``) is replaced by the watertight ray/triangle test it is
built on; see csrc/xlb_mesh.cuh.
"""

import numpy as np
import torch

from ... import _lib
from ...compute_backend import ComputeBackend
from ..operator import Operator, byref, check, lib, stream_ptr
from .indices_boundary_masker import mark_masks_written


class MeshBoundaryMasker(Operator):
    """Input preparation shared by the mesh maskers (mesh_boundary_masker.py:17-52, 181-214)."""

    def __init__(self, velocity_set=None, precision_policy=None, compute_backend=None):
        super().__init__(velocity_set, precision_policy, compute_backend)
        if self.velocity_set.d == 2:
            raise NotImplementedError("This Operator is not implemented in 2D!")  # mesh_boundary_masker.py:37-39

    def _prepare_kernel_inputs(self, bc, bc_mask, global_shape=None):
        assert bc.mesh_vertices is not None, f'Please provide the mesh vertices for {bc.__class__.__name__} BC using keyword "mesh_vertices"!'
        assert bc.indices is None, f"Please use IndicesBoundaryMasker operator if {bc.__class__.__name__} is imposed on known indices of the grid!"
        mesh_vertices = np.asarray(bc.mesh_vertices.detach().cpu() if isinstance(bc.mesh_vertices, torch.Tensor) else bc.mesh_vertices)
        assert mesh_vertices.ndim == 2 and mesh_vertices.shape[1] == self.velocity_set.d, (
            "Mesh points must be reshaped into an array (N, 3) where N indicates number of points!"
        )
        if mesh_vertices.shape[0] % 3 != 0:
            raise ValueError("mesh_vertices must hold three consecutive vertices per triangle (N divisible by 3)")
        grid_shape = tuple(global_shape) if global_shape is not None else tuple(bc_mask.shape[1:])
        mesh_min = np.min(mesh_vertices, axis=0)
        mesh_max = np.max(mesh_vertices, axis=0)
        if any(mesh_min < 0) or any(mesh_max >= grid_shape):
            raise ValueError(
                f"Mesh extents ({mesh_min}, {mesh_max}) exceed domain dimensions {grid_shape}. The mesh must be fully contained within the domain."
            )
        bc.__dict__.pop("mesh_vertices", None)  # mesh_boundary_masker.py:203-204
        tri = torch.from_numpy(np.ascontiguousarray(mesh_vertices, dtype=np.float32)).to(bc_mask.device)
        return tri, int(bc.id)


class MeshMaskerRay(MeshBoundaryMasker):
    @Operator.register_backend(ComputeBackend.CUDA)
    def cuda_implementation(self, bc, distances, bc_mask, missing_mask, start_index=None, global_shape=None):
        """``start_index`` / ``global_shape``: global coordinate of local cell 0 and global extents (slab decomposition)."""
        vs = self.velocity_set
        self._check_field(bc_mask, "bc_mask", 1)
        self._check_field(missing_mask, "missing_mask", vs.q)
        if bc_mask.dtype != torch.uint8 or missing_mask.dtype not in (torch.bool, torch.uint8):
            raise ValueError("bc_mask must be uint8 and missing_mask bool or uint8")
        tri, bc_id = self._prepare_kernel_inputs(bc, bc_mask, global_shape)
        local_shape = tuple(bc_mask.shape[1:])
        origin = _lib.int3(start_index if start_index is not None else (0, 0, 0))
        gshape = _lib.int3(global_shape if global_shape is not None else local_shape)
        dom = self._domain(missing_mask)
        keys = None
        if bc.needs_mesh_distance:
            if distances is None:
                raise ValueError("this boundary condition needs the mesh distances: pass a [q, nx, ny, nz] field (the reference passes f_1)")
            self._check_field(distances, "distances", vs.q)
            keys = torch.full((vs.q,) + local_shape, -1, dtype=torch.int64, device=bc_mask.device)  # all ones = no hit yet
        check(
            lib().xlb_mask_mesh_ray(
                byref(dom), tri.data_ptr(), tri.shape[0] // 3, origin, gshape, bc_id, bc_mask.data_ptr(), missing_mask.data_ptr(),
                None if keys is None else keys.data_ptr(), stream_ptr(),
            ),
            "MeshMaskerRay",
        )
        if keys is not None:
            hit = keys != -1
            weight = (keys & 0xFFFFFFFF).to(torch.int32).view(torch.float32)  # low word = bits of the distance weight
            distances[hit] = weight[hit].to(distances.dtype)  # ray.py:75-76
        torch.cuda.current_stream().synchronize()  # the triangle buffer is freed on return
        mark_masks_written(bc_mask, missing_mask)
        return distances, bc_mask, missing_mask
