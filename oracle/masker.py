"""Oracle masks and index helpers (TEST INFRASTRUCTURE -- see oracle/__init__.py).

Restates ``IndicesBoundaryMasker.jax_implementation``
(xlb/operator/boundary_masker/indices_boundary_masker.py:51-143) bit-exactly, plus
``Grid.bounding_box_indices`` (xlb/grid/grid.py:135-191) used to build the test cases.
"""

import numpy as np

from . import ops


def bounding_box_indices(shape, remove_edges=False):
    """grid.py:135-191.  Faces: left x=0, right x=nx-1, front y=0, back y=ny-1, bottom z=0, top z=nz-1."""
    origin = np.array([0, 0, 0])
    bounds = np.array(shape)
    if remove_edges:
        origin += 1
        bounds -= 1
    slice_x = slice(origin[0], bounds[0])
    slice_y = slice(origin[1], bounds[1])
    dim = len(bounds)
    grid = np.indices(shape)
    if dim == 2:
        nx, ny = shape
        box = {
            "bottom": grid[:, slice_x, 0],
            "top": grid[:, slice_x, ny - 1],
            "left": grid[:, 0, slice_y],
            "right": grid[:, nx - 1, slice_y],
        }
    else:
        nx, ny, nz = shape
        slice_z = slice(origin[2], bounds[2])
        box = {
            "bottom": grid[:, slice_x, slice_y, 0].reshape(3, -1),
            "top": grid[:, slice_x, slice_y, nz - 1].reshape(3, -1),
            "left": grid[:, 0, slice_y, slice_z].reshape(3, -1),
            "right": grid[:, nx - 1, slice_y, slice_z].reshape(3, -1),
            "front": grid[:, slice_x, 0, slice_z].reshape(3, -1),
            "back": grid[:, slice_x, ny - 1, slice_z].reshape(3, -1),
        }
    return {k: v.tolist() for k, v in box.items()}


def are_indices_in_interior(lat, indices, shape):
    """indices_boundary_masker.py:51-62."""
    d = lat.d
    shape_array = np.array(shape)
    return np.all((indices[:d] > 0) & (indices[:d] < shape_array[:d, np.newaxis] - 1), axis=0)


def indices_boundary_masker(lat, bclist, bc_mask, missing_mask, start_index=None, pad_x=1):
    """indices_boundary_masker.py:73-143.

    ``pad_x`` is ``jax.device_count()`` in the reference (:81-82); it only changes the width of
    the all-True pad ring and never the cropped result.
    Returns new (bc_mask uint8 [1,...], missing_mask bool [q,...]).
    """
    dim = missing_mask.ndim - 1
    grid_shape = bc_mask[0].shape
    pads = (pad_x, 1) if dim == 2 else (pad_x, 1, 1)
    shift = np.array(pads)[:, np.newaxis]
    if start_index is None:
        start_index = (0,) * dim
    bc_ext = np.pad(bc_mask[0], [(p, p) for p in pads], constant_values=0)
    mm_ext = np.pad(missing_mask.astype(bool), [(0, 0)] + [(p, p) for p in pads], constant_values=True)
    for bc in bclist:
        assert bc.indices is not None
        bc_indices = np.array(bc.indices)
        origin = np.array(start_index)[:, np.newaxis]
        if any(are_indices_in_interior(lat, bc_indices, grid_shape)):  # :107
            solid = bc_indices - origin + shift
            indices_shifted = bc.pad_indices() - origin + shift
            mm_ext[(slice(None),) + tuple(solid)] = True  # :119-125
        else:
            indices_shifted = bc_indices - origin + shift
        bc_ext[tuple(indices_shifted)] = bc.id  # :128, list order, last writer wins
        bc.indices = None  # :131
    mm_ext = ops.stream(lat, mm_ext)  # :134
    crop = tuple(slice(p, -p) for p in pads)
    missing_out = mm_ext[(slice(None),) + crop].copy()
    bc_out = bc_mask.copy()
    bc_out[0] = bc_ext[crop]
    return bc_out, missing_out
