"""Indices-based boundary masker on the device
(xlb/operator/boundary_masker/indices_boundary_masker.py:51-143, JAX semantics -- the parity target).

Outputs: ``bc_mask`` uint8 [1, ...] holding the id of the last BC in list order that tags a cell, and
``missing_mask`` bool [q, ...] with ``missing[l, x]`` set iff the pull source ``x - c_l`` lies outside the
domain or on a cell marked solid.  A BC whose index list touches the domain interior is an obstacle: all
its cells become solid and its id is written to the padded set ``indices + c_l`` (only if it
``needs_padding``).  Two kernels: ``xlb_mask_tag`` per BC, ``xlb_mask_missing`` once.
"""

import numpy as np
import torch

from ... import _lib
from ...compute_backend import ComputeBackend
from ..operator import Operator, byref, check, lib, stream_ptr

_generation = [0]


def mark_masks_written(*masks):
    """The kernels write the masks through raw pointers, which torch cannot see: bump each tensor's version counter (an
    in-place op on an empty view; no launch) and stamp a process-wide generation number, so that anything cached per mask
    content -- the stepper's BC cell list -- notices the new content even at a recycled address."""
    _generation[0] += 1
    for m in masks:
        m.view(-1)[:0].zero_()
        m._xlb_generation = _generation[0]


class IndicesBoundaryMasker(Operator):
    def __init__(self, velocity_set=None, precision_policy=None, compute_backend=None, grid=None):
        super().__init__(velocity_set, precision_policy, compute_backend)
        self.grid = grid

    def are_indices_in_interior(self, indices, shape):
        """True per index column that is strictly inside the domain (:51-62)."""
        d = self.velocity_set.d
        hi = np.array(shape)[:d, np.newaxis] - 1
        return np.all((indices[:d] > 0) & (indices[:d] < hi), axis=0)

    @Operator.register_backend(ComputeBackend.CUDA)
    def cuda_implementation(self, bclist, bc_mask, missing_mask, start_index=None, global_shape=None):
        """``start_index`` is the global coordinate of local cell 0 (slab decomposition); ``global_shape``
        defaults to the field shape."""
        vs = self.velocity_set
        d = vs.d
        self._check_field(bc_mask, "bc_mask", 1)
        self._check_field(missing_mask, "missing_mask", vs.q)
        if bc_mask.dtype != torch.uint8 or missing_mask.dtype not in (torch.bool, torch.uint8):
            raise ValueError("bc_mask must be uint8 and missing_mask bool or uint8")
        local_shape = tuple(bc_mask.shape[1:])
        global_shape = tuple(global_shape) if global_shape is not None else local_shape
        start_index = tuple(start_index) if start_index is not None else (0,) * d
        origin = _lib.int3([0] * (3 - d) + list(start_index))
        gshape = _lib.int3([1] * (3 - d) + list(global_shape))
        dom = self._domain(missing_mask)
        solid = torch.zeros(local_shape, dtype=torch.uint8, device=bc_mask.device)
        for bc in bclist:
            assert bc.indices is not None, f"Please specify indices associated with the {bc.__class__.__name__} BC!"
            idx = bc.indices
            if isinstance(idx, torch.Tensor):
                idx_np = idx.detach().cpu().numpy()
            else:
                idx_np = np.asarray(idx)
            idx_np = np.ascontiguousarray(idx_np.reshape(d, -1), dtype=np.int32)
            n = idx_np.shape[1]
            interior = bool(n) and bool(np.any(self.are_indices_in_interior(idx_np, global_shape)))
            idx_dev = torch.from_numpy(idx_np).to(bc_mask.device)
            pad = 1 if (interior and bc.needs_padding) else 0
            check(
                lib().xlb_mask_tag(
                    byref(dom), idx_dev.data_ptr(), n, origin, gshape, int(bc.id), 1 if interior else 0, pad, bc_mask.data_ptr(), solid.data_ptr(), stream_ptr()
                ),
                "IndicesBoundaryMasker",
            )
            bc.__dict__.pop("indices", None)  # :131
            del idx_dev
        check(lib().xlb_mask_missing(byref(dom), solid.data_ptr(), origin, gshape, missing_mask.data_ptr(), stream_ptr()), "IndicesBoundaryMasker")
        torch.cuda.current_stream().synchronize()  # index buffers above are freed on return
        mark_masks_written(bc_mask, missing_mask)
        return bc_mask, missing_mask
