"""xlb_mask_indices_static: the allocation-free variant of the indices masker (include/xlb_b200.h; no counterpart in the
reference).  Its per-cell ``solid`` scratch is a static __device__ array inside
the library instead of a caller-allocated buffer.  Checked here, through the C ABI, against the existing entry points
(xlb_mask_tag + xlb_mask_missing): bit-exact masks (the bar of every mask test), on two calls in a row whose second would
expose scratch left over from the first.

Collected last on purpose (``zz``): a failure here cannot hide other tests.
"""

import ctypes

import numpy as np
import pytest
import torch

from oracle import cases
from xlb_b200 import _lib

pytestmark = pytest.mark.gpu


def _domain(q, d, shape):
    dom = _lib.XlbDomain()
    nx, ny, nz = (1,) + tuple(shape) if d == 2 else tuple(shape)
    dom.d, dom.q, dom.nx, dom.ny, dom.nz = d, q, nx, ny, nz
    dom.x_periodic, dom.compute_dtype, dom.store_dtype = 1, 1, 1
    dom.q_stride = dom.mask_q_stride = nx * ny * nz
    return dom


def _bc_lists(case):
    """(id, int32 index array [d][n], interior?, needs_padding?) per BC, the host-side preparation of
    IndicesBoundaryMasker (indices_boundary_masker.py:100-128) restated with NumPy."""
    shape = np.array(case["shape"])
    d = len(shape)
    pads = {"HalfwayBounceBackBC", "ZouHeBC", "RegularizedBC"}  # the BCs with needs_padding (bc_halfway_bounce_back.py:60, bc_zouhe.py:146)
    out = []
    for i, spec in enumerate(case["bcs"]):
        idx = np.ascontiguousarray(np.array(spec["indices"], dtype=np.int32).reshape(d, -1))
        interior = bool(np.any(np.all((idx > 0) & (idx < shape[:, None] - 1), axis=0)))
        out.append((i + 1, idx, interior, spec["kind"] in pads))
    return out


def _existing(case, q, dom, origin, gshape, stream):
    """bc_mask, missing_mask through xlb_mask_tag / xlb_mask_missing with a caller-allocated solid array."""
    lib = _lib.lib()
    shape = tuple(case["shape"])
    bc = torch.zeros((1,) + shape, dtype=torch.uint8, device="cuda")
    mm = torch.zeros((q,) + shape, dtype=torch.uint8, device="cuda")
    solid = torch.zeros(shape, dtype=torch.uint8, device="cuda")
    keep = []
    for bc_id, idx, interior, pads in _bc_lists(case):
        dev = torch.from_numpy(idx).cuda()
        keep.append(dev)
        _lib.check(lib.xlb_mask_tag(ctypes.byref(dom), dev.data_ptr(), idx.shape[1], origin, gshape, bc_id, int(interior), int(interior and pads),
                                    bc.data_ptr(), solid.data_ptr(), stream), "xlb_mask_tag")
    _lib.check(lib.xlb_mask_missing(ctypes.byref(dom), solid.data_ptr(), origin, gshape, mm.data_ptr(), stream), "xlb_mask_missing")
    torch.cuda.synchronize()
    return bc, mm


def _static(case, q, dom, origin, gshape, stream, outputs, index_buffers):
    """The same through xlb_mask_indices_static.  Outputs and index buffers are allocated by the CALLER, before the calls."""
    lib = _lib.lib()
    bc, mm = outputs
    lists = _bc_lists(case)
    for k, ((bc_id, idx, interior, pads), dev) in enumerate(zip(lists, index_buffers)):
        rc = lib.xlb_mask_indices_static(ctypes.byref(dom), dev.data_ptr(), idx.shape[1], origin, gshape, bc_id, int(interior), int(interior and pads),
                                         int(k == 0), int(k == len(lists) - 1), bc.data_ptr(), mm.data_ptr(), stream)
        _lib.check(rc, "xlb_mask_indices_static")
    torch.cuda.synchronize()
    return bc, mm


def test_static_masker_equals_existing_entry_points_twice():
    # first call: many solid cells (a large sphere); second call: a smaller lattice with few -- scratch that was not cleared
    # would leave stale solid cells of the first call in the second call's missing_mask
    todo = [cases.sphere3d((48, 24, 24), radius=7), cases.sphere3d((40, 20, 20), lattice="D3Q19", collision="BGK", radius=2), cases.cavity2d(24)]
    stream = _lib.current_stream_ptr()
    for case in todo:
        d = len(case["shape"])
        q = {"D2Q9": 9, "D3Q19": 19, "D3Q27": 27}[case["lattice"]]
        dom = _domain(q, d, case["shape"])
        origin = _lib.int3([0, 0, 0])
        gshape = _lib.int3([1] * (3 - d) + list(case["shape"]))
        want_bc, want_mm = _existing(case, q, dom, origin, gshape, stream)
        # caller-side buffers first, so that nothing is allocated between here and the end of the static calls
        shape = tuple(case["shape"])
        outputs = (torch.zeros((1,) + shape, dtype=torch.uint8, device="cuda"), torch.zeros((q,) + shape, dtype=torch.uint8, device="cuda"))
        index_buffers = [torch.from_numpy(idx).cuda() for _, idx, _, _ in _bc_lists(case)]
        torch.cuda.synchronize()
        before = torch.cuda.memory_stats()["allocation.all.allocated"]
        n0 = _lib.launch_count()
        got_bc, got_mm = _static(case, q, dom, origin, gshape, stream, outputs, index_buffers)
        assert torch.cuda.memory_stats()["allocation.all.allocated"] == before, "the static entry point went through torch's allocator"
        assert _lib.launch_count() - n0 == len(case["bcs"]) + 1  # one tag kernel per BC + the missing pass
        assert torch.equal(got_bc, want_bc), case["name"]
        assert torch.equal(got_mm, want_mm), case["name"]
        # and both equal the oracle's masks (the existing path's own parity test, repeated here for the new one)
        ost, _ = cases.build_oracle(case)
        _, _, bc_o, mm_o = ost.prepare_fields()
        assert np.array_equal(got_bc.cpu().numpy(), bc_o) and np.array_equal(got_mm.cpu().numpy().astype(bool), mm_o)


def test_static_masker_rejects_blocks_beyond_its_scratch():
    lib = _lib.lib()
    dom = _domain(19, 3, (512, 256, 129))  # 16.9 M cells > XLB_MASK_STATIC_MAX_CELLS = 16 Mi
    assert 512 * 256 * 129 > _lib.MASK_STATIC_MAX_CELLS
    rc = lib.xlb_mask_indices_static(ctypes.byref(dom), None, 0, None, None, 1, 0, 0, 1, 0, None, None, _lib.current_stream_ptr())
    assert rc == _lib.EINVAL
    msg = lib.xlb_last_error().decode()
    assert "exceeds the static scratch" in msg and "XLB_MASK_STATIC_MAX_CELLS" in msg
    with pytest.raises(ValueError):
        _lib.check(rc, "xlb_mask_indices_static")
