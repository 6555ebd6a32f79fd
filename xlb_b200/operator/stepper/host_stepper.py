"""Step populations that live in HOST memory: an x-chunk pipeline over three CUDA streams.

    h2d stream     :  chunk k+1 (planes a-1 .. b of f_0, pinned host -> device)
    compute stream :  chunk k    (xlb_step on the staged planes)
    d2h stream     :  chunk k-1  (planes a .. b-1 of f_1, device -> pinned host)

so that one step costs about max(H2D, D2H, kernel) instead of their sum; PCIe is full duplex.  The masks and the BC
table stay resident on the device for the whole grid (they are O(1) bytes per cell and static); only the populations
travel.  This is the path ``bench.py`` times as ``e2e`` (host buffers in, host buffers out), and the natural way to
run a lattice whose two population buffers exceed HBM (the reference's experimental out-of-core arrays,
xlb/experimental/ooc, do the same with CuPy tiles).

The per-cell arithmetic is the same ``xlb_step_ex`` launch as the resident path, with the wrapped stepper's collision
parameters (LES constant, body force, BC kinds; pull from ghost planes, x not periodic inside a chunk; the global periodic
wrap is resolved when staging planes -1 and nx), so results are bit-identical to ``stepper(f_0, f_1, ...)`` on
device-resident fields.  Each chunk moves with ONE strided batch copy per direction (``xlb_copy_runs``: q runs of
``planes * ny * nz`` elements), not q separate copies.
"""

import ctypes

import numpy as np
import torch

from ... import _lib
from ..operator import byref, check, lib


class HostStreamedStepper:
    def __init__(self, stepper, chunk_planes=32):
        if stepper.velocity_set.d != 3:
            raise NotImplementedError("HostStreamedStepper: 3-D lattices only")
        if getattr(stepper.grid, "halo", 0):
            raise NotImplementedError("HostStreamedStepper: use a single-block grid (no slab decomposition)")
        self.stepper = stepper
        self.chunk_planes = int(chunk_planes)
        self._state = None

    def _setup(self, f_host):
        st = self.stepper
        q, nx, ny, nz = f_host.shape
        P = max(2, min(self.chunk_planes, nx))
        dev = torch.device("cuda", torch.cuda.current_device())
        bufs_in = [torch.empty((q, P + 2, ny, nz), dtype=f_host.dtype, device=dev) for _ in range(2)]
        bufs_out = [torch.empty((q, P + 2, ny, nz), dtype=f_host.dtype, device=dev) for _ in range(2)]
        streams = dict(h2d=torch.cuda.Stream(), d2h=torch.cuda.Stream())
        table, n_bc = st.build_bc_table((nx, ny, nz))
        for bc in st.boundary_conditions:  # prescribed-value arrays are addressed with chunk-local x
            row = np.zeros(1, dtype=_lib.BC_ENTRY_DTYPE)
            bc.bc_table_row(row[0], (nx, ny, nz))
            if int(row["pv"][0]) != 0 and int(row["pv_stride"][0][1]) != 0:
                raise NotImplementedError("HostStreamedStepper: prescribed profiles must be constant along x")
        # chunk list: [a, b) with the last chunk at least 2 planes thick (extrapolation outflow reads x-2)
        chunks = []
        a = 0
        while a < nx:
            b = min(a + P, nx)
            if nx - b == 1:
                b = nx if b - a < P else b - 1
            chunks.append((a, b))
            a = b
        # the wrapped stepper's collision parameters (LES constant, force, BC kinds) without its BC cell list, whose
        # indices refer to the resident block, not to a staged chunk
        cp = _lib.XlbCollisionParams()
        ctypes.memmove(ctypes.byref(cp), ctypes.byref(st._collision_params), ctypes.sizeof(cp))
        cp.bc_cells, cp.n_bc_cells = None, 0
        cp.missing_bits = None  # the packed words are indexed by the resident block's cells as well
        cp.halfway_fraction = 0.0
        self._state = dict(shape=tuple(f_host.shape), P=P, cp=cp, bufs_in=bufs_in, bufs_out=bufs_out, streams=streams, table=table, n_bc=n_bc,
                           chunks=chunks,
                           ev_in=[torch.cuda.Event() for _ in range(2)], ev_done=[torch.cuda.Event() for _ in range(2)],
                           ev_out=[torch.cuda.Event() for _ in range(2)], ev_free_in=[torch.cuda.Event() for _ in range(2)])
        return self._state

    @staticmethod
    def _copy_planes(dst, dst_plane, src, src_plane, n_planes, stream):
        """planes [src_plane, src_plane + n) of every population of ``src`` -> planes [dst_plane, ...) of ``dst``: one batch copy."""
        q = src.shape[0]
        plane = src.shape[2] * src.shape[3]
        esz = src.element_size()
        rc = lib().xlb_copy_runs(
            dst.data_ptr() + dst_plane * plane * esz, dst.stride(0) * esz, src.data_ptr() + src_plane * plane * esz, src.stride(0) * esz,
            n_planes * plane * esz, q, _lib.c_void_p(stream.cuda_stream),
        )
        check(rc, "HostStreamedStepper")

    def _stage(self, f_host, buf, a, b, stream):
        """planes a-1 .. b (periodic) of every population -> buf[:, 0 : b-a+2]"""
        nx = f_host.shape[1]
        lo, hi = a - 1, b + 1
        s0, s1 = max(lo, 0), min(hi, nx)
        self._copy_planes(buf, s0 - lo, f_host, s0, s1 - s0, stream)
        if lo < 0:
            self._copy_planes(buf, 0, f_host, nx - 1, 1, stream)
        if hi > nx:
            self._copy_planes(buf, hi - lo - 1, f_host, 0, 1, stream)

    def __call__(self, f0_host, f1_host, bc_mask, missing_mask, omega, timestep):
        """``f0_host`` / ``f1_host``: pinned host tensors ``[q, nx, ny, nz]`` in the store dtype; masks on the device."""
        st = self.stepper
        if f0_host.is_cuda or f1_host.is_cuda:
            raise ValueError("HostStreamedStepper: populations must be host tensors (use the stepper itself for device fields)")
        if not (f0_host.is_pinned() and f1_host.is_pinned()):
            raise ValueError("HostStreamedStepper: host populations must be pinned (torch.empty(..., pin_memory=True))")
        if f0_host.dtype != st.store_dtype or f0_host.shape != f1_host.shape or not f0_host.is_contiguous() or not f1_host.is_contiguous():
            raise ValueError("HostStreamedStepper: f_0 / f_1 must be contiguous, same shape, store dtype")
        s = self._state if (self._state and self._state["shape"] == tuple(f0_host.shape)) else self._setup(f0_host)
        q, nx, ny, nz = f0_host.shape
        P = s["P"]
        cur = torch.cuda.current_stream()
        h2d, d2h = s["streams"]["h2d"], s["streams"]["d2h"]
        vs = st.velocity_set
        plane = ny * nz
        esz_mask = 1
        h2d.wait_stream(cur)
        d2h.wait_stream(cur)
        for k, (a, b) in enumerate(s["chunks"]):
            i = k % 2
            n = b - a
            # ---- stage in (buffer i is free once the kernel of chunk k-2 has consumed it)
            with torch.cuda.stream(h2d):
                if k >= 2:
                    h2d.wait_event(s["ev_done"][i])
                self._stage(f0_host, s["bufs_in"][i], a, b, h2d)
                s["ev_in"][i].record(h2d)
            # ---- compute (output buffer i is free once chunk k-2 has been copied out)
            cur.wait_event(s["ev_in"][i])
            if k >= 2:
                cur.wait_event(s["ev_out"][i])
            dom = _lib.XlbDomain()
            dom.d, dom.q = vs.d, vs.q
            dom.nx, dom.ny, dom.nz = P + 2, ny, nz
            dom.x_periodic = 0
            dom.compute_dtype = st.precision_policy.compute_precision.abi_code
            dom.store_dtype = st.precision_policy.store_precision.abi_code
            dom.q_stride = (P + 2) * plane
            dom.mask_q_stride = nx * plane
            n_bc = s["n_bc"]
            off = (a - 1) * plane * esz_mask  # chunk-local plane 0 == global plane a-1 (never dereferenced for ghosts)
            rc = lib().xlb_step_ex(
                byref(dom), st.collision.abi_model, byref(s["cp"]), s["bufs_in"][i].data_ptr(), s["bufs_out"][i].data_ptr(),
                (bc_mask.data_ptr() + off) if n_bc else None, (missing_mask.data_ptr() + off) if n_bc else None,
                s["table"].data_ptr() if n_bc else None, n_bc, float(omega), int(timestep), 1, n + 1, st.kernel_variant,
                _lib.c_void_p(cur.cuda_stream),
            )
            check(rc, "HostStreamedStepper")
            s["ev_done"][i].record(cur)
            # ---- stage out
            with torch.cuda.stream(d2h):
                d2h.wait_event(s["ev_done"][i])
                self._copy_planes(f1_host, a, s["bufs_out"][i], 1, n, d2h)
                s["ev_out"][i].record(d2h)
        cur.wait_stream(d2h)
        cur.wait_stream(h2d)
        return f0_host, f1_host
