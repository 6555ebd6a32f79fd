"""Slab domain decomposition over the GPUs of one node (xlb/distribute/distribute.py:10-105 of the reference).

The reference shards every field along x over all devices (jax_grid.py:22-36) and, after a local periodic roll,
repairs the two wrapped planes with a periodic ``lax.ppermute`` ring that moves only the populations crossing the
cut: those with c_x = +1 (``velocity_set.right_indices``) to the right neighbour, those with c_x = -1
(``left_indices``) to the left one (distribute.py:23-44).

Here: one process per GPU (torchrun), rank r owns global planes [x0, x0 + n) and stores them with one ghost plane on
each side, ``[q, n + 2, ny, nz]`` (2-D: ghost rows, ``[q, n + 2, ny]``).  x is the slowest spatial axis, so a
population's halo plane is one contiguous run: no pack / unpack kernels.

The exchange is FUSED INTO THE STEP KERNEL (``halo="peer"``, the default on GPUs): every rank maps its ring neighbours'
population buffers into its own address space (CUDA IPC over NVLink / NVSwitch, ``xlb_ipc_*``), and the launch that
updates the two edge planes of the slab stores the outgoing populations not only into its own ``f_1`` but straight into
the neighbours' ghost planes (``XlbHaloPush``; the kernel's PUSH instance).  One time step on a rank is then

    edge branch (high priority) : xlb_halo_wait -> step(first + last owned plane, one launch, push) -> xlb_halo_signal
    main branch                 : step(interior planes)                                             -> join

with no collective call, no communication kernel and no host synchronisation: the NVLink stores are posted while the
interior is being updated, ``xlb_halo_signal`` publishes "my push number k has landed" to the neighbours' flags and
``xlb_halo_wait`` makes the next edge launch wait for both neighbours' pushes (which, because a rank signals only
after its own edge launch, also orders every push after the neighbour's last read of the ghost planes it overwrites).
The whole step is kernels only, so each buffer parity is captured once into a CUDA graph and replayed (``graph=True``).

``halo="nccl"`` keeps the collective version -- ncclSend/Recv of the 2 x n_out edge planes on a communication stream,
overlapped with the interior launch -- as the baseline and for buffers that were not set up by ``prepare_fields``; the
same ``HaloExchange`` object drives gloo in the CPU tests.  The ring is periodic like the reference's (rank P-1 <->
rank 0): global streaming is periodic, walls are boundary conditions.
"""

import ctypes

import numpy as np
import torch
import torch.distributed as dist

from .. import _lib
from ..compute_backend import ComputeBackend
from ..default_config import DefaultConfig


class SlabDecomposition:
    """Which x planes a rank owns, and who its ring neighbours are."""

    def __init__(self, rank=None, world=None, group=None, planes=None):
        """``planes``: owned x planes per rank (a list of ``world`` counts summing to nx) instead of the even split -- for
        domains whose cost per plane is not uniform (a large obstacle makes the planes it crosses dearer: every rank waits for
        the slowest one each step); ``balanced_planes`` derives such a list from a per-plane cost."""
        self.group = group
        self.planes = None if planes is None else [int(n) for n in planes]
        if world is None:
            world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
        if rank is None:
            rank = dist.get_rank(group) if dist.is_available() and dist.is_initialized() else 0
        if not (0 <= rank < world):
            raise ValueError(f"rank {rank} outside world of size {world}")
        self.rank, self.world = int(rank), int(world)
        self.left = (self.rank - 1) % self.world   # periodic ring (distribute.py:23-24)
        self.right = (self.rank + 1) % self.world

    def local_range(self, nx, rank=None):
        """(first owned global plane, number of owned planes) of ``rank`` (default: this rank); remainders go to the
        lowest ranks."""
        rank = self.rank if rank is None else int(rank)
        if self.planes is not None:
            if len(self.planes) != self.world or sum(self.planes) != int(nx):
                raise ValueError(f"planes={self.planes}: need {self.world} counts that sum to nx = {nx}")
            n = self.planes[rank]
            if n < 2 and self.world > 1:
                raise ValueError(f"slab of {n} plane(s) on rank {rank}: need at least 2 owned planes per rank")
            return sum(self.planes[:rank]), n
        base, rem = divmod(int(nx), self.world)
        n = base + (1 if rank < rem else 0)
        x0 = rank * base + min(rank, rem)
        if n < 2 and self.world > 1:
            raise ValueError(f"slab of {n} plane(s) on rank {rank}: need at least 2 owned planes per rank")
        return x0, n

    def make_grid(self, global_shape, backend_config=None):
        """A ``CudaGrid`` describing this rank's slab of ``global_shape`` (2-D or 3-D; fields get one ghost plane / row per
        side of the first axis, the axis the reference shards, jax_grid.py:22-36)."""
        from ..grid.cuda_grid import CudaGrid

        global_shape = tuple(int(s) for s in global_shape)
        if self.world == 1:
            return CudaGrid(global_shape, backend_config=backend_config)
        if len(global_shape) not in (2, 3):
            raise ValueError("slab decomposition: 2-D (D2Q9) or 3-D (D3Q19 / D3Q27) lattices")
        x0, n = self.local_range(global_shape[0])
        grid = CudaGrid((n,) + global_shape[1:], backend_config=backend_config)
        grid.global_shape = global_shape
        grid.nDevices = self.world
        grid.rank = self.rank
        grid.x_offset = x0
        grid.halo = 1
        grid.decomposition = self
        return grid


def balanced_planes(plane_cost, world, min_planes=2):
    """Owned planes per rank such that every rank gets about the same share of ``sum(plane_cost)`` (one relative cost per
    global x plane).  Contiguous slabs, at least ``min_planes`` each; the cut after rank r is the first plane at which the
    running cost reaches (r + 1) / world of the total."""
    cost = np.asarray(plane_cost, dtype=np.float64).reshape(-1)
    nx, world = cost.size, int(world)
    if world < 1 or nx < world * min_planes or np.any(cost <= 0):
        raise ValueError("balanced_planes: need positive costs and at least min_planes planes per rank")
    acc = np.cumsum(cost)
    cuts = [0]
    for r in range(1, world):
        x = int(np.searchsorted(acc, acc[-1] * r / world, side="left")) + 1
        x = max(x, cuts[-1] + min_planes)          # room for this rank ...
        x = min(x, nx - (world - r) * min_planes)  # ... and for the ones to come
        cuts.append(x)
    cuts.append(nx)
    return [cuts[r + 1] - cuts[r] for r in range(world)]


class HaloExchange:
    """The periodic ring exchange of the outgoing populations of the two edge planes (distribute.py:26-44) as a
    collective: batched send / recv (NCCL on GPUs, gloo in the CPU tests).  Used for the initial state, by the
    ``halo="nccl"`` schedule, and as the fallback for buffers the peer mapping does not know."""

    def __init__(self, decomposition, velocity_set):
        self.dec = decomposition
        self.right_indices = [int(i) for i in velocity_set.right_indices]  # c_x = +1 -> travel to the right neighbour
        self.left_indices = [int(i) for i in velocity_set.left_indices]    # c_x = -1 -> travel to the left neighbour
        self._ops = {}  # buffer address -> P2POp list (views and ops are built once per buffer, not per step)

    def _build_ops(self, f):
        n = f.shape[1] - 2  # owned planes
        dec = self.dec
        ops = []
        # order matters when left == right (world == 2): messages between one pair match in issue order
        for l in self.right_indices:
            ops.append(dist.P2POp(dist.isend, f[l, n], dec.right, group=dec.group))       # my last owned plane
        for l in self.left_indices:
            ops.append(dist.P2POp(dist.isend, f[l, 1], dec.left, group=dec.group))        # my first owned plane
        for l in self.right_indices:
            ops.append(dist.P2POp(dist.irecv, f[l, 0], dec.left, group=dec.group))        # left ghost plane
        for l in self.left_indices:
            ops.append(dist.P2POp(dist.irecv, f[l, n + 1], dec.right, group=dec.group))   # right ghost plane
        return ops

    def bytes_per_step(self, f):
        plane = f[0, 0].numel() * f.element_size()
        return plane * (len(self.right_indices) + len(self.left_indices))

    def exchange(self, f):
        """Fill the ghost planes of ``f`` from the ring neighbours.  Asynchronous on the current CUDA stream
        (NCCL); blocking for CPU tensors (gloo, used by the CPU tests)."""
        if self.dec.world == 1:
            n = f.shape[1] - 2
            f[self.right_indices, 0] = f[self.right_indices, n]
            f[self.left_indices, n + 1] = f[self.left_indices, 1]
            return
        key = (f.data_ptr(), tuple(f.shape), f.dtype)
        hit = self._ops.get(key)
        if hit is None or hit[0] is not f:  # keep the tensor alive with its views; a recycled address gets fresh ops
            hit = (f, self._build_ops(f))
            self._ops[key] = hit
        reqs = dist.batch_isend_irecv(hit[1])
        for r in reqs:
            r.wait()


class PeerHalo:
    """Peer mapping of the ring neighbours' population buffers and flags (CUDA IPC), and the ``XlbHaloPush`` /
    signal / wait arguments derived from it.  Built once per ``prepare_fields``; all ranks must build it together."""

    N_WORDS = 8  # sync words per rank: [0] pushes sent, [1] pushes waited for, [2] wait timed out, [4] flag from left, [5] flag from right

    def __init__(self, dec, buffers, ghost_dst):
        """``buffers``: this rank's population buffers (same order on every rank).  ``ghost_dst``: (ghost plane of the LEFT
        neighbour that receives my first owned plane, ghost plane of the RIGHT neighbour that receives my last one)."""
        lib = _lib.lib()
        self.dec = dec
        self.buffers = list(buffers)
        self.ghost_dst = (int(ghost_dst[0]), int(ghost_dst[1]))
        dev = buffers[0].device
        self.sync = torch.zeros(self.N_WORDS, dtype=torch.int64, device=dev)
        mine = [self.sync] + self.buffers
        # per tensor: 64-byte IPC handle, byte offset inside its allocation, population stride (elements; ranks whose slabs
        # differ in thickness have different strides, and the pushing kernel needs the RECEIVER's)
        blob = torch.zeros((len(mine), 80), dtype=torch.uint8)
        for i, t in enumerate(mine):
            handle = (ctypes.c_ubyte * 64)()
            offset = ctypes.c_int64(0)
            _lib.check(lib.xlb_ipc_export(t.data_ptr(), handle, ctypes.byref(offset)), "xlb_ipc_export")
            blob[i, :64] = torch.frombuffer(bytearray(handle), dtype=torch.uint8)
            blob[i, 64:] = torch.tensor([offset.value, t.stride(0) if t.dim() > 1 else 0], dtype=torch.int64).view(torch.uint8)
        blob = blob.to(dev)
        everyone = [torch.empty_like(blob) for _ in range(dec.world)]
        dist.all_gather(everyone, blob, group=dec.group)
        self._bases = {}  # handle bytes -> mapped base address (one cudaIpcOpenMemHandle per distinct allocation)
        self.left, self.left_qs = self._map(everyone[dec.left].cpu())
        self.right, self.right_qs = self._map(everyone[dec.right].cpu())
        dist.barrier(group=dec.group)  # nobody starts pushing before every mapping exists

    def _map(self, blob):
        lib = _lib.lib()
        ptrs, strides = [], []
        for row in blob:
            handle = bytes(row[:64].tolist())
            offset, stride = (int(v) for v in row[64:].view(torch.int64).tolist())
            strides.append(stride)
            base = self._bases.get(handle)
            if base is None:
                dptr, b = ctypes.c_void_p(), ctypes.c_void_p()
                _lib.check(lib.xlb_ipc_import(handle, 0, ctypes.byref(dptr), ctypes.byref(b)), "xlb_ipc_import")
                base = self._bases[handle] = b.value
            ptrs.append(base + offset)
        return ptrs, strides  # [sync, buffer 0, buffer 1, ...] of that neighbour

    def index_of(self, f):
        for i, b in enumerate(self.buffers):
            if b.data_ptr() == f.data_ptr():
                return i
        return None

    def push_args(self, f_1, first_owned, last_owned):
        """``XlbHaloPush`` for a step that writes ``f_1`` (one of the mapped buffers)."""
        i = self.index_of(f_1)
        if i is None:
            return None
        h = _lib.XlbHaloPush()
        h.peer_f1[0], h.peer_f1[1] = self.left[1 + i], self.right[1 + i]
        h.peer_q_stride[0], h.peer_q_stride[1] = self.left_qs[1 + i], self.right_qs[1 + i]
        h.src[0], h.src[1] = int(first_owned), int(last_owned)
        h.dst[0], h.dst[1] = self.ghost_dst
        return h

    def signal(self):
        """After the edge launch: publish my push count to both neighbours (their flag "from right" / "from left")."""
        me = self.sync.data_ptr()
        _lib.check(_lib.lib().xlb_halo_signal(me + 0, self.left[0] + 8 * 5, self.right[0] + 8 * 4, _lib.current_stream_ptr()), "xlb_halo_signal")

    def wait(self):
        """Before the edge launch: both neighbours' pushes of the previous step have landed in my ghost planes."""
        me = self.sync.data_ptr()
        _lib.check(_lib.lib().xlb_halo_wait(me + 8 * 1, me + 8 * 4, me + 8 * 5, _lib.current_stream_ptr()), "xlb_halo_wait")

    def timed_out(self):
        """True if a wait ever gave up (a neighbour stopped stepping); synchronises."""
        return bool(self.sync[2].item() != 0)

    def close(self):
        for base in self._bases.values():
            _lib.lib().xlb_ipc_close(base)
        self._bases = {}


class DistributedStepper:
    """``distribute(stepper, ...)`` result: same call signature as the stepper, one slab per rank."""

    def __init__(self, stepper, grid, velocity_set, overlap=True, halo="auto", graph=True):
        self.stepper = stepper
        self.grid = grid
        self.velocity_set = velocity_set
        self.overlap = bool(overlap)
        self.dec = getattr(grid, "decomposition", None) or SlabDecomposition(0, 1)
        self.halo = HaloExchange(self.dec, velocity_set)
        if halo not in ("auto", "peer", "nccl"):
            raise ValueError("distribute: halo must be 'auto', 'peer' or 'nccl'")
        self.halo_mode = halo
        self.use_graph = bool(graph)
        self.peer = None
        self._side = None       # second stream of the fork (edge branch for the peer schedule, comm stream for NCCL)
        self._ev_fork = self._ev_join = None
        self._graphs = {}       # (buffers, omega) -> [calls seen, CUDAGraph or None]
        self.launches_per_step = None

    def __getattr__(self, name):  # boundary_conditions, collision, macroscopic, ... of the wrapped stepper
        return getattr(self.stepper, name)

    # ---- setup
    def prepare_fields(self, initializer=None):
        f_0, f_1, bc_mask, missing_mask = self.stepper.prepare_fields(initializer)
        if self.dec.world > 1:
            self._exchange_any(f_0)
            torch.cuda.current_stream().synchronize()
            if self.halo_mode in ("auto", "peer") and f_0.is_cuda:
                self._setup_peer(f_0, f_1)
        return f_0, f_1, bc_mask, missing_mask

    def exchange(self, f):
        """Collective ghost fill (NCCL / gloo ring) of a 3-D ``[q, n+2, ny, nz]`` or 2-D ``[q, n+2, ny]`` field: what
        ``prepare_fields`` does for the initial state; call it again after overwriting the owned cells of ``f_0`` by hand."""
        self.halo.exchange(f if f.dim() == 4 else f.unsqueeze(-1))

    _exchange_any = exchange

    def _setup_peer(self, f_0, f_1):
        if self.peer is not None:
            self.peer.close()
            self.peer = None
        self._graphs = {}
        n_global = self.grid.global_shape[0]
        n_left = self.dec.local_range(n_global, self.dec.left)[1]
        try:
            self.peer = PeerHalo(self.dec, [f_0, f_1], ghost_dst=(n_left + 1, 0))
        except (RuntimeError, NotImplementedError, ValueError) as exc:
            ok = 0
            err = exc
        else:
            ok = 1
            err = None
        flag = torch.tensor([ok], device=f_0.device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=self.dec.group)  # all ranks take the same path
        if not bool(flag.item()):
            if self.peer is not None:
                self.peer.close()
                self.peer = None
            if self.halo_mode == "peer":
                raise RuntimeError(f"distribute: peer mapping of the neighbours' buffers failed ({err}); use halo='nccl'")

    # ---- one step
    def __call__(self, f_0, f_1, bc_mask, missing_mask, omega, timestep):
        if self.dec.world == 1:
            return self.stepper(f_0, f_1, bc_mask, missing_mask, omega, timestep)
        if self.peer is not None and self.peer.index_of(f_1) is not None and self.peer.index_of(f_0) is not None:
            if not self.use_graph:
                self._step_peer(f_0, f_1, bc_mask, missing_mask, omega, timestep)
                return f_0, f_1
            key = (f_0.data_ptr(), f_1.data_ptr(), bc_mask.data_ptr(), missing_mask.data_ptr(), float(omega))
            slot = self._graphs.setdefault(key, [0, None])
            slot[0] += 1
            if slot[1] is not None:
                slot[1].replay()
            elif slot[0] < 2 or torch.cuda.is_current_stream_capturing():
                self._step_peer(f_0, f_1, bc_mask, missing_mask, omega, timestep)  # first call: builds plans / lists eagerly
            else:
                g = torch.cuda.CUDAGraph()
                c0 = _lib.launch_count()
                with torch.cuda.graph(g):
                    self._step_peer(f_0, f_1, bc_mask, missing_mask, omega, timestep)
                self.launches_per_step = _lib.launch_count() - c0
                slot[1] = g
                g.replay()  # a capture does not execute: this replay IS the step
            return f_0, f_1
        self._step_nccl(f_0, f_1, bc_mask, missing_mask, omega, timestep)
        return f_0, f_1

    def _ranges(self, f_0):
        """(edge range, interior range or None, first owned, last owned) along the slab axis."""
        n = f_0.shape[1] - 2
        if self.velocity_set.d == 3:
            edges = _lib.XlbStepRange(1, max(n - 1, 1), 2 if n > 1 else 1, 0, 1, 0)
            inner = _lib.XlbStepRange(2, 1, n - 2, 0, 1, 0) if n > 2 else None
        else:  # D2Q9 runs as nx = 1: the user's x axis is the kernel's y axis
            edges = _lib.XlbStepRange(0, 1, 1, 1, max(n - 1, 1), 2 if n > 1 else 1)
            inner = _lib.XlbStepRange(0, 1, 1, 2, 1, n - 2) if n > 2 else None
        return edges, inner, 1, n

    def _fork(self):
        if self._side is None:
            self._side = torch.cuda.Stream(priority=-1)  # the edge branch / the exchange goes first
            self._ev_fork = torch.cuda.Event()
            self._ev_join = torch.cuda.Event()
        return self._side

    def _step_peer(self, f_0, f_1, bc_mask, missing_mask, omega, timestep):
        st, peer = self.stepper, self.peer
        edges, inner, first, last = self._ranges(f_0)
        push = peer.push_args(f_1, first, last)
        cur = torch.cuda.current_stream()
        if not self.overlap or inner is None:
            peer.wait()
            st.launch(f_0, f_1, bc_mask, missing_mask, omega, timestep, edges, push)
            peer.signal()
            if inner is not None:
                st.launch(f_0, f_1, bc_mask, missing_mask, omega, timestep, inner)
            return
        side = self._fork()
        self._ev_fork.record(cur)
        side.wait_event(self._ev_fork)
        with torch.cuda.stream(side):
            peer.wait()
            st.launch(f_0, f_1, bc_mask, missing_mask, omega, timestep, edges, push)
            peer.signal()
            self._ev_join.record(side)
        st.launch(f_0, f_1, bc_mask, missing_mask, omega, timestep, inner)
        cur.wait_event(self._ev_join)

    def _step_nccl(self, f_0, f_1, bc_mask, missing_mask, omega, timestep):
        st = self.stepper
        edges, inner, _, _ = self._ranges(f_0)
        cur = torch.cuda.current_stream()
        if not self.overlap or inner is None or not f_0.is_cuda:
            st.launch(f_0, f_1, bc_mask, missing_mask, omega, timestep)
            self._exchange_any(f_1)
            return
        side = self._fork()
        st.launch(f_0, f_1, bc_mask, missing_mask, omega, timestep, edges)
        self._ev_fork.record(cur)
        side.wait_event(self._ev_fork)
        with torch.cuda.stream(side):
            self._exchange_any(f_1)
            self._ev_join.record(side)
        st.launch(f_0, f_1, bc_mask, missing_mask, omega, timestep, inner)
        cur.wait_event(self._ev_join)

    def close(self):
        """Unmap the neighbours' buffers (collective: every rank must call it before any rank frees its fields)."""
        if self.peer is not None:
            torch.cuda.synchronize()
            dist.barrier(group=self.dec.group)  # every push into my buffers has landed, every neighbour is done with the mapping
            self.peer.close()
            self.peer = None
        self._graphs = {}

    # ---- helpers for tests / post-processing
    def owned(self, field):
        """View of the owned planes (ghosts stripped)."""
        return field if self.dec.world == 1 else field[:, 1:-1]

    def gather(self, field):
        """All ranks' owned planes concatenated along x (on every rank)."""
        own = self.owned(field).contiguous()
        if self.dec.world == 1:
            return own
        sizes = [self.dec.local_range(self.grid.global_shape[0], r)[1] for r in range(self.dec.world)]
        parts = [torch.empty((own.shape[0], s) + tuple(own.shape[2:]), dtype=own.dtype, device=own.device) for s in sizes]
        if own.dtype == torch.bool:
            parts = [p.to(torch.uint8) for p in parts]
            dist.all_gather(parts, own.to(torch.uint8), group=self.dec.group)
            return torch.cat(parts, dim=1).to(torch.bool)
        dist.all_gather(parts, own, group=self.dec.group)
        return torch.cat(parts, dim=1)


def distribute(operator, grid, velocity_set, num_results=1, ops="permute", overlap=True, halo="auto", graph=True):
    """Same call as the reference's ``distribute`` (distribute.py:82-105).  ``operator`` must be an
    ``IncompressibleNavierStokesStepper`` built on a grid from ``SlabDecomposition.make_grid`` (or any grid when
    running on one rank).  ``num_results`` / ``ops`` are accepted for signature parity.  ``halo``: "peer" = the exchange
    fused into the edge-plane launch over peer-mapped memory, "nccl" = send / recv on a communication stream, "auto" =
    peer where the mapping succeeds on every rank.  ``graph``: replay each buffer parity of the peer schedule as a CUDA graph."""
    from ..operator.stepper import IncompressibleNavierStokesStepper

    if ops != "permute":
        raise NotImplementedError(f"distribute: ops={ops!r} (the reference implements only 'permute', distribute.py:31-41)")
    if not isinstance(operator, IncompressibleNavierStokesStepper):
        raise NotImplementedError("distribute: only IncompressibleNavierStokesStepper is distributed by the CUDA backend")
    if DefaultConfig.default_backend not in (None, ComputeBackend.CUDA):
        raise ValueError("distribute: CUDA backend only")
    return DistributedStepper(operator, grid, velocity_set, overlap=overlap, halo=halo, graph=graph)
