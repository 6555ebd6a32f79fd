"""x-slab domain decomposition with an NCCL halo ring (xlb/distribute/distribute.py:10-105 of the reference).

The reference shards every field along x over all devices (jax_grid.py:22-36) and, after a local periodic roll,
repairs the two wrapped planes with a periodic ``lax.ppermute`` ring that moves only the populations crossing the
cut: those with c_x = +1 (``velocity_set.right_indices``) to the right neighbour, those with c_x = -1
(``left_indices``) to the left one (distribute.py:23-44).

Here: one process per GPU (torchrun), rank r owns global planes [x0, x0 + n) and stores them with one ghost plane on
each side, ``[q, n + 2, ny, nz]``.  Because x is the slowest spatial axis, a population's halo plane is one contiguous
run of ny*nz elements: no pack / unpack kernels.  One time step on a rank:

    compute stream : step(first owned plane), step(last owned plane) --event--> step(interior planes) ----.
    comm stream    :                                   wait --> ncclSend/Recv of the 2 x n_out planes ----+--> join

i.e. the outgoing planes of f_1 travel over NVLink into the neighbours' ghost planes of f_1 while the interior is
being updated (``overlap=True``; with ``overlap=False`` the exchange is serialised after one whole-slab launch).
The ring is periodic like the reference's (rank P-1 <-> rank 0): global streaming is periodic, walls are BCs.
"""

import torch
import torch.distributed as dist

from ..compute_backend import ComputeBackend
from ..default_config import DefaultConfig


class SlabDecomposition:
    """Which x planes a rank owns, and who its ring neighbours are."""

    def __init__(self, rank=None, world=None, group=None):
        self.group = group
        if world is None:
            world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
        if rank is None:
            rank = dist.get_rank(group) if dist.is_available() and dist.is_initialized() else 0
        if not (0 <= rank < world):
            raise ValueError(f"rank {rank} outside world of size {world}")
        self.rank, self.world = int(rank), int(world)
        self.left = (self.rank - 1) % self.world   # periodic ring (distribute.py:23-24)
        self.right = (self.rank + 1) % self.world

    def local_range(self, nx):
        """(first owned global plane, number of owned planes); remainders go to the lowest ranks."""
        base, rem = divmod(int(nx), self.world)
        n = base + (1 if self.rank < rem else 0)
        x0 = self.rank * base + min(self.rank, rem)
        if n < 2 and self.world > 1:
            raise ValueError(f"slab of {n} plane(s) on rank {self.rank}: need at least 2 owned planes per rank")
        return x0, n

    def make_grid(self, global_shape, backend_config=None):
        """A ``CudaGrid`` describing this rank's slab of ``global_shape`` (fields get one ghost plane per side)."""
        from ..grid.cuda_grid import CudaGrid

        global_shape = tuple(int(s) for s in global_shape)
        if self.world == 1:
            return CudaGrid(global_shape, backend_config=backend_config)
        if len(global_shape) != 3:
            raise NotImplementedError("slab decomposition is implemented for 3-D lattices (D3Q19 / D3Q27)")
        x0, n = self.local_range(global_shape[0])
        grid = CudaGrid((n,) + global_shape[1:], backend_config=backend_config)
        grid.global_shape = global_shape
        grid.nDevices = self.world
        grid.rank = self.rank
        grid.x_offset = x0
        grid.halo = 1
        grid.decomposition = self
        return grid


class HaloExchange:
    """The periodic ring exchange of the outgoing populations of the two edge planes (distribute.py:26-44)."""

    def __init__(self, decomposition, velocity_set):
        self.dec = decomposition
        self.right_indices = [int(i) for i in velocity_set.right_indices]  # c_x = +1 -> travel to the right neighbour
        self.left_indices = [int(i) for i in velocity_set.left_indices]    # c_x = -1 -> travel to the left neighbour

    def _ops(self, f):
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
        reqs = dist.batch_isend_irecv(self._ops(f))
        for r in reqs:
            r.wait()


class DistributedStepper:
    """``distribute(stepper, ...)`` result: same call signature as the stepper, one slab per rank."""

    def __init__(self, stepper, grid, velocity_set, overlap=True):
        self.stepper = stepper
        self.grid = grid
        self.velocity_set = velocity_set
        self.overlap = bool(overlap)
        self.dec = getattr(grid, "decomposition", None) or SlabDecomposition(0, 1)
        self.halo = HaloExchange(self.dec, velocity_set)
        self._comm_stream = None
        self._ev_edges = None
        self._ev_comm = None

    def __getattr__(self, name):  # boundary_conditions, collision, macroscopic, ... of the wrapped stepper
        return getattr(self.stepper, name)

    def prepare_fields(self, initializer=None):
        f_0, f_1, bc_mask, missing_mask = self.stepper.prepare_fields(initializer)
        if self.dec.world > 1:
            self.halo.exchange(f_0)
            torch.cuda.current_stream().synchronize()
        return f_0, f_1, bc_mask, missing_mask

    def __call__(self, f_0, f_1, bc_mask, missing_mask, omega, timestep):
        st = self.stepper
        if self.dec.world == 1:
            return st(f_0, f_1, bc_mask, missing_mask, omega, timestep)
        n = f_0.shape[1] - 2
        cur = torch.cuda.current_stream()
        if not self.overlap:
            st.launch_range(f_0, f_1, bc_mask, missing_mask, omega, timestep, 1, n + 1)
            self.halo.exchange(f_1)
            return f_0, f_1
        if self._comm_stream is None:
            self._comm_stream = torch.cuda.Stream()
            self._ev_edges = torch.cuda.Event()
            self._ev_comm = torch.cuda.Event()
        st.launch_range(f_0, f_1, bc_mask, missing_mask, omega, timestep, 1, 2)
        st.launch_range(f_0, f_1, bc_mask, missing_mask, omega, timestep, n, n + 1)
        self._ev_edges.record(cur)
        self._comm_stream.wait_event(self._ev_edges)
        with torch.cuda.stream(self._comm_stream):
            self.halo.exchange(f_1)
            self._ev_comm.record(self._comm_stream)
        if n > 2:
            st.launch_range(f_0, f_1, bc_mask, missing_mask, omega, timestep, 2, n)
        cur.wait_event(self._ev_comm)
        return f_0, f_1

    # ---- helpers for tests / post-processing
    def owned(self, field):
        """View of the owned planes (ghosts stripped)."""
        return field if self.dec.world == 1 else field[:, 1:-1]

    def gather(self, field):
        """All ranks' owned planes concatenated along x (on every rank)."""
        own = self.owned(field).contiguous()
        if self.dec.world == 1:
            return own
        sizes = [SlabDecomposition(r, self.dec.world).local_range(self.grid.global_shape[0])[1] for r in range(self.dec.world)]
        parts = [torch.empty((own.shape[0], s) + tuple(own.shape[2:]), dtype=own.dtype, device=own.device) for s in sizes]
        if own.dtype == torch.bool:
            parts = [p.to(torch.uint8) for p in parts]
            dist.all_gather(parts, own.to(torch.uint8), group=self.dec.group)
            return torch.cat(parts, dim=1).to(torch.bool)
        dist.all_gather(parts, own, group=self.dec.group)
        return torch.cat(parts, dim=1)


def distribute(operator, grid, velocity_set, num_results=1, ops="permute", overlap=True):
    """Same call as the reference's ``distribute`` (distribute.py:82-105).  ``operator`` must be an
    ``IncompressibleNavierStokesStepper`` built on a grid from ``SlabDecomposition.make_grid`` (or any grid when
    running on one rank).  ``num_results`` / ``ops`` are accepted for signature parity."""
    from ..operator.stepper import IncompressibleNavierStokesStepper

    if ops != "permute":
        raise NotImplementedError(f"distribute: ops={ops!r} (the reference implements only 'permute', distribute.py:31-41)")
    if not isinstance(operator, IncompressibleNavierStokesStepper):
        raise NotImplementedError("distribute: only IncompressibleNavierStokesStepper is distributed by the CUDA backend")
    if DefaultConfig.default_backend not in (None, ComputeBackend.CUDA):
        raise ValueError("distribute: CUDA backend only")
    return DistributedStepper(operator, grid, velocity_set, overlap=overlap)
