"""Multi-rank plumbing on CPU (gloo, world_size 2 and 3): slab partition, periodic ring, halo contents.
The same HaloExchange object drives NCCL on the GPUs; the kernels are not involved here."""

import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import ops
from oracle.distribute import sharded_stream
from oracle.lattice import make_lattice


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, lattice, shape, out_dir, planes=None):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from xlb_b200.distribute import HaloExchange, SlabDecomposition

        lat = make_lattice(lattice)
        rng = np.random.default_rng(7)
        f = rng.uniform(0, 1, (lat.q,) + shape).astype(np.float32)
        dec = SlabDecomposition(planes=planes)
        assert (dec.rank, dec.world) == (rank, world)
        x0, n = dec.local_range(shape[0])
        if planes is not None:
            assert n == planes[rank] and x0 == sum(planes[:rank])
        slab = np.full((lat.q, n + 2) + shape[1:], np.nan, np.float32)
        slab[:, 1:-1] = f[:, x0 : x0 + n]
        t = torch.from_numpy(slab)
        HaloExchange(dec, lat).exchange(t)
        got = t.numpy()
        ri, li = lat.right_indices, lat.left_indices
        assert np.array_equal(got[ri, 0], f[ri, (x0 - 1) % shape[0]])
        assert np.array_equal(got[li, n + 1], f[li, (x0 + n) % shape[0]])
        # ghosts of the other populations stay untouched (only outgoing populations travel)
        others = [l for l in range(lat.q) if l not in set(ri.tolist())]
        assert np.isnan(got[others, 0]).all()
        # pull-stream of the owned planes from slab + ghosts == global periodic stream
        glob = ops.stream(lat, f)
        local = np.empty((lat.q, n) + shape[1:], np.float32)
        for l in range(lat.q):
            cx, cy, cz = (int(v) for v in lat.c[:, l])
            src = got[l, 1 - cx : 1 - cx + n]
            local[l] = np.roll(src, (cy, cz), axis=(1, 2))
        assert np.array_equal(local, glob[:, x0 : x0 + n])
        open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,lattice,shape", [(2, "D3Q19", (8, 5, 4)), (3, "D3Q27", (10, 4, 5)), (2, "D3Q27", (4, 3, 3))])
def test_halo_ring_gloo(world, lattice, shape, tmp_path):
    mp.spawn(_worker, args=(world, _free_port(), lattice, shape, str(tmp_path)), nprocs=world, join=True)
    assert all((tmp_path / f"ok{r}").exists() for r in range(world))


def test_halo_ring_gloo_with_given_planes(tmp_path):
    """Slabs of unequal plane counts (cost-balanced decomposition): same ring, same result."""
    world, planes = 3, [5, 2, 4]
    mp.spawn(_worker, args=(world, _free_port(), "D3Q19", (11, 4, 3), str(tmp_path), planes), nprocs=world, join=True)
    assert all((tmp_path / f"ok{r}").exists() for r in range(world))


def test_balanced_planes():
    from xlb_b200.distribute import SlabDecomposition, balanced_planes

    assert balanced_planes(np.ones(16), 8) == [2] * 8
    assert sum(balanced_planes(np.ones(17), 8)) == 17 and min(balanced_planes(np.ones(17), 8)) >= 2
    # an expensive region in the middle: the ranks that hold it get fewer planes, every share is within one plane's cost of even
    cost = np.ones(96)
    cost[40:56] = 4.0
    planes = balanced_planes(cost, 4)
    assert sum(planes) == 96 and min(planes) >= 2
    cuts = np.cumsum([0] + planes)
    shares = [cost[cuts[r] : cuts[r + 1]].sum() for r in range(4)]
    assert max(shares) - min(shares) <= 2 * cost.max() and max(shares) < 1.1 * cost.sum() / 4
    assert max(planes) > 96 // 4 > min(planes)
    # a cost so concentrated that the even-cost cut would leave a rank without planes: the minimum is kept
    spike = np.ones(12)
    spike[0] = 1000.0
    p = balanced_planes(spike, 4)
    assert sum(p) == 12 and min(p) >= 2
    with pytest.raises(ValueError):
        balanced_planes(np.ones(7), 4)
    # the decomposition takes the list as it is, and checks it against nx
    dec = SlabDecomposition(1, 4, planes=planes)
    assert dec.local_range(96) == (planes[0], planes[1])
    with pytest.raises(ValueError):
        dec.local_range(97)
    with pytest.raises(ValueError):
        SlabDecomposition(0, 2, planes=[1, 7]).local_range(8)


def _worker_2d(rank, world, port, shape, out_dir):
    """D2Q9 slabs: ghost ROWS of a [q, n+2, ny] field (examples/cfd/lid_driven_cavity_2d_distributed.py:22)."""
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from xlb_b200.distribute import HaloExchange, SlabDecomposition

        lat = make_lattice("D2Q9")
        f = np.random.default_rng(11).uniform(0, 1, (lat.q,) + shape).astype(np.float32)
        dec = SlabDecomposition()
        x0, n = dec.local_range(shape[0])
        slab = np.full((lat.q, n + 2, shape[1]), np.nan, np.float32)
        slab[:, 1:-1] = f[:, x0 : x0 + n]
        t = torch.from_numpy(slab)
        HaloExchange(dec, lat).exchange(t.unsqueeze(-1))  # what DistributedStepper.exchange does for 2-D fields
        got = t.numpy()
        glob = ops.stream(lat, f)
        local = np.empty((lat.q, n, shape[1]), np.float32)
        for l in range(lat.q):
            cx, cy = (int(v) for v in lat.c[:, l])
            local[l] = np.roll(got[l, 1 - cx : 1 - cx + n], cy, axis=1)
        assert np.array_equal(local, glob[:, x0 : x0 + n])
        open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,shape", [(2, (8, 6)), (3, (10, 5))])
def test_halo_ring_gloo_2d(world, shape, tmp_path):
    mp.spawn(_worker_2d, args=(world, _free_port(), shape, str(tmp_path)), nprocs=world, join=True)
    assert all((tmp_path / f"ok{r}").exists() for r in range(world))


def test_slab_partition():
    from xlb_b200.distribute import SlabDecomposition

    for world in (1, 2, 3, 4, 8):
        for nx in (16, 19, 64):
            got = [SlabDecomposition(r, world).local_range(nx) for r in range(world)]
            assert got[0][0] == 0 and sum(n for _, n in got) == nx
            for (a, n), (b, _) in zip(got, got[1:]):
                assert a + n == b
    d = SlabDecomposition(0, 4)
    assert (d.left, d.right) == (3, 1)
    assert d.local_range(19, rank=3) == (15, 4) and d.local_range(19, rank=0) == (0, 5)
    with pytest.raises(ValueError):
        SlabDecomposition(1, 8).local_range(8)


@pytest.mark.parametrize("lattice,shape,n", [("D3Q19", (8, 5, 4), 2), ("D3Q27", (12, 4, 3), 4)])
def test_oracle_ring_equals_global_stream(lattice, shape, n):
    """oracle/distribute.py (distribute.py:18-46 restated) == the global periodic roll."""
    lat = make_lattice(lattice)
    f = np.random.default_rng(1).uniform(0, 1, (lat.q,) + shape)
    assert np.array_equal(sharded_stream(lat, f, n), ops.stream(lat, f))


def test_bench_plane_cost_follows_the_obstacle_of_its_workloads():
    """bench.py hands the ranks of an obstacle workload slabs of equal COST: its per-plane cost must count the same sphere
    bench.build_case tags (same centre, same radius), and the resulting slabs must be lighter where the sphere is."""
    import bench
    from xlb_b200.distribute import balanced_planes

    for kind, shape in (("windtunnel", (256, 96, 96)), ("sphere", (240, 96, 96))):
        nx, ny, nz = shape
        r, cx = (min(ny // 8, nx // 4), nx // 3) if kind == "windtunnel" else (ny // 12, nx // 6)
        cy, cz = ny // 2, nz // 2
        y, z = np.meshgrid(np.arange(ny), np.arange(nz), indexing="ij")
        solid = np.array([np.count_nonzero((x - cx) ** 2 + (y - cy) ** 2 + (z - cz) ** 2 < r * r) for x in range(nx)])
        cost = bench.obstacle_plane_cost(kind, shape)
        assert cost.shape == (nx,) and np.all(cost >= ny * nz)
        extra = (cost - ny * nz) / (bench.HALFWAY_CELL_COST - 1.0)  # the model's solid cells per plane (pi * (r^2 - t^2))
        assert np.array_equal(extra > 0, solid > 0) or np.count_nonzero((extra > 0) != (solid > 0)) <= 2  # same planes, up to the two tips
        assert abs(extra.sum() - solid.sum()) <= 0.05 * solid.sum()
        planes = balanced_planes(cost, 8)
        cuts = np.cumsum([0] + planes)
        holder = int(np.searchsorted(cuts, cx, side="right") - 1)  # the rank that holds the sphere's centre plane
        assert planes[holder] == min(planes) and sum(planes) == nx
