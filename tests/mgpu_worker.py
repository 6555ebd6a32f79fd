"""torchrun worker: N-rank x-slab run == single-GPU run of the same global problem, bitwise.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tests/mgpu_worker.py
"""

import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def main():
    from oracle import cases
    from tests.util import build_product, to_dev
    from xlb_b200.distribute import SlabDecomposition, distribute

    rank = int(os.environ["RANK"])
    local_rank = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local_rank)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    world = dist.get_world_size()
    todo = [
        (cases.cavity3d(0, shape=(8 * world, 12, 16), omega=1.7), 25),
        (cases.sphere3d((12 * world if world > 2 else 36, 24, 24)), 25),
        (cases.cavity3d(0, shape=(6 * world + 1, 10, 8), lattice="D3Q27", collision="KBC", omega=1.8, policy="FP64FP64"), 10),  # uneven slabs
        (cases.cavity3d(0, shape=(4 * world, 16, 16), policy="FP32FP16", omega=1.5), 10),
        (cases.sphere3d((12 * world if world > 2 else 36, 24, 24), collision="SmagorinskyLESBGK", omega=1.9, force=(2e-5, 0.0, 1e-5)), 10),
        # 2-D slabs (the reference's only shipped use of distribute: examples/cfd/lid_driven_cavity_2d_distributed.py:22)
        (cases.cavity2d(0, shape=(8 * world + 1, 24), omega=1.6), 25),
        (cases.cavity2d(0, shape=(6 * world, 32), collision="KBC", policy="FP64FP64", omega=1.7), 10),
    ]

    def uneven(nx):
        """Plane counts that differ by 4 between neighbours (what a cost-balanced decomposition hands out, balanced_planes)."""
        planes = [nx // world] * world
        planes[0] += nx - sum(planes)
        for r in range(0, world - 1, 2):
            planes[r] += 2
            planes[r + 1] -= 2
        return planes

    sphere = cases.sphere3d((12 * world if world > 2 else 36, 24, 24), lattice="D3Q27", collision="KBC", omega=1.7)
    todo.append((sphere, 10, uneven(sphere["shape"][0])))
    # (halo exchange, overlap schedule, CUDA-graph replay): the fused peer push (default), the same serialised and eager,
    # and the NCCL send/recv schedule
    modes = [("peer", True, True), ("peer", False, False), ("nccl", True, False)]
    for case, nsteps, *rest in todo:
        planes = rest[0] if rest else None
        for halo, overlap, graph in modes:
            # single-GPU reference of the global problem, on every rank
            st, _ = build_product(case)
            f_0, f_1, bc_mask, missing = st.prepare_fields()
            f_0 = to_dev(cases.random_state(case, seed=3))
            for i in range(nsteps):
                f_0, f_1 = st(f_0, f_1, bc_mask, missing, case["omega"], i)
                f_0, f_1 = f_1, f_0
            ref, ref_bc, ref_mm = f_0.clone(), bc_mask.clone(), missing.clone()
            # distributed run
            dec = SlabDecomposition(planes=planes)
            # with the overlap schedule the slab stepper also runs the BC-cell-list path (forced: these blocks are far
            # below the size where "auto" would pick it), so its per-plane slicing is covered by the bitwise comparison
            st, _ = build_product(case, backend_config={"bc_cell_list": True} if overlap else None, decomposition=dec)
            dst = distribute(st, st.grid, st.velocity_set, overlap=overlap, halo=halo, graph=graph)
            g_0, g_1, g_bc, g_mm = dst.prepare_fields()
            x0, n = dec.local_range(case["shape"][0])
            assert torch.equal(dst.owned(g_bc), ref_bc[:, x0 : x0 + n]), "bc_mask slab differs"
            assert torch.equal(dst.owned(g_mm), ref_mm[:, x0 : x0 + n]), "missing_mask slab differs"
            init = to_dev(cases.random_state(case, seed=3))
            g_0[:, 1:-1] = init[:, x0 : x0 + n]
            dst.exchange(g_0)
            assert (dst.peer is not None) == (halo == "peer"), "peer mapping not established"
            for i in range(nsteps):
                g_0, g_1 = dst(g_0, g_1, g_bc, g_mm, case["omega"], i)
                g_0, g_1 = g_1, g_0
            torch.cuda.synchronize()
            same = torch.equal(dst.owned(g_0), ref[:, x0 : x0 + n])
            full = dst.gather(g_0)
            same_all = torch.equal(full, ref)
            if not (same and same_all):
                d = (dst.owned(g_0).double() - ref[:, x0 : x0 + n].double()).abs().max().item()
                raise AssertionError(f"rank {rank}: {case['name']} halo={halo} overlap={overlap}: distributed != single GPU (max abs diff {d:.3e})")
            if halo == "peer":
                assert not dst.peer.timed_out()
                if graph:
                    assert dst.launches_per_step is not None and len(dst._graphs) == 2, "the peer schedule was not replayed from CUDA graphs"
            # MomentumTransfer on the obstacle: slab partial sums all-reduced == single-GPU value (to FP64 summation order)
            from xlb_b200.operator.boundary_condition import HalfwayBounceBackBC
            from xlb_b200.operator.force import MomentumTransfer

            solids = [b for b in st.boundary_conditions if isinstance(b, HalfwayBounceBackBC)]
            if solids and overlap and halo == "peer" and len(case["shape"]) == 3:
                st1, _ = build_product(case)
                solid1 = [b for b in st1.boundary_conditions if isinstance(b, HalfwayBounceBackBC)][-1]
                f_ref = MomentumTransfer(solid1)(ref, f_1, ref_bc, ref_mm)
                f_dist = MomentumTransfer(solids[-1], grid=st.grid)(g_0, g_1, g_bc, g_mm)
                err = (f_ref.double() - f_dist.double()).abs().max().item() / f_ref.double().abs().max().item()
                assert err < 1e-6, f"rank {rank}: MomentumTransfer distributed {f_dist.tolist()} != single {f_ref.tolist()}"
            dst.close()
            if rank == 0:
                print(f"OK {case['name']} world={world} halo={halo} overlap={overlap} graph={graph} planes={planes or 'even'} steps={nsteps}: bitwise equal to the single-GPU run", flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
