"""torchrun worker: the reference's own distributed example, examples/cfd/lid_driven_cavity_2d_distributed.py (its only shipped
use of ``distribute``), run on N ranks with nothing changed but the backend it selects, against the same problem on one GPU.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tests/mgpu_example_worker.py

The script calls the plain ``grid_factory`` and ``distribute(stepper, grid, velocity_set)``: with torch.distributed initialised,
``grid_factory`` hands every rank its slab (what the reference's JAX grid does over its devices, jax_grid.py:22-36).
"""

import os
import sys
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def main():
    from oracle import ref_examples
    from tests.test_reference_examples import _exec, _switch_backend

    rank = int(os.environ["RANK"])
    local_rank = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local_rank)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    world = dist.get_world_size()

    import xlb_b200 as xlb
    import xlb_b200.compat as compat
    from xlb_b200.compute_backend import ComputeBackend
    from xlb_b200.operator.boundary_condition.boundary_condition_registry import boundary_condition_registry
    from xlb_b200.precision_policy import PrecisionPolicy

    compat.install_as_xlb()
    calls = []
    utils = types.ModuleType("xlb.utils")
    utils.save_image = lambda *a, **k: calls.append("save_image")
    utils.save_fields_vtk = lambda *a, **k: calls.append("save_fields_vtk")
    utils.warp_array_to_jax = lambda a: a
    sys.modules["xlb.utils"] = utils
    jnp = types.ModuleType("jax.numpy")
    jnp.ndarray = torch.Tensor
    jax = types.ModuleType("jax")
    jax.numpy = jnp
    sys.modules["jax"], sys.modules["jax.numpy"] = jax, jnp

    def src(name):
        path = ref_examples.path_of(name)
        if path is None:
            raise SystemExit(f"{name}: oracle/_ref/examples is empty (run __graft_entry__.build() where /root/reference is)")
        return _switch_backend(open(path).read())

    base_ns = _exec(src("lid_driven_cavity_2d.py"), "lid_driven_cavity_2d")
    base = types.ModuleType("lid_driven_cavity_2d")  # the distributed script imports its base class from the sibling file
    base.__dict__.update(base_ns)
    sys.modules["lid_driven_cavity_2d"] = base
    dist_ns = _exec(src("lid_driven_cavity_2d_distributed.py"), "lid_driven_cavity_2d_distributed")

    n, u_lid, steps = 16 * world + 3, 0.05, 60  # uneven slabs
    omega = 1.0 / (3.0 * (u_lid * (n - 1) / 200.0) + 0.5)
    pol = PrecisionPolicy.FP32FP32
    vs = xlb.velocity_set.D2Q9(precision_policy=pol, compute_backend=ComputeBackend.CUDA)

    boundary_condition_registry.next_id = 1
    sim = dist_ns["LidDrivenCavity2D_distributed"](omega, u_lid, (n, n), vs, ComputeBackend.CUDA, pol)
    assert sim.grid.halo == 1 and sim.grid.nDevices == world, "grid_factory did not hand out slabs under torch.distributed"
    assert sim.stepper.peer is not None, "the halo push over peer memory was not set up"
    sim.run(num_steps=steps, post_process_interval=30)
    torch.cuda.synchronize()
    assert calls.count("save_image") == 3
    got = sim.stepper.gather(sim.f_0)
    got_bc = sim.stepper.gather(sim.bc_mask)

    # the same problem as one block on this rank
    from xlb_b200.grid import grid_factory
    from xlb_b200.operator.boundary_condition import EquilibriumBC, HalfwayBounceBackBC
    from xlb_b200.operator.stepper import IncompressibleNavierStokesStepper

    boundary_condition_registry.next_id = 1
    grid = grid_factory((n, n), compute_backend=ComputeBackend.CUDA, backend_config={"slab": False})
    box, box_ne = grid.bounding_box_indices(), grid.bounding_box_indices(remove_edges=True)
    walls = np.unique(np.array([box["bottom"][i] + box["left"][i] + box["right"][i] for i in range(2)]), axis=-1).tolist()
    bc_top = EquilibriumBC(rho=1.0, u=(u_lid, 0.0), indices=box_ne["top"])  # constructed first (id 1), listed second: as in the script (:52-56)
    bcs = [HalfwayBounceBackBC(indices=walls), bc_top]
    st = IncompressibleNavierStokesStepper(grid=grid, boundary_conditions=bcs, collision_type="BGK")
    f_0, f_1, bc_mask, missing = st.prepare_fields()
    for i in range(steps):
        f_0, f_1 = st(f_0, f_1, bc_mask, missing, omega, i)
        f_0, f_1 = f_1, f_0
    torch.cuda.synchronize()
    assert torch.equal(got_bc, bc_mask), f"rank {rank}: gathered bc_mask differs from the single-block one"
    assert torch.equal(got, f_0), f"rank {rank}: distributed example != single block (max abs diff {(got - f_0).abs().max().item():.3e})"
    sim.stepper.close()
    if rank == 0:
        print(f"OK lid_driven_cavity_2d_distributed.py world={world} {n}x{n} steps={steps}: bitwise equal to the single-GPU run", flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
