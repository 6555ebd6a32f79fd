#!/usr/bin/env python
"""Wind tunnel with a MESH obstacle -- the set-up of the reference's examples/cfd/windtunnel_3d.py (grid and BCs :66-106,
KBC stepper :114-119, drag from momentum exchange :133-170), written against ``xlb_b200``.

The obstacle is a triangle mesh voxelised on the device by the ray-cast masker (``MeshVoxelizationMethod("RAY")``,
windtunnel_3d.py:79): pass ``--stl file.stl`` (binary or ASCII; the reference ships DrivAer-Notchback.stl separately) or
let the script generate an ellipsoidal bluff body.  Prints C_D / C_L every ``--interval`` steps.

    python examples/windtunnel_3d.py --shape 512 128 128 --steps 2000 --interval 500
"""

import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np

import xlb_b200 as xlb
from xlb_b200.compute_backend import ComputeBackend
from xlb_b200.helper import ellipsoid_soup, fit_to_grid, read_stl
from xlb_b200.operator.boundary_condition import ExtrapolationOutflowBC, FullwayBounceBackBC, HalfwayBounceBackBC, RegularizedBC
from xlb_b200.operator.boundary_masker import MeshVoxelizationMethod
from xlb_b200.operator.force import MomentumTransfer
from xlb_b200.operator.macroscopic import Macroscopic
from xlb_b200.operator.stepper import IncompressibleNavierStokesStepper
from xlb_b200.precision_policy import PrecisionPolicy


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", type=int, nargs=3, default=(512, 128, 128))
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--interval", type=int, default=500)
    ap.add_argument("--stl", default=None, help="obstacle mesh (binary / ASCII STL); default: a generated ellipsoid")
    ap.add_argument("--wind_speed", type=float, default=0.02)  # windtunnel_3d.py:41
    ap.add_argument("--reynolds", type=float, default=50000.0)  # windtunnel_3d.py:42
    args = ap.parse_args()

    policy = PrecisionPolicy.FP32FP32
    vs = xlb.velocity_set.D3Q27(precision_policy=policy, compute_backend=ComputeBackend.CUDA)
    xlb.init(velocity_set=vs, default_backend=ComputeBackend.CUDA, default_precision_policy=policy)
    grid_shape = tuple(args.shape)
    grid = xlb.grid.grid_factory(grid_shape)
    box, box_no_edge = grid.bounding_box_indices(), grid.bounding_box_indices(remove_edges=True)
    walls = [box["bottom"][i] + box["top"][i] + box["front"][i] + box["back"][i] for i in range(vs.d)]
    walls = np.unique(np.array(walls), axis=-1).tolist()

    mesh = read_stl(args.stl) if args.stl else ellipsoid_soup(radii=(2.0, 0.9, 0.7), subdivisions=4)
    car_vertices, dx, car_cross_section = fit_to_grid(mesh, grid_shape)  # windtunnel_3d.py:83-100
    if not args.stl:
        car_cross_section *= np.pi / 4.0  # frontal area of an ellipse, not of its bounding rectangle

    # windtunnel_3d.py:44-46: viscosity from the Reynolds number based on the tunnel's x extent
    clength = grid_shape[0] - 1
    visc = args.wind_speed * clength / args.reynolds
    omega = 1.0 / (3.0 * visc + 0.5)

    bc_car = HalfwayBounceBackBC(mesh_vertices=car_vertices, voxelization_method=MeshVoxelizationMethod("RAY"))
    bcs = [
        FullwayBounceBackBC(indices=walls),
        RegularizedBC("velocity", prescribed_value=(args.wind_speed, 0.0, 0.0), indices=box_no_edge["left"]),
        ExtrapolationOutflowBC(indices=box_no_edge["right"]),
        bc_car,
    ]
    stepper = IncompressibleNavierStokesStepper(grid=grid, boundary_conditions=bcs, collision_type="KBC")
    f_0, f_1, bc_mask, missing_mask = stepper.prepare_fields()
    n_car = int((bc_mask[0] == bc_car.id).sum().item())
    print(f"grid {grid_shape}, {mesh.shape[0] // 3} triangles -> {n_car} boundary cells, omega = {omega:.5f}, frontal area = {car_cross_section:.1f} cells")
    momentum_transfer = MomentumTransfer(bc_car)
    macro = Macroscopic()

    for step in range(args.steps):
        f_0, f_1 = stepper(f_0, f_1, bc_mask, missing_mask, omega, step)
        f_0, f_1 = f_1, f_0
        if step % args.interval == 0 or step == args.steps - 1:
            force = momentum_transfer(f_0, f_1, bc_mask, missing_mask)
            cd = 2.0 * force[0].item() / (args.wind_speed**2 * car_cross_section)  # windtunnel_3d.py:153-156
            cl = 2.0 * force[2].item() / (args.wind_speed**2 * car_cross_section)
            rho, u = macro(f_0)
            print(f"step {step}: CD={cd:.4f} CL={cl:.5f}  rho in [{rho.min().item():.5f}, {rho.max().item():.5f}]  max|u|={u.norm(dim=0).max().item():.5f}")


if __name__ == "__main__":
    main()
