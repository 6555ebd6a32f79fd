#!/usr/bin/env python
"""Flow past a sphere with drag / lift from momentum exchange -- the set-up of the reference's
examples/cfd/flow_past_sphere_3d.py (geometry :38-54, Poiseuille inlet :58-81, BCs :107-113, force :139-150),
written against ``xlb_b200``.  Prints C_D / C_L every ``--interval`` steps and can dump (rho, u) as ``.npy``.

    python examples/flow_past_sphere_3d.py --shape 256 128 128 --steps 2000 --interval 500
"""

import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np
import torch

import xlb_b200 as xlb
from xlb_b200.compute_backend import ComputeBackend
from xlb_b200.operator.boundary_condition import ExtrapolationOutflowBC, FullwayBounceBackBC, HalfwayBounceBackBC, RegularizedBC
from xlb_b200.operator.force import MomentumTransfer
from xlb_b200.operator.macroscopic import Macroscopic
from xlb_b200.operator.stepper import IncompressibleNavierStokesStepper
from xlb_b200.precision_policy import PrecisionPolicy


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", type=int, nargs=3, default=(256, 128, 128))
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--interval", type=int, default=500)
    ap.add_argument("--velocity_set", default="D3Q19", choices=["D3Q19", "D3Q27"])
    ap.add_argument("--collision", default="BGK", choices=["BGK", "KBC", "SmagorinskyLESBGK"])
    ap.add_argument("--omega", type=float, default=1.6)
    ap.add_argument("--u_max", type=float, default=0.04)
    ap.add_argument("--dump", default=None, help="directory for rho/u .npy snapshots")
    args = ap.parse_args()

    policy = PrecisionPolicy.FP32FP32
    vs = getattr(xlb.velocity_set, args.velocity_set)(precision_policy=policy, compute_backend=ComputeBackend.CUDA)
    xlb.init(velocity_set=vs, default_backend=ComputeBackend.CUDA, default_precision_policy=policy)
    nx, ny, nz = args.shape
    grid = xlb.grid.grid_factory((nx, ny, nz))
    box, box_no_edge = grid.bounding_box_indices(), grid.bounding_box_indices(remove_edges=True)
    walls = np.unique(np.concatenate([np.array(box[k]) for k in ("bottom", "top", "front", "back")], axis=1), axis=-1)

    radius = ny // 12
    cx, cy, cz = nx // 6, ny // 2, nz // 2
    xs, ys, zs = (np.arange(c - radius, c + radius + 1) for c in (cx, cy, cz))
    X, Y, Z = np.meshgrid(xs, ys, zs, indexing="ij")
    inside = (X - cx) ** 2 + (Y - cy) ** 2 + (Z - cz) ** 2 < radius**2
    sphere = np.stack([X[inside], Y[inside], Z[inside]])

    def inlet_profile():
        # parabolic (Poiseuille-like) profile over the inlet plane, zero on the walls
        yy, zz = np.meshgrid(np.arange(ny), np.arange(nz), indexing="ij")
        r2 = (2.0 * (yy - (ny - 1) / 2.0) / (ny - 1)) ** 2 + (2.0 * (zz - (nz - 1) / 2.0) / (nz - 1)) ** 2
        ux = args.u_max * np.maximum(0.0, 1.0 - r2)
        return torch.from_numpy(np.stack([ux, np.zeros_like(ux), np.zeros_like(ux)]).astype(np.float32)).cuda()

    bc_sphere = HalfwayBounceBackBC(indices=sphere)
    bcs = [
        FullwayBounceBackBC(indices=walls),
        RegularizedBC("velocity", profile=inlet_profile, indices=box_no_edge["left"]),
        ExtrapolationOutflowBC(indices=box_no_edge["right"]),
        bc_sphere,
    ]
    stepper = IncompressibleNavierStokesStepper(grid=grid, boundary_conditions=bcs, collision_type=args.collision)
    f_0, f_1, bc_mask, missing_mask = stepper.prepare_fields()
    momentum_transfer = MomentumTransfer(bc_sphere)
    macro = Macroscopic()
    area = np.pi * radius**2

    for step in range(args.steps):
        f_0, f_1 = stepper(f_0, f_1, bc_mask, missing_mask, args.omega, step)
        f_0, f_1 = f_1, f_0
        if step % args.interval == 0 or step == args.steps - 1:
            force = momentum_transfer(f_0, f_1, bc_mask, missing_mask)
            cd = 2.0 * force[0].item() / (args.u_max**2 * area)
            cl = 2.0 * force[2].item() / (args.u_max**2 * area)
            rho, u = macro(f_0)
            print(f"step {step}: CD={cd:.4f} CL={cl:.5f}  rho in [{rho.min().item():.5f}, {rho.max().item():.5f}]  max|u|={u.norm(dim=0).max().item():.5f}")
            if args.dump:
                os.makedirs(args.dump, exist_ok=True)
                np.save(os.path.join(args.dump, f"rho_{step:07d}.npy"), rho.cpu().numpy())
                np.save(os.path.join(args.dump, f"u_{step:07d}.npy"), u.cpu().numpy())


if __name__ == "__main__":
    main()
