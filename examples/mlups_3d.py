#!/usr/bin/env python
"""Lid-driven cavity throughput benchmark on the CUDA backend -- the workload of the reference's
examples/performance/mlups_3d.py (BCs :194-204, loop :227-229, MLUPS :265-268), written against ``xlb_b200``.

    python examples/mlups_3d.py 512 200 --precision fp32/fp32 --velocity_set D3Q19 --collision BGK
"""

import argparse
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np
import torch

import xlb_b200 as xlb
from xlb_b200.compute_backend import ComputeBackend
from xlb_b200.operator.boundary_condition import EquilibriumBC, FullwayBounceBackBC
from xlb_b200.operator.macroscopic import Macroscopic
from xlb_b200.operator.stepper import IncompressibleNavierStokesStepper
from xlb_b200.precision_policy import PrecisionPolicy

POLICIES = {"fp32/fp32": "FP32FP32", "fp64/fp64": "FP64FP64", "fp64/fp32": "FP64FP32", "fp32/fp16": "FP32FP16", "fp64/fp16": "FP64FP16"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("cube_edge", type=int)
    ap.add_argument("num_steps", type=int)
    ap.add_argument("--precision", default="fp32/fp32", choices=sorted(POLICIES))
    ap.add_argument("--velocity_set", default="D3Q19", choices=["D3Q19", "D3Q27"])
    ap.add_argument("--collision", default="BGK", choices=["BGK", "KBC", "SmagorinskyLESBGK"])
    ap.add_argument("--omega", type=float, default=1.0)
    args = ap.parse_args()

    policy = PrecisionPolicy[POLICIES[args.precision]]
    vs = getattr(xlb.velocity_set, args.velocity_set)(precision_policy=policy, compute_backend=ComputeBackend.CUDA)
    xlb.init(velocity_set=vs, default_backend=ComputeBackend.CUDA, default_precision_policy=policy)
    n = args.cube_edge
    grid = xlb.grid.grid_factory((n, n, n))
    box, box_no_edge = grid.bounding_box_indices(), grid.bounding_box_indices(remove_edges=True)
    walls = np.unique(np.concatenate([np.array(box[k]) for k in ("bottom", "left", "right", "front", "back")], axis=1), axis=-1)
    bcs = [EquilibriumBC(rho=1.0, u=(0.02, 0.0, 0.0), indices=box_no_edge["top"]), FullwayBounceBackBC(indices=walls)]
    stepper = IncompressibleNavierStokesStepper(grid=grid, boundary_conditions=bcs, collision_type=args.collision)
    f_0, f_1, bc_mask, missing_mask = stepper.prepare_fields()

    for i in range(10):  # warm-up
        f_0, f_1 = stepper(f_0, f_1, bc_mask, missing_mask, args.omega, i)
        f_0, f_1 = f_1, f_0
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(args.num_steps):
        f_0, f_1 = stepper(f_0, f_1, bc_mask, missing_mask, args.omega, i)
        f_0, f_1 = f_1, f_0
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0

    rho, u = Macroscopic()(f_0)
    print(f"grid {n}^3  {args.velocity_set} {args.collision} {policy.name}: {args.num_steps} steps in {dt:.3f} s")
    print(f"MLUPS: {n**3 * args.num_steps / dt / 1e6:.1f}")
    print(f"rho in [{rho.min().item():.6f}, {rho.max().item():.6f}], max |u_x| = {u[0].abs().max().item():.6f}")


if __name__ == "__main__":
    main()
