"""Test helpers: turn a neutral case description (oracle/cases.py) into the PRODUCT's objects.

Only tests import this; it is the single place where both the oracle and ``xlb_b200`` are visible.
"""

import numpy as np
import torch

import xlb_b200 as xlb
from xlb_b200.compute_backend import ComputeBackend
from xlb_b200.grid import grid_factory
from xlb_b200.operator.boundary_condition import (
    DoNothingBC,
    EquilibriumBC,
    ExtrapolationOutflowBC,
    FullwayBounceBackBC,
    HalfwayBounceBackBC,
    RegularizedBC,
    ZouHeBC,
)
from xlb_b200.operator.boundary_condition.boundary_condition_registry import boundary_condition_registry
from xlb_b200.operator.stepper import IncompressibleNavierStokesStepper
from xlb_b200.precision_policy import PrecisionPolicy

TOL_ONE_STEP = {"FP32FP32": 1e-6, "FP64FP64": 1e-13, "FP64FP32": 1e-7, "FP32FP16": 2.0**-10, "FP64FP16": 2.0**-10}
TOL_N_STEP = {"FP32FP32": 1e-5, "FP64FP64": 1e-12}


def init_xlb(lattice, policy):
    pol = PrecisionPolicy[policy]
    vs_cls = {"D2Q9": xlb.velocity_set.D2Q9, "D3Q19": xlb.velocity_set.D3Q19, "D3Q27": xlb.velocity_set.D3Q27}[lattice]
    vs = vs_cls(precision_policy=pol, compute_backend=ComputeBackend.CUDA)
    xlb.init(velocity_set=vs, default_backend=ComputeBackend.CUDA, default_precision_policy=pol)
    return vs, pol


def build_product(case, backend_config=None, decomposition=None):
    """-> (stepper, [bc objects]) with ids 1.. in list order (registry reset so ids match the oracle's).
    ``decomposition`` (a SlabDecomposition) builds the stepper on this rank's x-slab of the case."""
    init_xlb(case["lattice"], case["policy"])
    boundary_condition_registry.next_id = 1
    if decomposition is not None:
        grid = decomposition.make_grid(case["shape"])
    else:
        grid = grid_factory(case["shape"], compute_backend=ComputeBackend.CUDA, backend_config={"slab": False})  # the whole block, also under torchrun
    bcs = []
    for spec in case["bcs"]:
        kind = spec["kind"]
        idx = [list(c) for c in spec["indices"]]
        if kind == "EquilibriumBC":
            b = EquilibriumBC(rho=spec["rho"], u=spec["u"], indices=idx)
        elif kind == "FullwayBounceBackBC":
            b = FullwayBounceBackBC(indices=idx)
        elif kind == "HalfwayBounceBackBC":
            b = HalfwayBounceBackBC(indices=idx, prescribed_value=spec.get("prescribed_value"))
        elif kind == "DoNothingBC":
            b = DoNothingBC(indices=idx)
        elif kind == "ExtrapolationOutflowBC":
            b = ExtrapolationOutflowBC(indices=idx)
        elif kind in ("ZouHeBC", "RegularizedBC"):
            ZouHeBC_ = ZouHeBC if kind == "ZouHeBC" else RegularizedBC
            if "profile_array" in spec:
                arr = spec["profile_array"]
                b = ZouHeBC_(spec["bc_type"], profile=lambda arr=arr: arr, indices=idx)
            else:
                b = ZouHeBC_(spec["bc_type"], prescribed_value=spec["prescribed_value"], indices=idx)
        else:
            raise ValueError(kind)
        bcs.append(b)
    stepper = IncompressibleNavierStokesStepper(
        grid=grid, boundary_conditions=bcs, collision_type=case["collision"], backend_config=backend_config or {},
        force_vector=None if case.get("force") is None else np.asarray(case["force"], dtype=np.float64),
    )
    return stepper, bcs


def to_dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def to_np(t):
    return t.detach().cpu().numpy()


def rel_err(a, b):
    """max |a-b| / max |b|   (relative to max|f|)."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))
