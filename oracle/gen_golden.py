"""Generate tests/golden/*.npz by running the UNMODIFIED reference (hsalehipour/XLB, JAX code path) under
``oracle/refshim`` (NumPy stand-in for jax).  TEST INFRASTRUCTURE; runs only in the build container:

    python -m oracle.gen_golden            # rewrites tests/golden/

Each fixture holds, for one case of ``oracle/cases.py``:
  bc_mask, missing_mask (packbits)           IndicesBoundaryMasker via stepper.prepare_fields()
  f_init                                     initial populations returned by prepare_fields()
  f_one                                      one stepper() call from cases.random_state(case, seed=0)
  f_n, n_steps                               n_steps stepper() calls + swaps from f_init
The random input state is regenerated from its seed by the tests (not stored).
"""

import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def golden_cases():
    from . import cases

    return [
        (cases.periodic((8, 6, 4)), 5),
        (cases.periodic((6, 6, 6), lattice="D3Q27", collision="KBC", policy="FP64FP64"), 5),
        (cases.cavity3d(12, omega=1.0), 20),
        (cases.cavity3d(12, omega=1.7), 20),
        (cases.cavity3d(12, omega=1.7, policy="FP64FP64"), 20),
        (cases.cavity3d(12, omega=1.7, policy="FP32FP16"), 20),
        (cases.cavity3d(0, shape=(8, 10, 12), lattice="D3Q27", collision="KBC", omega=1.8), 20),
        (cases.cavity2d(20), 30),
        (cases.cavity2d(20, collision="KBC", policy="FP64FP64"), 30),
        (cases.sphere3d((28, 16, 16)), 30),
        (cases.sphere3d((28, 16, 16), policy="FP64FP64"), 30),
        (cases.sphere3d((28, 16, 16), profile=True), 10),
        (cases.sphere3d((24, 16, 12), lattice="D3Q19", collision="BGK", inlet="pressure", outlet="pressure"), 20),
        (cases.sphere3d((24, 16, 16), outlet="donothing", moving_sphere=(0.01, 0.0, 0.005)), 10),
        (cases.sphere3d((24, 14, 14), policy="FP64FP64", inlet="pressure", outlet="extrapolation"), 10),
        (cases.sphere3d((28, 16, 16), inlet="regularized_velocity", profile=True), 30),
        (cases.sphere3d((24, 14, 14), policy="FP64FP64", inlet="regularized_pressure", lattice="D3Q19", collision="BGK"), 20),
        # Smagorinsky LES and exact-difference forcing as extra collision cases
        (cases.cavity3d(12, omega=1.9, collision="SmagorinskyLESBGK"), 20),
        (cases.sphere3d((28, 16, 16), collision="SmagorinskyLESBGK", omega=1.9, inlet="regularized_velocity"), 30),
        (cases.cavity2d(20, collision="SmagorinskyLESBGK", policy="FP64FP64", omega=1.9), 30),
        (cases.periodic((8, 6, 8), omega=1.4, force=(1e-4, 0.0, -5e-5)), 10),
        (cases.sphere3d((28, 16, 16), omega=1.7, force=(2e-5, 0.0, 0.0)), 30),
        (cases.cavity3d(12, omega=1.8, collision="SmagorinskyLESBGK", policy="FP64FP64", force=(0.0, 1e-4, 0.0)), 20),
        (cases.cavity2d(20, collision="KBC", omega=1.7, force=(1e-4, -1e-4)), 30),
    ]


def build_reference(case):
    """Reference objects for a case description (the same mapping tests/util.build_product does for the product)."""
    import xlb
    from xlb.compute_backend import ComputeBackend
    from xlb.grid import grid_factory
    from xlb.operator.boundary_condition import (
        DoNothingBC, EquilibriumBC, ExtrapolationOutflowBC, FullwayBounceBackBC, HalfwayBounceBackBC, RegularizedBC, ZouHeBC,
    )
    from xlb.operator.boundary_condition.boundary_condition_registry import boundary_condition_registry
    from xlb.operator.stepper import IncompressibleNavierStokesStepper
    from xlb.precision_policy import PrecisionPolicy
    import jax
    import jax.numpy as jnp

    pol = PrecisionPolicy[case["policy"]]
    jax.config.update("jax_enable_x64", False)  # the reference switches it on itself for FP64FP64 / FP64FP32
    vs = getattr(xlb.velocity_set, case["lattice"])(precision_policy=pol, compute_backend=ComputeBackend.JAX)
    xlb.init(velocity_set=vs, default_backend=ComputeBackend.JAX, default_precision_policy=pol)
    boundary_condition_registry.next_id = 1
    grid = grid_factory(case["shape"], compute_backend=ComputeBackend.JAX)
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
                b = ZouHeBC_(spec["bc_type"], profile=lambda arr=arr: jnp.array(arr), indices=idx)
            else:
                b = ZouHeBC_(spec["bc_type"], prescribed_value=spec["prescribed_value"], indices=idx)
        else:
            raise ValueError(kind)
        bcs.append(b)
    force = None if case.get("force") is None else jnp.array(case["force"])
    st = IncompressibleNavierStokesStepper(grid=grid, boundary_conditions=bcs, collision_type=case["collision"], force_vector=force)
    st._golden_bcs = bcs
    return st


def run_reference(case, n_steps):
    import jax.numpy as jnp

    from . import cases

    st = build_reference(case)
    f_0, f_1, bc_mask, missing = st.prepare_fields()
    out = {"bc_mask": np.asarray(bc_mask), "missing_mask": np.asarray(missing), "f_init": np.asarray(f_0)}
    fin = jnp.array(cases.random_state(case, seed=0))
    _, f_one = st(fin, f_1, bc_mask, missing, case["omega"], 0)
    out["f_one"] = np.asarray(f_one)
    a, b = f_0, f_1
    for i in range(n_steps):
        a, b = st(a, b, bc_mask, missing, case["omega"], i)
        a, b = b, a
    out["f_n"] = np.asarray(a)
    # MomentumTransfer on the HalfwayBounceBackBC obstacle (flow_past_sphere_3d.py:139-172) after the n steps
    from xlb.compute_backend import ComputeBackend
    from xlb.operator.boundary_condition import HalfwayBounceBackBC
    from xlb.operator.force.momentum_transfer import MomentumTransfer

    solids = [bc for bc in st._golden_bcs if isinstance(bc, HalfwayBounceBackBC)]
    if solids:
        mt = MomentumTransfer(solids[-1], compute_backend=ComputeBackend.JAX)
        out["force_n"] = np.asarray(mt(a, b, bc_mask, missing))
    return out


def main():
    from . import refshim

    refshim.install()
    from . import cases

    os.makedirs(GOLDEN, exist_ok=True)
    index = {}
    for case, n_steps in golden_cases():
        ref = run_reference(case, n_steps)
        # immediate cross-check of the restatement (reported, not asserted: tests/test_oracle_golden.py asserts)
        ost, _ = cases.build_oracle(case)
        f0, f1, bc_o, mm_o = ost.prepare_fields()
        _, one = ost(cases.random_state(case, seed=0), None, bc_o, mm_o, case["omega"], 0)
        fn, _ = ost.run(f0.astype(ost.sdt), f1, bc_o, mm_o, case["omega"], n_steps)
        scale = float(np.max(np.abs(ref["f_one"].astype(np.float64))))
        e1 = float(np.max(np.abs(one.astype(np.float64) - ref["f_one"].astype(np.float64)))) / scale
        en = float(np.max(np.abs(fn.astype(np.float64) - ref["f_n"].astype(np.float64)))) / scale
        ok_masks = bool(np.array_equal(bc_o, ref["bc_mask"]) and np.array_equal(mm_o, ref["missing_mask"]))
        extra = {}
        ef = ""
        if "force_n" in ref:
            from . import bc as obc
            from .force import momentum_transfer

            solid = [b for b in ost.boundary_conditions if isinstance(b, obc.HalfwayBounceBackBC)][-1]
            fo = momentum_transfer(ost.lat, solid, fn, bc_o, mm_o)
            extra["force_n"] = ref["force_n"]
            ef = f" force={ref['force_n']} force_equal={bool(np.array_equal(fo, ref['force_n']))}"
        print(f"{case['name']:70s} masks_equal={ok_masks} one_step_err={e1:.2e} n_step_err={en:.2e} dtypes={ref['f_one'].dtype},{ref['f_init'].dtype}{ef}")
        path = os.path.join(GOLDEN, case["name"] + ".npz")
        np.savez_compressed(
            path,
            bc_mask=ref["bc_mask"],
            missing_bits=np.packbits(ref["missing_mask"].astype(np.uint8)),
            missing_shape=np.array(ref["missing_mask"].shape),
            f_init=ref["f_init"],
            f_one=ref["f_one"],
            f_n=ref["f_n"],
            n_steps=np.array(n_steps),
            **extra,
        )
        index[case["name"]] = {"n_steps": n_steps, "bytes": os.path.getsize(path)}
    with open(os.path.join(GOLDEN, "INDEX.json"), "w") as fh:
        json.dump({"generator": "python -m oracle.gen_golden (reference run under oracle/refshim)", "cases": index}, fh, indent=1)
    print("total bytes", sum(v["bytes"] for v in index.values()))


if __name__ == "__main__":
    sys.exit(main())
