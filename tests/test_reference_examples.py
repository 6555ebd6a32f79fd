"""Drop-in proof (north_star "existing examples run by switching only the backend"): the reference's OWN
example sources are executed against ``xlb_b200`` and their results compared with the oracle.

What is changed in a script, and nothing else:
  * the backend it selects: every ``ComputeBackend.JAX`` becomes ``ComputeBackend.CUDA`` (the "switch");
  * for the module-level script (flow_past_sphere_3d.py) the three run-size assignments ``grid_shape`` / ``num_steps`` /
    ``post_process_interval``; class-based examples are instantiated with small arguments instead.
What is supplied from outside the script:
  * ``import xlb`` resolves to ``xlb_b200`` (``xlb_b200.compat.install_as_xlb``: the same module objects, no copy);
  * stand-ins for what is out of scope or absent on this image: ``xlb.utils`` (VTK / image writers: no-ops that record their
    calls), ``warp`` and ``jax.numpy`` (only the handful of names the scripts touch outside the kernels).

The sources are not committed (reference code stays out of the repo): ``oracle/ref_examples.py`` copies them verbatim into
the git-ignored ``oracle/_ref/examples/`` (``__graft_entry__.build()`` does it whenever /root/reference is present).
"""

import os
import re
import sys
import types

import numpy as np
import pytest

from oracle import cases, ref_examples

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _example(name):
    path = ref_examples.path_of(name)
    if path is None:
        pytest.skip(f"{name}: oracle/_ref/examples is empty and /root/reference is not here (run __graft_entry__.build() where it is)")
    with open(path) as fh:
        return fh.read()


def _switch_backend(src):
    out, n = re.subn(r"ComputeBackend\.JAX\b", "ComputeBackend.CUDA", src)
    assert n >= 1, "the script never names the backend it selects?"
    return out


class _Recorder:
    def __init__(self):
        self.calls = []

    def __call__(self, name):
        def fn(*args, **kwargs):
            self.calls.append(name)

        return fn


@pytest.fixture
def reference_environment():
    """``xlb`` -> xlb_b200, plus the out-of-scope / absent modules the scripts import.  Everything is removed afterwards."""
    import torch

    import xlb_b200.compat as compat

    saved = {k: sys.modules.get(k) for k in ("warp", "jax", "jax.numpy", "xlb.utils", "lid_driven_cavity_2d")}
    compat.install_as_xlb()
    rec = _Recorder()
    utils = types.ModuleType("xlb.utils")
    utils.save_image = rec("save_image")
    utils.save_fields_vtk = rec("save_fields_vtk")
    utils.warp_array_to_jax = lambda a: a
    utils.ToJAX = lambda *a, **k: (lambda f: f)
    sys.modules["xlb.utils"] = utils
    sys.modules["xlb"].utils = utils

    wp = types.ModuleType("warp")  # only what the example scripts use outside kernels
    wp.func = lambda f: f
    wp.vec3i = tuple
    wp.vec = lambda value, length=1: np.asarray(value)
    wp.max = np.maximum
    wp.synchronize = torch.cuda.synchronize if torch.cuda.is_available() else (lambda: None)
    sys.modules["warp"] = wp

    jnp = types.ModuleType("jax.numpy")  # post-processing of torch tensors
    jnp.ndarray = torch.Tensor
    jnp.sqrt = torch.sqrt
    for name in ("arange", "meshgrid", "maximum", "zeros_like", "stack"):  # flow_past_sphere_3d.py:62-79 builds its inlet profile with these
        setattr(jnp, name, getattr(np, name))
    jax = types.ModuleType("jax")
    jax.numpy = jnp
    sys.modules["jax"], sys.modules["jax.numpy"] = jax, jnp
    try:
        yield rec
    finally:
        if hasattr(sys.modules.get("xlb"), "utils"):
            del sys.modules["xlb"].utils
        compat.uninstall()
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v


def _exec(src, name, extra=None):
    ns = {"__name__": name, "__file__": name + ".py"}
    ns.update(extra or {})
    exec(compile(src, name + ".py", "exec"), ns)
    return ns


def test_backend_switch_is_the_only_edit():
    """CPU check of the harness itself: the switch touches nothing but the backend selection."""
    for rel, _ in ref_examples.EXAMPLES:
        src = _example(os.path.basename(rel))
        if "ComputeBackend.JAX" not in src:
            continue
        new = _switch_backend(src)
        a, b = src.splitlines(), new.splitlines()
        assert len(a) == len(b)
        for x, y in zip(a, b):
            assert x == y or x.replace("ComputeBackend.JAX", "ComputeBackend.CUDA") == y


def test_alias_resolves_to_the_same_modules():
    """``import xlb`` after install_as_xlb() gives xlb_b200's own module objects; out-of-scope sub-packages stay missing."""
    import xlb_b200
    import xlb_b200.compat as compat

    compat.install_as_xlb()
    try:
        import xlb
        from xlb.operator.force.momentum_transfer import MomentumTransfer  # noqa: F401
        from xlb.operator.stepper import IncompressibleNavierStokesStepper

        assert xlb is xlb_b200
        assert sys.modules["xlb.operator.stepper"] is sys.modules["xlb_b200.operator.stepper"]
        assert IncompressibleNavierStokesStepper.__module__ == "xlb_b200.operator.stepper.nse_stepper"
        with pytest.raises(ModuleNotFoundError):
            import xlb.utils  # noqa: F401
    finally:
        compat.uninstall()
    assert "xlb" not in sys.modules


@pytest.mark.gpu
def test_lid_driven_cavity_2d_runs_unchanged(reference_environment):
    """examples/cfd/lid_driven_cavity_2d.py: the LidDrivenCavity2D class as shipped, 48 x 48, 40 steps, post-processing on;
    the populations equal the oracle's (the unmodified reference's) after the same 40 steps within the FP32 N-step bar."""
    import xlb_b200 as xlb
    from tests.util import rel_err, to_np
    from xlb_b200.compute_backend import ComputeBackend
    from xlb_b200.operator.boundary_condition.boundary_condition_registry import boundary_condition_registry
    from xlb_b200.precision_policy import PrecisionPolicy

    ns = _exec(_switch_backend(_example("lid_driven_cavity_2d.py")), "lid_driven_cavity_2d")
    n, u_lid, steps = 48, 0.05, 40
    omega = 1.0 / (3.0 * (u_lid * (n - 1) / 200.0) + 0.5)  # the script's own Re = 200 recipe (:121-125)
    pol = PrecisionPolicy.FP32FP32
    vs = xlb.velocity_set.D2Q9(precision_policy=pol, compute_backend=ComputeBackend.CUDA)
    boundary_condition_registry.next_id = 1
    sim = ns["LidDrivenCavity2D"](omega, u_lid, (n, n), vs, ComputeBackend.CUDA, pol)
    sim.run(num_steps=steps, post_process_interval=20)
    assert reference_environment.calls.count("save_fields_vtk") == 3 and reference_environment.calls.count("save_image") == 3
    case = cases.cavity2d(n, omega=omega, u_lid=u_lid)
    ost, _ = cases.build_oracle(case)
    f0, f1, bc_o, mm_o = ost.prepare_fields()
    # the script CONSTRUCTS the lid first (id 1) and the walls second (id 2) but LISTS them [walls, lid]; the oracle case numbers
    # them in list order.  Same masks up to that relabelling (the write order, which decides overlaps, is the list order in both).
    relabel = np.array([0, 2, 1], dtype=np.uint8)
    assert np.array_equal(to_np(sim.bc_mask), relabel[bc_o]) and np.array_equal(to_np(sim.missing_mask), mm_o)
    ref, _ = ost.run(f0.astype(ost.sdt), f1, bc_o, mm_o, omega, steps)
    assert rel_err(to_np(sim.f_0), ref) <= 1e-5


@pytest.mark.gpu
def test_flow_past_sphere_3d_runs_unchanged(reference_environment):
    """examples/cfd/flow_past_sphere_3d.py, a module-level script: RegularizedBC inlet fed by the script's own per-index
    (Warp-style) Poiseuille profile, extrapolation outflow, half-way sphere, MomentumTransfer and Macroscopic in its
    post-processing.  Reduced to 64 x 24 x 24 and 30 steps through its three run-size assignments."""
    from tests.util import rel_err, to_np
    from xlb_b200.operator.boundary_condition.boundary_condition_registry import boundary_condition_registry

    src = _switch_backend(_example("flow_past_sphere_3d.py"))
    shape, steps = (64, 24, 24), 30
    for pattern, repl in ((r"^grid_shape = .*$", f"grid_shape = {shape}"), (r"^num_steps = .*$", f"num_steps = {steps}"),
                          (r"^post_process_interval = .*$", "post_process_interval = 10")):
        src, n = re.subn(pattern, repl, src, flags=re.M)
        assert n == 1, pattern
    boundary_condition_registry.next_id = 1
    ns = _exec(src, "flow_past_sphere_3d")
    assert reference_environment.calls.count("save_image") == 4  # steps 0, 10, 20, 29
    case = cases.sphere3d(shape, lattice="D3Q19", collision="BGK", omega=1.6, u_max=0.04, inlet="regularized_velocity", profile=True)
    ost, _ = cases.build_oracle(case)
    f0, f1, bc_o, mm_o = ost.prepare_fields()
    # constructed: inlet (1), walls (2), outlet (3), sphere (4); listed [walls, inlet, outlet, sphere] like the oracle case
    relabel = np.array([0, 2, 1, 3, 4], dtype=np.uint8)
    assert np.array_equal(to_np(ns["bc_mask"]), relabel[bc_o]) and np.array_equal(to_np(ns["missing_mask"]), mm_o)
    ref, _ = ost.run(f0.astype(ost.sdt), f1, bc_o, mm_o, 1.6, steps)
    assert rel_err(to_np(ns["f_0"]), ref) <= 1e-5


@pytest.mark.gpu
def test_mlups_3d_runs_unchanged(reference_environment, capsys):
    """examples/performance/mlups_3d.py: its own ``benchmark(args)`` (init_xlb -> run_simulation -> MLUPS statistics) at 32^3.
    The script picks its backend from a command-line word; the switch is the ``ComputeBackend`` member handed to it."""
    import argparse

    import xlb_b200 as xlb
    from xlb_b200 import _lib
    from xlb_b200.compute_backend import ComputeBackend
    from xlb_b200.precision_policy import PrecisionPolicy

    ns = _exec(_switch_backend(_example("mlups_3d.py")), "mlups_3d")  # its "jax" word now maps to CUDA, its distribute() branch runs
    pol = PrecisionPolicy.FP32FP32
    args = argparse.Namespace(
        cube_edge=32, num_steps=20, compute_backend=ComputeBackend.CUDA, precision="fp32/fp32", precision_policy=pol, gpu_devices=[0],
        velocity_set=xlb.velocity_set.D3Q19(precision_policy=pol, compute_backend=ComputeBackend.CUDA), collision_model="BGK", occ="none",
        report=False, export_final_velocity=False, measure_scalability=False, repetitions=2,
    )
    n0 = _lib.launch_count()
    stats = ns["benchmark"](args)
    assert _lib.launch_count() - n0 >= 10 + 2 * 20  # warm-up + two repetitions of the step loop went through libxlb_b200.so
    assert len(stats["raw_mlups"]) == 2 and stats["mean_mlups"] > 0
    ns["print_summary_with_stats"](args, stats)
    assert "MLUPs" in capsys.readouterr().out


@pytest.mark.gpu
def test_per_index_profile_equals_array_profile():
    """The Warp-style ``profile(index)`` of flow_past_sphere_3d.py:89-102 (one inflow speed per inlet cell, u = -value * normal,
    bc_zouhe.py:332-337; XlbBcEntry.pv_mode = 1) gives bit for bit the step of the JAX-style array profile (:62-79)."""
    import torch

    from tests.util import build_product, to_dev
    from xlb_b200.operator.boundary_condition import RegularizedBC
    from xlb_b200.operator.boundary_condition.boundary_condition_registry import boundary_condition_registry
    from xlb_b200.operator.stepper import IncompressibleNavierStokesStepper

    shape, u_max = (40, 24, 20), 0.04
    case = cases.sphere3d(shape, lattice="D3Q19", collision="BGK", omega=1.6, u_max=u_max, inlet="regularized_velocity", profile=True)
    ref, _ = build_product(case)
    f_0, f_1, bc_mask, missing = ref.prepare_fields()
    f_0.copy_(to_dev(cases.random_state(case, seed=4)))

    H_y, H_z = float(shape[1] - 1), float(shape[2] - 1)

    def bc_profile_warp(index):  # the script's wp.func body (:89-102) in the arithmetic of cases.poiseuille_profile (FP64, then stored)
        y_c, z_c = index[1] - H_y / 2.0, index[2] - H_z / 2.0
        r2 = (2.0 * y_c / H_y) ** 2.0 + (2.0 * z_c / H_z) ** 2.0
        return u_max * np.maximum(0.0, 1.0 - r2)

    boundary_condition_registry.next_id = 1
    alt, bcs = build_product(case)
    inlet_id = [b.id for b in bcs if isinstance(b, RegularizedBC)][0]
    inlet_idx = case["bcs"][1]["indices"]
    new_inlet = RegularizedBC("velocity", profile=bc_profile_warp, indices=inlet_idx)
    new_inlet.id = inlet_id
    new_bcs = [new_inlet if b.id == inlet_id else b for b in bcs]
    st = IncompressibleNavierStokesStepper(grid=alt.grid, boundary_conditions=new_bcs, collision_type="BGK")
    g_1 = torch.zeros_like(f_1)
    for variant_stepper, out in ((ref, f_1), (st, g_1)):
        variant_stepper(f_0, out, bc_mask, missing, 1.6, 0)
    torch.cuda.synchronize()
    assert torch.equal(f_1, g_1)
