"""Parity and self-consistency at the sizes BASELINE.json quotes the metric on ("next round" item 1).

* one oracle step at 128^3 D3Q19 BGK FP32 (configs[0]) and at 384x128x128 D3Q27 KBC with the configs[2] BC list;
* VARIANT_AUTO == VARIANT_GENERIC bitwise at 512^3 D3Q19 (FP32 and FP16 storage; population offsets beyond 2^31 elements),
  at 256^3 FP64, and on a 1536 x 1536 cross-section (D3Q27 KBC, BC cell list on: non-power-of-two rows, multi-wave grids,
  the dense list launch);
* mass conservation over 20 steps in a closed 128^3 box.

The big cases are generated on the device (no oracle run is feasible there); what they check is that the tuned kernels
perform the per-cell arithmetic of the generic engine, which the small-grid tests tie to the oracle.
"""

import numpy as np
import pytest
import torch

import bench
from oracle import cases
from tests.util import TOL_ONE_STEP, build_product, rel_err, to_dev, to_np
from xlb_b200 import _lib

pytestmark = pytest.mark.gpu


def _device_random_state(vs, shape, dtype, seed=0, amp=0.01):
    """f = w_l (1 + amp U(-1, 1)) generated on the device (SURVEY 8(d) parity inputs, at sizes NumPy cannot hold comfortably)."""
    g = torch.Generator(device="cuda")
    g.manual_seed(seed)
    f = torch.empty((vs.q,) + tuple(shape), dtype=dtype, device="cuda")
    for l in range(vs.q):
        r = torch.rand(tuple(shape), generator=g, device="cuda", dtype=torch.float32)
        f[l] = (float(vs.w[l]) * (1.0 + amp * (2.0 * r - 1.0))).to(dtype)
        del r
    return f


@pytest.mark.parametrize("omega", [1.0, 1.7])
def test_one_step_parity_cavity_128(omega):
    """BASELINE configs[0]: examples/performance/mlups_3d.py at 128^3, D3Q19 BGK FP32FP32."""
    case = cases.cavity3d(128, omega=omega)
    ost, _ = cases.build_oracle(case)
    _, _, bc_o, mm_o = ost.prepare_fields()
    st, _ = build_product(case)
    _, f_1, bc_mask, missing = st.prepare_fields()
    assert np.array_equal(to_np(bc_mask), bc_o) and np.array_equal(to_np(missing), mm_o)
    f0 = cases.random_state(case, seed=1)
    _, ref = ost(f0, None, bc_o, mm_o, omega, 0)
    st(to_dev(f0), f_1, bc_mask, missing, omega, 0)
    err = rel_err(to_np(f_1), ref)
    assert err <= TOL_ONE_STEP["FP32FP32"], f"128^3 one-step rel err {err:.3e}"


def test_one_step_parity_sphere_384x128x128():
    """BASELINE configs[2]: D3Q27 KBC flow past a sphere, Zou-He velocity inlet + extrapolation outflow + half-way sphere
    + full-way walls, one step against the oracle; KBC bar = 4e-6 (see BASELINE.md tolerances)."""
    case = cases.sphere3d((384, 128, 128))
    ost, _ = cases.build_oracle(case)
    _, _, bc_o, mm_o = ost.prepare_fields()
    st, _ = build_product(case)
    _, f_1, bc_mask, missing = st.prepare_fields()
    assert np.array_equal(to_np(bc_mask), bc_o) and np.array_equal(to_np(missing), mm_o)
    f0 = cases.random_state(case, seed=2)
    _, ref = ost(f0, None, bc_o, mm_o, case["omega"], 0)
    f_0 = to_dev(f0)
    st(f_0, f_1, bc_mask, missing, case["omega"], 0)
    err = rel_err(to_np(f_1), ref)
    assert err <= 4 * TOL_ONE_STEP["FP32FP32"], f"384x128x128 one-step rel err {err:.3e}"
    # the same step through the generic per-cell engine and with the BC cell list forced on: bitwise
    gen, _ = build_product(case, backend_config={"kernel_variant": _lib.VARIANT_GENERIC})
    _, g_1, bc_g, mm_g = gen.prepare_fields()
    gen(f_0, g_1, bc_g, mm_g, case["omega"], 0)
    assert torch.equal(f_1, g_1)
    lst, _ = build_product(case, backend_config={"bc_cell_list": True})
    _, h_1, bc_l, mm_l = lst.prepare_fields()
    assert lst._cells[1] is not None
    lst(f_0, h_1, bc_l, mm_l, case["omega"], 0)
    assert torch.equal(f_1, h_1)


def _auto_equals_generic(workload, shape=None, steps=2, seed=3, backend_config=None):
    from xlb_b200.operator.boundary_condition.boundary_condition_registry import boundary_condition_registry

    boundary_condition_registry.next_id = 1
    auto, vs, pol, shape, omega = bench.build_case(workload, shape=shape, variant=_lib.VARIANT_AUTO, backend_config=backend_config)
    f_0, f_1, bc_mask, missing = auto.prepare_fields()
    del f_0
    f_0 = _device_random_state(vs, shape, f_1.dtype, seed=seed)
    # the generic stepper runs on the masks of the first one (no second 2.7 GB missing_mask at 512^3): its BC objects must
    # get the same ids, or its table would not know the ids stored in bc_mask and every wall would be updated as fluid
    boundary_condition_registry.next_id = 1
    gen, _, _, _, _ = bench.build_case(workload, shape=shape, variant=_lib.VARIANT_GENERIC)
    assert [b.id for b in gen.boundary_conditions] == [b.id for b in auto.boundary_conditions]
    assert int(bc_mask.max()) == max(b.id for b in auto.boundary_conditions)
    g_1 = torch.empty_like(f_1)
    n0 = _lib.launch_count()
    for t in range(steps):
        auto(f_0, f_1, bc_mask, missing, omega, t)
        gen(f_0, g_1, bc_mask, missing, omega, t)
        torch.cuda.synchronize()
        assert torch.equal(f_1, g_1), f"{workload} {shape}: AUTO != GENERIC at step {t}"
        assert bool(torch.isfinite(f_1.float().sum()))
        f_0, f_1 = f_1, f_0
    assert _lib.launch_count() - n0 >= 2 * steps
    return auto


@pytest.mark.parametrize("workload", ["cavity512_d3q19_bgk_fp32", "cavity512_d3q19_bgk_fp16"])
def test_auto_equals_generic_512(workload):
    """BASELINE configs[1] (and its FP16-storage twin): 19 * 512^3 = 2.5e9 elements per buffer, so population offsets
    exceed 32 bits; 262 144 blocks = hundreds of waves."""
    _auto_equals_generic(workload, shape=(512, 512, 512))


def test_auto_equals_generic_256_fp64():
    _auto_equals_generic("cavity512_d3q19_bgk_fp64", shape=(256, 256, 256))


def test_auto_equals_generic_1536_cross_section():
    """BASELINE configs[4] cross-section (1536 x 1536, 8 planes), D3Q27 KBC wind tunnel with the BC cell list: rows of
    768 vectors (not a power of two), 4.7 M listed cells through step_cells_kernel, bulk-staged fused kernel."""
    st = _auto_equals_generic("windtunnel192x1536x1536_d3q27_kbc_fp32", shape=(8, 1536, 1536), backend_config={"bc_cell_list": True})
    assert st._cells[1] is not None and 2 * 1534 * 1534 - 100 < st._cells[1].numel() <= 2 * 1534 * 1534  # inlet + outlet planes (minus the few inlet cells the obstacle re-tags at this plane count)


def test_auto_equals_generic_sphere384_bench_workload():
    """The bench's own configs[2] builder (bench.build_case), so that the timed workload itself is covered."""
    _auto_equals_generic("sphere384_d3q27_kbc_fp32", steps=3)


def test_mass_conservation_closed_box_128():
    """A closed box (six full-way bounce-back walls): streaming permutes, BGK conserves mass and the full-way rule swaps
    populations, so sum(f) over ALL cells is invariant up to FP32 rounding; 20 steps, relative drift <= 1e-6."""
    from xlb_b200.compute_backend import ComputeBackend
    from xlb_b200.grid import grid_factory
    from xlb_b200.operator.boundary_condition import FullwayBounceBackBC
    from xlb_b200.operator.stepper import IncompressibleNavierStokesStepper
    from tests.util import init_xlb

    vs, _ = init_xlb("D3Q19", "FP32FP32")
    grid = grid_factory((128, 128, 128), compute_backend=ComputeBackend.CUDA)
    box = grid.bounding_box_indices()
    walls = np.unique(np.concatenate([np.array(box[k]) for k in ("bottom", "top", "left", "right", "front", "back")], axis=1), axis=-1)
    st = IncompressibleNavierStokesStepper(grid=grid, boundary_conditions=[FullwayBounceBackBC(indices=walls)], collision_type="BGK")
    _, f_1, bc_mask, missing = st.prepare_fields()
    f_0 = _device_random_state(vs, (128, 128, 128), torch.float32, seed=4)
    m0 = float(f_0.double().sum())
    for t in range(20):
        f_0, f_1 = st(f_0, f_1, bc_mask, missing, 1.7, t)
        f_0, f_1 = f_1, f_0
    m1 = float(f_0.double().sum())
    assert abs(m1 - m0) / m0 <= 1e-6, f"mass drift {(m1 - m0) / m0:.3e} over 20 steps"
