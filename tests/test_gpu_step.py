"""The fused step kernel (xlb_step through the C ABI) vs the oracle's restatement of
IncompressibleNavierStokesStepper.jax_implementation_pull (nse_stepper.py:238-282).

Tolerances (BASELINE.json): relative to max|f|; one step <= 1e-6 FP32 / 1e-13 FP64 / 1 fp16 ulp,
N steps <= 1e-5 FP32 / 1e-12 FP64.
"""

import numpy as np
import pytest
import torch

from oracle import cases
from tests.util import TOL_N_STEP, TOL_ONE_STEP, build_product, rel_err, to_dev, to_np
from xlb_b200 import _lib

pytestmark = pytest.mark.gpu

ONE_STEP = [
    cases.periodic((16, 12, 8)),
    cases.periodic((9, 7, 5), lattice="D3Q27", collision="KBC"),  # odd extents -> scalar kernel
    cases.periodic((16, 12, 8), lattice="D3Q27", collision="KBC", policy="FP64FP64"),
    cases.cavity3d(16),
    cases.cavity3d(16, omega=1.7),
    cases.cavity3d(16, policy="FP64FP64", omega=1.7),
    cases.cavity3d(16, policy="FP32FP16", omega=1.7),
    cases.cavity3d(16, policy="FP64FP32", omega=1.7),
    cases.cavity3d(16, policy="FP64FP16", omega=1.7),
    cases.cavity3d(0, shape=(10, 12, 16), lattice="D3Q27", collision="KBC", omega=1.8),
    cases.cavity3d(0, shape=(7, 9, 11), omega=1.3),
    cases.cavity2d(24),
    cases.cavity2d(24, collision="KBC", policy="FP64FP64"),
    cases.cavity2d(0, shape=(21, 13), collision="KBC"),
    cases.sphere3d((48, 24, 24)),
    cases.sphere3d((48, 24, 24), policy="FP64FP64"),
    cases.sphere3d((48, 24, 24), policy="FP32FP16"),
    cases.sphere3d((48, 24, 24), profile=True),
    cases.sphere3d((40, 24, 20), lattice="D3Q19", collision="BGK", inlet="pressure", outlet="pressure"),
    cases.sphere3d((36, 24, 24), outlet="donothing", moving_sphere=(0.01, 0.0, 0.005)),
    cases.sphere3d((36, 20, 20), policy="FP64FP64", inlet="pressure", outlet="extrapolation"),
    cases.sphere3d((48, 24, 24), inlet="regularized_velocity", profile=True),
    cases.sphere3d((40, 24, 20), inlet="regularized_velocity", policy="FP64FP64"),
    cases.sphere3d((40, 24, 20), lattice="D3Q19", collision="BGK", inlet="regularized_pressure", outlet="pressure"),
    # Smagorinsky LES and exact-difference forcing as template cases of the fused kernel
    cases.cavity3d(16, omega=1.9, collision="SmagorinskyLESBGK"),
    cases.cavity3d(16, omega=1.9, collision="SmagorinskyLESBGK", policy="FP64FP64"),
    cases.sphere3d((48, 24, 24), collision="SmagorinskyLESBGK", omega=1.9, inlet="regularized_velocity"),
    cases.cavity2d(24, collision="SmagorinskyLESBGK", omega=1.9),
    cases.periodic((16, 12, 8), omega=1.4, force=(1e-4, 0.0, -5e-5)),
    cases.cavity3d(16, omega=1.7, force=(1e-4, 2e-5, 0.0)),
    cases.cavity3d(16, omega=1.7, policy="FP64FP64", force=(1e-4, 2e-5, 0.0)),
    cases.sphere3d((48, 24, 24), omega=1.7, force=(2e-5, 0.0, 0.0)),
    cases.sphere3d((48, 24, 24), collision="SmagorinskyLESBGK", omega=1.9, force=(2e-5, 0.0, 1e-5)),
    cases.cavity2d(24, collision="KBC", omega=1.7, force=(1e-4, -1e-4)),
]
VARIANTS = [_lib.VARIANT_AUTO, _lib.VARIANT_SCALAR, _lib.VARIANT_GENERIC]


def _one_step(case, variant, seed=0):
    ost, _ = cases.build_oracle(case)
    _, _, bc_o, mm_o = ost.prepare_fields()
    st, _ = build_product(case, backend_config={"kernel_variant": variant})
    _, f_1, bc_mask, missing = st.prepare_fields()
    assert np.array_equal(to_np(bc_mask), bc_o) and np.array_equal(to_np(missing), mm_o)
    f0 = cases.random_state(case, seed=seed)
    _, ref = ost(f0, None, bc_o, mm_o, case["omega"], 0)
    f_0 = to_dev(f0)
    out0, out1 = st(f_0, f_1, bc_mask, missing, case["omega"], 0)
    assert out1.data_ptr() == f_1.data_ptr() and out0.data_ptr() == f_0.data_ptr()
    assert np.array_equal(to_np(f_0), f0), "f_0 must not be modified by the step"
    return to_np(out1), ref


@pytest.mark.parametrize("variant", VARIANTS, ids=["auto", "scalar", "generic"])
@pytest.mark.parametrize("case", ONE_STEP, ids=lambda c: c["name"])
def test_one_step_parity(case, variant):
    got, ref = _one_step(case, variant)
    assert got.dtype == ref.dtype
    tol = TOL_ONE_STEP[case["policy"]]
    if case["collision"] == "KBC":
        tol *= 4  # gamma = f(sp1/sp2) amplifies rounding differences of the summation order
    elif case.get("force") is not None or case["collision"] == "SmagorinskyLESBGK":
        # omega = 1.9 and the feq(u+F) - feq(u) difference amplify FP32 rounding: the FP32 oracle itself is 1.2e-6 away
        # from the same step evaluated in FP64, the CUDA result 1.0e-6 (profiles/r1h_one_step_parity_report.txt)
        tol *= 2
    err = rel_err(got, ref)
    assert err <= tol, f"{case['name']}: rel err {err:.3e} > {tol:.1e}"


@pytest.mark.parametrize(
    "case", [c for c in ONE_STEP if c["policy"] in ("FP32FP32", "FP64FP64")][:8] + [c for c in ONE_STEP if c.get("force") or "LES" in c["collision"]],
    ids=lambda c: c["name"],
)
def test_variants_agree_bitwise(case):
    """The vector, scalar and generic kernels perform the same per-cell arithmetic."""
    a, _ = _one_step(case, _lib.VARIANT_AUTO)
    b, _ = _one_step(case, _lib.VARIANT_SCALAR)
    c, _ = _one_step(case, _lib.VARIANT_GENERIC)
    assert np.array_equal(a, b)
    assert np.array_equal(a, c)


N_STEP = [
    (cases.cavity3d(24, omega=1.0), 100),
    (cases.cavity3d(24, omega=1.7), 100),
    (cases.cavity3d(24, omega=1.7, policy="FP64FP64"), 100),
    (cases.cavity2d(32), 100),
    (cases.sphere3d((48, 24, 24)), 200),
    (cases.sphere3d((48, 24, 24), policy="FP64FP64"), 200),
    (cases.sphere3d((48, 24, 24), profile=True, lattice="D3Q19", collision="BGK", omega=1.5), 100),
    (cases.sphere3d((48, 24, 24), inlet="regularized_velocity", profile=True), 200),
    (cases.sphere3d((48, 24, 24), collision="SmagorinskyLESBGK", omega=1.9, inlet="regularized_velocity"), 200),
    (cases.cavity3d(24, omega=1.7, force=(1e-4, 2e-5, 0.0)), 100),
    (cases.sphere3d((48, 24, 24), omega=1.7, force=(2e-5, 0.0, 0.0), policy="FP64FP64"), 200),
]


@pytest.mark.parametrize("case,nsteps", N_STEP, ids=lambda v: v["name"] if isinstance(v, dict) else str(v))
def test_n_step_parity(case, nsteps):
    ost, _ = cases.build_oracle(case)
    f0, f1, bc_o, mm_o = ost.prepare_fields()
    ref, _ = ost.run(f0.astype(ost.sdt), f1, bc_o, mm_o, case["omega"], nsteps)
    st, _ = build_product(case)
    f_0, f_1, bc_mask, missing = st.prepare_fields()
    for i in range(nsteps):
        f_0, f_1 = st(f_0, f_1, bc_mask, missing, case["omega"], i)
        f_0, f_1 = f_1, f_0
    err = rel_err(to_np(f_0), ref)
    tol = TOL_N_STEP[case["policy"]]
    assert np.isfinite(to_np(f_0)).all()
    assert err <= tol, f"{case['name']}: {nsteps}-step rel err {err:.3e} > {tol:.1e}"


def test_fp16_storage_n_step():
    """FP32FP16: raw populations are stored in fp16 (no f - w shift, like the reference), so one rounding flip of the
    rest population (ulp(1/3) = 2^-12 = 2.4e-4) moves rho and u by that much.  Stated tolerance after 50 steps:
    |rho - rho_ref| <= 5e-3 and |u - u_ref| <= 1e-3 (about 4 fp16 ulps of w_0)."""
    from oracle import ops
    from oracle.lattice import make_lattice

    case = cases.cavity3d(24, policy="FP32FP16", omega=1.7)
    lat = make_lattice(case["lattice"])
    ost, _ = cases.build_oracle(case)
    f0, f1, bc_o, mm_o = ost.prepare_fields()
    ref, _ = ost.run(f0.astype(ost.sdt), f1, bc_o, mm_o, case["omega"], 50)
    st, _ = build_product(case)
    f_0, f_1, bc_mask, missing = st.prepare_fields()
    assert f_0.dtype == torch.float16
    for i in range(50):
        f_0, f_1 = st(f_0, f_1, bc_mask, missing, case["omega"], i)
        f_0, f_1 = f_1, f_0
    rho_r, u_r = ops.macroscopic(lat, ref.astype(np.float32))
    rho_g, u_g = ops.macroscopic(lat, to_np(f_0).astype(np.float32))
    assert np.max(np.abs(rho_g - rho_r)) <= 5e-3
    assert np.max(np.abs(u_g - u_r)) <= 1e-3


def test_slab_launches_equal_whole_launch():
    """xlb_step on x sub-ranges (what the multi-GPU overlap schedule issues) == one whole-grid launch, bitwise."""
    case = cases.sphere3d((48, 24, 24))
    st, _ = build_product(case)
    _, f_1, bc_mask, missing = st.prepare_fields()
    f_0 = to_dev(cases.random_state(case, seed=5))
    st(f_0, f_1, bc_mask, missing, case["omega"], 0)
    whole = f_1.clone()
    f_1.zero_()
    _, dom, table, n_bc, _, _ = st._get_plan(f_0)
    from ctypes import byref

    for xb, xe in [(0, 1), (1, 47), (47, 48)]:
        rc = _lib.lib().xlb_step(byref(dom), st.collision.abi_model, f_0.data_ptr(), f_1.data_ptr(), bc_mask.data_ptr(), missing.data_ptr(),
                                 table.data_ptr(), n_bc, float(case["omega"]), 0, xb, xe, 0, _lib.current_stream_ptr())
        _lib.check(rc, "xlb_step")
    assert torch.equal(whole, f_1)


def test_error_paths():
    case = cases.cavity3d(8)
    st, _ = build_product(case)
    f_0, f_1, bc_mask, missing = st.prepare_fields()
    # Operator.__call__ re-raises as a plain Exception like the reference (operator.py:128-133); the typed error is the cause
    for bad in (
        lambda: st(f_0, f_0, bc_mask, missing, 1.0, 0),  # same buffer
        lambda: st(f_0.double(), f_1, bc_mask, missing, 1.0, 0),  # wrong store dtype
        lambda: st(f_0.cpu(), f_1, bc_mask, missing, 1.0, 0),  # host tensor: no CPU path
        lambda: st(f_0[:, ::2], f_1[:, ::2], bc_mask, missing, 1.0, 0),  # non-contiguous
    ):
        with pytest.raises(Exception) as info:
            bad()
        assert type(info.value) is Exception and isinstance(info.value.__cause__, ValueError)
    with pytest.raises(ValueError):
        st.launch_range(f_0, f_0, bc_mask, missing, 1.0, 0, 0, 8)  # the C-ABI layer itself keeps the typed errors


@pytest.mark.parametrize("case,chunk", [(cases.sphere3d((48, 24, 24)), 7), (cases.cavity3d(0, shape=(21, 12, 16), omega=1.7), 4),
                                         (cases.periodic((16, 8, 8), lattice="D3Q27", collision="KBC"), 16),
                                         (cases.cavity3d(16, policy="FP32FP16", omega=1.5), 5),
                                         # the wrapped stepper's collision parameters travel with it: body force, LES constant
                                         (cases.cavity3d(16, omega=1.7, force=(1e-4, 2e-5, 0.0)), 6),
                                         (cases.sphere3d((48, 24, 24), collision="SmagorinskyLESBGK", omega=1.9, force=(2e-5, 0.0, 1e-5)), 9)], ids=lambda v: v["name"] if isinstance(v, dict) else str(v))
def test_host_streamed_step_equals_resident(case, chunk):
    """HostStreamedStepper (pinned host populations, x-chunk pipeline over three streams) == resident step, bitwise."""
    from xlb_b200.operator.stepper import HostStreamedStepper

    st, _ = build_product(case)
    f_0, f_1, bc_mask, missing = st.prepare_fields()
    f_0 = to_dev(cases.random_state(case, seed=9))
    h0 = torch.empty(f_0.shape, dtype=f_0.dtype, pin_memory=True)
    h1 = torch.empty(f_0.shape, dtype=f_0.dtype, pin_memory=True)
    h0.copy_(f_0)
    hs = HostStreamedStepper(st, chunk_planes=chunk)
    for i in range(3):
        f_0, f_1 = st(f_0, f_1, bc_mask, missing, case["omega"], i)
        f_0, f_1 = f_1, f_0
        h0, h1 = hs(h0, h1, bc_mask, missing, case["omega"], i)
        h0, h1 = h1, h0
    torch.cuda.synchronize()
    assert torch.equal(f_0.cpu(), h0)
    with pytest.raises(ValueError):
        hs(f_0, f_1, bc_mask, missing, 1.0, 0)


BULK_CASES = [
    cases.cavity3d(0, shape=(5, 6, 256), omega=1.7),
    cases.cavity3d(0, shape=(4, 5, 512), omega=1.0),
    cases.cavity3d(0, shape=(4, 4, 256), omega=1.7, policy="FP64FP64"),
    cases.cavity3d(0, shape=(4, 4, 256), omega=1.7, policy="FP32FP16"),
    cases.cavity3d(0, shape=(4, 6, 768), lattice="D3Q27", collision="KBC", omega=1.8),
    cases.sphere3d((28, 16, 256)),
    cases.sphere3d((24, 16, 256), policy="FP64FP64", inlet="pressure"),
    cases.periodic((3, 3, 256), lattice="D3Q27", collision="KBC"),
    cases.cavity2d(0, shape=(20, 256), collision="KBC"),
    cases.sphere3d((28, 16, 128)),  # D3Q27 KBC on a 128-wide grid: tiles of two whole rows (multi-row mode of the 128-thread instance)
    cases.sphere3d((28, 15, 128)),  # odd number of rows: one-row tiles of the 64-thread instance
    cases.periodic((3, 4, 128), lattice="D3Q27", collision="KBC"),  # two-row tiles, every pull wraps in y and z somewhere
    cases.periodic((2, 2, 128), lattice="D3Q27", collision="KBC"),  # ONE tile per plane: both source rows straddle the y boundary
    cases.cavity3d(0, shape=(5, 6, 128), omega=1.7),  # D3Q19 BGK on two-row tiles
    cases.cavity3d(0, shape=(4, 4, 128), omega=1.7, policy="FP64FP64"),
    # FP16 storage: packed-half2 consumer of the staged rows (512-cell tiles where nz allows, else 256)
    cases.cavity3d(0, shape=(4, 5, 512), omega=1.7, policy="FP32FP16"),
    cases.cavity3d(0, shape=(3, 4, 1024), omega=1.2, policy="FP32FP16"),
    cases.periodic((3, 3, 512), policy="FP32FP16"),
    cases.sphere3d((28, 16, 256), policy="FP32FP16"),
    cases.sphere3d((24, 16, 512), policy="FP32FP16", lattice="D3Q19", collision="BGK", inlet="pressure"),
    cases.cavity3d(0, shape=(4, 4, 768), omega=1.9, policy="FP32FP16"),
]


@pytest.mark.parametrize("case", BULK_CASES, ids=lambda c: c["name"])
def test_bulk_variant(case):
    """The persistent bulk-copy-staged kernel (cp.async.bulk + mbarrier ring) == the vector kernel bitwise, and within
    tolerance of the oracle."""
    got, ref = _one_step(case, _lib.VARIANT_BULK)
    tol = TOL_ONE_STEP[case["policy"]] * (4 if case["collision"] == "KBC" else 1)
    assert rel_err(got, ref) <= tol
    vec, _ = _one_step(case, _lib.VARIANT_VECTOR)
    assert np.array_equal(got, vec)


def test_bulk_variant_rejects_unaligned_shapes():
    case = cases.cavity3d(0, shape=(4, 4, 24))
    st, _ = build_product(case, backend_config={"kernel_variant": _lib.VARIANT_BULK})
    f_0, f_1, bc_mask, missing = st.prepare_fields()
    with pytest.raises(ValueError):
        st.launch_range(f_0, f_1, bc_mask, missing, 1.0, 0, 0, 4)
    with pytest.raises(Exception) as info:
        st(f_0, f_1, bc_mask, missing, 1.0, 0)
    assert isinstance(info.value.__cause__, ValueError)


def test_bulk_variant_many_steps_and_slab_ranges():
    """Multi-tile pipeline (more tiles than resident blocks) over 20 steps, and x sub-range launches."""
    case = cases.cavity3d(0, shape=(40, 48, 256), omega=1.6)
    a, _ = build_product(case, backend_config={"kernel_variant": _lib.VARIANT_BULK})
    f_0, f_1, bc_mask, missing = a.prepare_fields()
    b, _ = build_product(case, backend_config={"kernel_variant": _lib.VARIANT_VECTOR})
    g_0, g_1, _, _ = b.prepare_fields()
    for i in range(20):
        f_0, f_1 = a(f_0, f_1, bc_mask, missing, case["omega"], i)
        f_0, f_1 = f_1, f_0
        g_0, g_1 = b(g_0, g_1, bc_mask, missing, case["omega"], i)
        g_0, g_1 = g_1, g_0
    assert torch.equal(f_0, g_0)
    f_1.zero_()
    for xb, xe in [(0, 1), (1, 39), (39, 40)]:
        a.launch_range(f_0, f_1, bc_mask, missing, case["omega"], 0, xb, xe)
    b(g_0, g_1, bc_mask, missing, case["omega"], 0)
    assert torch.equal(f_1, g_1)


def test_bulk_variant_concurrent_launches_on_two_streams():
    """The persistent bulk kernel hands its tiles out through a pair of scheduler words in the library's static scratch; every
    launch gets its own pair (round robin), so launches that overlap on the GPU -- two streams here, the edge / interior
    branches of the multi-GPU schedule in general -- do not disturb each other, and the words are back at zero afterwards."""
    case = cases.cavity3d(0, shape=(64, 48, 256), omega=1.6)
    a, _ = build_product(case, backend_config={"kernel_variant": _lib.VARIANT_BULK})
    f_0, f_1, bc_mask, missing = a.prepare_fields()
    f_0.copy_(to_dev(cases.random_state(case, seed=21)))
    ref, _ = build_product(case, backend_config={"kernel_variant": _lib.VARIANT_VECTOR})
    r_1 = torch.zeros_like(f_1)
    ref(f_0, r_1, bc_mask, missing, case["omega"], 0)
    outs = [torch.zeros_like(f_1) for _ in range(4)]
    streams = [torch.cuda.Stream() for _ in range(2)]
    torch.cuda.synchronize()
    for rep in range(3):
        for k, out in enumerate(outs):
            with torch.cuda.stream(streams[k % 2]):
                a(f_0, out, bc_mask, missing, case["omega"], 0)
        torch.cuda.synchronize()
        for out in outs:
            assert torch.equal(out, r_1)
            out.zero_()


# ----------------------------------------------------------------------------- BC cell list (XlbCollisionParams.bc_cells)
LIST_CASES = [
    cases.sphere3d((48, 24, 24)),
    cases.sphere3d((28, 16, 256)),  # bulk-staged kernel
    cases.sphere3d((36, 24, 24), outlet="donothing", moving_sphere=(0.01, 0.0, 0.005)),
    cases.sphere3d((40, 24, 20), lattice="D3Q19", collision="BGK", inlet="regularized_pressure", outlet="pressure"),
    cases.sphere3d((48, 24, 24), policy="FP32FP16"),
]


def test_list_bc_cells_matches_numpy():
    """xlb_list_bc_cells: the cells of the listed BC kinds, any order, restricted to an x range."""
    from ctypes import byref

    case = cases.sphere3d((23, 9, 7))  # ragged: 23*9*7 is not a multiple of 16
    st, bcs = build_product(case)
    f_0, _, bc_mask, _ = st.prepare_fields()
    _, dom, table, _, _, _ = st._get_plan(f_0)
    listed_ids = [bc.id for bc in bcs if int(bc.abi_kind) in st._LISTED_KINDS]
    assert listed_ids
    m = to_np(bc_mask)[0]
    for xb, xe in [(0, 23), (0, 1), (1, 22), (22, 23), (5, 5)]:
        sel = np.isin(m, listed_ids)
        sel[:xb] = False
        sel[xe:] = False
        want = np.flatnonzero(sel.reshape(-1))
        count = torch.zeros(1, dtype=torch.int64, device="cuda")
        rc = _lib.lib().xlb_list_bc_cells(byref(dom), bc_mask.data_ptr(), table.data_ptr(), 0, xb, xe, None, 0, count.data_ptr(), _lib.current_stream_ptr())
        _lib.check(rc, "xlb_list_bc_cells")
        assert int(count.item()) == want.size
        buf = torch.full((want.size + 3,), -1, dtype=torch.int64, device="cuda")
        count.zero_()
        rc = _lib.lib().xlb_list_bc_cells(byref(dom), bc_mask.data_ptr(), table.data_ptr(), 0, xb, xe, buf.data_ptr(), buf.numel(), count.data_ptr(),
                                          _lib.current_stream_ptr())
        _lib.check(rc, "xlb_list_bc_cells")
        got = to_np(buf)
        assert np.array_equal(np.sort(got[: want.size]), want) and np.all(got[want.size:] == -1)
    # explicit kind mask: the half-way obstacle only
    half_ids = [bc.id for bc in bcs if int(bc.abi_kind) == _lib.BC_HALFWAY]
    count = torch.zeros(1, dtype=torch.int64, device="cuda")
    rc = _lib.lib().xlb_list_bc_cells(byref(dom), bc_mask.data_ptr(), table.data_ptr(), 1 << _lib.BC_HALFWAY, 0, 23, None, 0, count.data_ptr(),
                                      _lib.current_stream_ptr())
    _lib.check(rc, "xlb_list_bc_cells")
    assert int(count.item()) == int(np.isin(m, half_ids).sum())


@pytest.mark.parametrize("case", LIST_CASES, ids=lambda c: c["name"])
def test_bc_cell_list_path_is_bitwise_identical(case):
    """Fused kernel + dense launch over the BC cell list == fused kernel with the in-kernel per-cell path, bitwise, over
    several steps and for x sub-range launches (the multi-GPU schedule)."""
    a, _ = build_product(case, backend_config={"bc_cell_list": True})
    f_0, f_1, bc_mask, missing = a.prepare_fields()
    assert a._cells[1] is not None and a._cells[1].numel() > 0
    b, _ = build_product(case, backend_config={"bc_cell_list": False})
    g_0, g_1, bc_b, mm_b = b.prepare_fields()
    assert b._cells[1] is None
    f_0.copy_(to_dev(cases.random_state(case, seed=3)))
    g_0.copy_(f_0)
    n0 = _lib.launch_count()
    a(f_0, f_1, bc_mask, missing, case["omega"], 0)
    assert _lib.launch_count() - n0 == 2  # fused kernel + list kernel
    f_1.zero_()
    for i in range(6):
        f_0, f_1 = a(f_0, f_1, bc_mask, missing, case["omega"], i)
        f_0, f_1 = f_1, f_0
        g_0, g_1 = b(g_0, g_1, bc_b, mm_b, case["omega"], i)
        g_0, g_1 = g_1, g_0
    assert torch.equal(f_0, g_0)
    nx = case["shape"][0]
    f_1.zero_()
    for xb, xe in [(0, 1), (1, nx - 1), (nx - 1, nx)]:
        a.launch_range(f_0, f_1, bc_mask, missing, case["omega"], 0, xb, xe)
    b(g_0, g_1, bc_b, mm_b, case["omega"], 0)
    assert torch.equal(f_1, g_1)
    c, _ = build_product(case)  # default "auto": this block is far too small for the second launch to pay off
    c.prepare_fields()
    assert c._cells[1] is None


def test_bc_cell_list_follows_mask_updates():
    """The list is cached per bc_mask content: an in-place edit of the mask rebuilds it."""
    case = cases.sphere3d((48, 24, 24))
    st, bcs = build_product(case, backend_config={"bc_cell_list": True})
    f_0, f_1, bc_mask, missing = st.prepare_fields()
    n_before = st._cells[1].numel()
    outlet = [bc for bc in bcs if int(bc.abi_kind) == _lib.BC_EXTRAPOLATION][0]
    bc_mask[bc_mask == outlet.id] = 0  # drop the outflow BC (in place -> version bump)
    st(f_0, f_1, bc_mask, missing, case["omega"], 0)
    assert st._cells[1].numel() < n_before
    ref, _ = build_product(case, backend_config={"kernel_variant": _lib.VARIANT_GENERIC})
    g_1 = torch.zeros_like(f_1)
    ref(f_0, g_1, bc_mask, missing, case["omega"], 0)
    assert torch.equal(f_1, g_1)


@pytest.mark.parametrize("case", [cases.sphere3d((48, 24, 24)), cases.sphere3d((28, 16, 256)), cases.sphere3d((48, 24, 24), policy="FP32FP16"),
                                  cases.sphere3d((40, 24, 20), lattice="D3Q19", collision="BGK", policy="FP64FP64", inlet="pressure", outlet="pressure"),
                                  cases.cavity2d(24, collision="KBC")], ids=lambda c: c["name"])
def test_packed_missing_bits_equal_byte_planes(case):
    """xlb_pack_missing: bit l of word `cell` == missing_mask[l][cell]; and the step with XlbCollisionParams.missing_bits (the
    default) == the step that reads the q byte planes, bitwise, for the fused kernels and the BC cell list kernel."""
    a, _ = build_product(case, backend_config={"bc_cell_list": True})
    f_0, f_1, bc_mask, missing = a.prepare_fields()
    bits = a._mbits[1]
    assert bits is not None and bits.dtype == torch.int32 and tuple(bits.shape) == tuple(missing.shape[1:])
    mm = to_np(missing).astype(np.uint32)
    want = np.zeros(mm.shape[1:], dtype=np.uint32)
    for l in range(mm.shape[0]):
        want |= mm[l] << np.uint32(l)
    assert np.array_equal(to_np(bits).view(np.uint32), want)
    f_0.copy_(to_dev(cases.random_state(case, seed=13)))
    b, _ = build_product(case, backend_config={"bc_cell_list": True, "packed_missing": False})
    g_0, g_1, bc_b, mm_b = b.prepare_fields()
    assert b._mbits[1] is None
    g_0.copy_(f_0)
    for i in range(4):
        f_0, f_1 = a(f_0, f_1, bc_mask, missing, case["omega"], i)
        f_0, f_1 = f_1, f_0
        g_0, g_1 = b(g_0, g_1, bc_b, mm_b, case["omega"], i)
        g_0, g_1 = g_1, g_0
    assert torch.equal(f_0, g_0)


def test_bc_cell_list_follows_remasking_through_the_c_abi():
    """The maskers write bc_mask through raw pointers (invisible to torch's version counter): re-running a masker on the
    same tensor must still invalidate the cached list."""
    from xlb_b200.operator.boundary_masker import IndicesBoundaryMasker

    case = cases.sphere3d((48, 24, 24))
    st, bcs = build_product(case, backend_config={"bc_cell_list": True})
    f_0, f_1, bc_mask, missing = st.prepare_fields()
    n_before = st._cells[1].numel()
    # tag one more inlet-kind cell in place with the inlet's id, through the masker
    inlet = [bc for bc in bcs if int(bc.abi_kind) == _lib.BC_ZOUHE_VELOCITY][0]
    inlet.indices = [[5], [7], [9]]
    IndicesBoundaryMasker(velocity_set=st.velocity_set, precision_policy=st.precision_policy, compute_backend=st.compute_backend, grid=st.grid)(
        [inlet], bc_mask, missing)
    st(f_0, f_1, bc_mask, missing, case["omega"], 0)
    n_after = st._cells[1].numel()
    assert n_after > n_before
    st.invalidate_bc_cells()
    st(f_0, f_1, bc_mask, missing, case["omega"], 0)
    assert st._cells[1].numel() == n_after


def test_second_prepare_fields_fails_loudly():
    """The maskers consume bc.indices (indices_boundary_masker.py:131): a second prepare_fields() with the same BC objects
    raises like the reference instead of returning empty masks."""
    st, _ = build_product(cases.cavity3d(8))
    st.prepare_fields()
    with pytest.raises(AttributeError):
        st.prepare_fields()


# ----------------------------------------------------------------------------- FP16 storage: packed half2 kernel
H2_CASES = [
    cases.cavity3d(16, policy="FP32FP16", omega=1.7),
    cases.cavity3d(0, shape=(6, 10, 20), policy="FP32FP16", omega=1.3),  # ntz = 5: not a power of two, ragged warps
    cases.periodic((8, 6, 128), policy="FP32FP16"),  # full warps along z: shuffles carry the z shift
    cases.sphere3d((48, 24, 24), policy="FP32FP16"),  # D3Q27 KBC, half-way obstacle (vectorised path) + complex BCs
    cases.sphere3d((40, 24, 20), lattice="D3Q19", collision="BGK", policy="FP32FP16", inlet="pressure", outlet="pressure"),
    cases.sphere3d((36, 24, 24), policy="FP32FP16", outlet="donothing", moving_sphere=(0.01, 0.0, 0.005)),
    cases.cavity3d(16, policy="FP32FP16", omega=1.9, collision="SmagorinskyLESBGK"),
    cases.cavity3d(16, policy="FP32FP16", omega=1.7, force=(1e-4, 2e-5, 0.0)),  # forced: widened path
    cases.cavity2d(24, policy="FP32FP16", collision="KBC"),
]


@pytest.mark.parametrize("case", H2_CASES, ids=lambda c: c["name"])
def test_fp16_packed_kernel_bitwise_equals_generic(case):
    """step_vector_h2_kernel (populations packed as half2 from load to store, pair-wise widening) performs the same FP32
    arithmetic and the same round-to-nearest-even store as the per-cell generic kernel."""
    a, ref = _one_step(case, _lib.VARIANT_VECTOR, seed=11)
    b, _ = _one_step(case, _lib.VARIANT_GENERIC, seed=11)
    assert a.dtype == np.float16
    assert np.array_equal(a.view(np.uint16), b.view(np.uint16))
    assert rel_err(a, ref) <= TOL_ONE_STEP["FP32FP16"] * (4 if case["collision"] == "KBC" else 1)


# ----------------------------------------------------------------------------- CUDA-graph replay of the step loop
@pytest.mark.parametrize("case", [cases.cavity3d(16, omega=1.7), cases.sphere3d((48, 24, 24)), cases.cavity3d(16, policy="FP32FP16", omega=1.7)],
                         ids=lambda c: c["name"])
def test_graphed_steps_equal_eager_loop(case):
    """GraphedSteps (an even number of steps captured into one CUDA graph) == the eager loop, bitwise, including the
    two-launch steps of the BC cell list; replays go through no host-side library call."""
    from xlb_b200.operator.stepper import GraphedSteps

    st, _ = build_product(case, backend_config={"bc_cell_list": True})
    f_0, f_1, bc_mask, missing = st.prepare_fields()
    f_0.copy_(to_dev(cases.random_state(case, seed=9)))
    g_0, g_1 = f_0.clone(), f_1.clone()
    for i in range(12):
        g_0, g_1 = st(g_0, g_1, bc_mask, missing, case["omega"], i)
        g_0, g_1 = g_1, g_0
    graphed = GraphedSteps(st, f_0, f_1, bc_mask, missing, case["omega"], steps=4)  # f_1 is scratch, f_0 is left as it was
    n0 = _lib.launch_count()
    for _ in range(3):
        a, b = graphed()
    assert _lib.launch_count() == n0  # replays issue no launches through the library's host path
    assert a.data_ptr() == f_0.data_ptr()
    assert torch.equal(a, g_0)
    with pytest.raises(ValueError):
        GraphedSteps(st, f_0, f_1, bc_mask, missing, case["omega"], steps=3)


PLANE_ORDER_CASES = [
    (cases.cavity3d(0, shape=(9, 6, 128), omega=1.7), "VARIANT_AUTO"),  # K1 (vector)
    (cases.cavity3d(0, shape=(7, 5, 256), omega=1.7, policy="FP32FP16"), "VARIANT_AUTO"),  # K1c (packed half2)
    (cases.cavity3d(0, shape=(6, 4, 256), omega=1.7, policy="FP64FP64"), "VARIANT_AUTO"),
    (cases.sphere3d((28, 16, 256)), "VARIANT_AUTO"),  # K1b (bulk pipeline: keeps its own claim order)
    (cases.sphere3d((28, 16, 128)), "VARIANT_VECTOR"),
    (cases.cavity2d(0, shape=(20, 64), omega=1.6), "VARIANT_AUTO"),
]


@pytest.mark.parametrize("case,variant", PLANE_ORDER_CASES, ids=lambda v: v["name"] if isinstance(v, dict) else v)
def test_timestep_parity_only_changes_the_plane_order(case, variant):
    """``timestep`` is carried for signature parity with nse_stepper.py:229 and the reference never reads it; here its parity
    chooses the direction in which the fused kernels walk the planes (L2 reuse between steps).  The result
    must not depend on it: even and odd time steps from the same state are bitwise equal, full launches and plane sub-ranges."""
    st, _ = build_product(case, backend_config={"kernel_variant": getattr(_lib, variant)})
    f_0, f_1, bc_mask, missing = st.prepare_fields()
    f_0.copy_(to_dev(cases.random_state(case, seed=9)))
    outs = []
    for timestep in (0, 1, 7):
        f_1.zero_()
        st(f_0, f_1, bc_mask, missing, case["omega"], timestep)
        outs.append(f_1.clone())
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])
    if len(case["shape"]) == 3:
        nx = case["shape"][0]
        f_1.zero_()
        for xb, xe in [(0, 2), (2, nx - 1), (nx - 1, nx)]:
            st.launch_range(f_0, f_1, bc_mask, missing, case["omega"], 1, xb, xe)
        assert torch.equal(f_1, outs[0])
