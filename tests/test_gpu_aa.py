"""In-place (AA pattern) stepping on ONE population buffer (csrc/xlb_step_aa.cuh, ``xlb_step_inplace``; north_star:
"optionally the in-place AA pattern to halve memory") against the two-buffer kernels: bit-identical after every even number
of steps, for periodic boxes and cavities (no BC / FullwayBounceBack / Equilibrium), all lattices, vector and scalar paths."""

import numpy as np
import pytest
import torch

from oracle import cases
from tests.util import TOL_N_STEP, build_product, rel_err, to_dev, to_np
from xlb_b200 import _lib

pytestmark = pytest.mark.gpu


def _fullway_cavity2d(n, policy="FP32FP32", collision="BGK", omega=1.6):
    """D2Q9 cavity with full-way walls (the shipped 2-D example uses half-way walls, which need the second buffer)."""
    from oracle.masker import bounding_box_indices

    shape = (n, n)
    box, box_ne = bounding_box_indices(shape), bounding_box_indices(shape, remove_edges=True)
    walls = np.unique(np.array([box["bottom"][i] + box["left"][i] + box["right"][i] for i in range(2)]), axis=-1).tolist()
    return dict(name=f"cavity2d_fullway_{n}_{collision}_{policy}", lattice="D2Q9", policy=policy, shape=shape, collision=collision, omega=omega, force=None,
                bcs=[dict(kind="FullwayBounceBackBC", indices=walls), dict(kind="EquilibriumBC", rho=1.0, u=(0.05, 0.0), indices=box_ne["top"])])


AA_CASES = [
    cases.periodic((16, 12, 8)),
    cases.periodic((9, 7, 5), lattice="D3Q27", collision="KBC"),  # odd extents: one cell per thread
    cases.periodic((6, 5, 128), lattice="D3Q27", collision="KBC"),  # full warps along z: the shuffles carry the z shift of loads AND stores
    cases.periodic((4, 6, 256), policy="FP64FP64"),
    cases.cavity3d(16, omega=1.7),
    cases.cavity3d(0, shape=(10, 12, 16), lattice="D3Q27", collision="KBC", omega=1.8),
    cases.cavity3d(0, shape=(7, 9, 11), omega=1.3),
    cases.cavity3d(16, policy="FP64FP64", omega=1.7),
    cases.cavity3d(16, policy="FP32FP16", omega=1.7),
    cases.cavity3d(0, shape=(5, 6, 256), omega=1.9, collision="SmagorinskyLESBGK"),
    cases.periodic((20, 24), lattice="D2Q9"),
    _fullway_cavity2d(24),
    _fullway_cavity2d(20, collision="KBC", policy="FP64FP64"),
]


@pytest.mark.parametrize("case", AA_CASES, ids=lambda c: c["name"])
def test_in_place_equals_two_buffers_on_even_steps(case):
    two, _ = build_product(case)
    f_0, f_1, bc_mask, missing = two.prepare_fields()
    one, _ = build_product(case, backend_config={"in_place": True})
    a_0, a_1, bc_a, mm_a = one.prepare_fields()
    assert a_0 is a_1 and torch.equal(bc_a, bc_mask)
    init = to_dev(cases.random_state(case, seed=17))
    f_0.copy_(init)
    a_0.copy_(init)
    n0 = _lib.launch_count()
    for i in range(8):
        f_0, f_1 = two(f_0, f_1, bc_mask, missing, case["omega"], i)
        f_0, f_1 = f_1, f_0
        a_0, a_1 = one(a_0, a_1, bc_a, mm_a, case["omega"], i)
        a_0, a_1 = a_1, a_0  # the reference's loop, harmless on one buffer
        assert one.aa_parity == (i + 1) % 2
        if i % 2 == 1:
            assert torch.equal(a_0, f_0), f"{case['name']}: in-place != two-buffer after {i + 1} steps"
    assert _lib.launch_count() - n0 == 16
    assert torch.isfinite(a_0.float()).all()


def test_in_place_n_step_parity_with_the_oracle():
    """100 steps of the 24^3 cavity on one buffer against the oracle (the N-step bar of the two-buffer path)."""
    case = cases.cavity3d(24, omega=1.7)
    ost, _ = cases.build_oracle(case)
    f0, f1, bc_o, mm_o = ost.prepare_fields()
    ref, _ = ost.run(f0.astype(ost.sdt), f1, bc_o, mm_o, case["omega"], 100)
    st, _ = build_product(case, backend_config={"in_place": True})
    f, _, bc_mask, missing = st.prepare_fields()
    for i in range(100):
        st(f, f, bc_mask, missing, case["omega"], i)
    assert rel_err(to_np(f), ref) <= TOL_N_STEP["FP32FP32"]


def test_in_place_rejects_what_needs_a_second_buffer():
    with pytest.raises(NotImplementedError):
        build_product(cases.sphere3d((48, 24, 24)), backend_config={"in_place": True})  # half-way, Zou-He, extrapolation
    with pytest.raises(NotImplementedError):
        build_product(cases.cavity3d(16, force=(1e-4, 0.0, 0.0)), backend_config={"in_place": True})
    st, _ = build_product(cases.cavity3d(8), backend_config={"in_place": True})
    f, _, bc_mask, missing = st.prepare_fields()
    with pytest.raises(ValueError):
        st.cuda_implementation(f, f.clone(), bc_mask, missing, 1.0, 0)
    # the C layer checks the BC kinds itself
    dom = st._domain(f)
    cp = _lib.XlbCollisionParams()
    cp.bc_kind_mask = 1 << _lib.BC_HALFWAY
    table, n_bc = st.build_bc_table(tuple(f.shape[1:]))
    import ctypes

    rc = _lib.lib().xlb_step_inplace(ctypes.byref(dom), 0, ctypes.byref(cp), f.data_ptr(), bc_mask.data_ptr(), table.data_ptr(), n_bc, 1.0, 0, _lib.current_stream_ptr())
    assert rc == _lib.EUNSUPPORTED
    rc = _lib.lib().xlb_step_inplace(ctypes.byref(dom), 0, ctypes.byref(cp), f.data_ptr(), None, None, 0, 1.0, 2, _lib.current_stream_ptr())
    assert rc == _lib.EINVAL


def test_in_place_at_512_cubed():
    """BASELINE configs[1] on one buffer (10.2 GB instead of 20.4 GB of populations): two steps == the two-buffer kernels."""
    import bench
    from xlb_b200.operator.boundary_condition.boundary_condition_registry import boundary_condition_registry

    boundary_condition_registry.next_id = 1
    two, vs, _, shape, omega = bench.build_case("cavity512_d3q19_bgk_fp32")
    f_0, f_1, bc_mask, missing = two.prepare_fields()
    g = torch.Generator(device="cuda")
    g.manual_seed(5)
    for l in range(vs.q):
        f_0[l] = float(vs.w[l]) * (1.0 + 0.01 * (2.0 * torch.rand(shape, generator=g, device="cuda") - 1.0))
    boundary_condition_registry.next_id = 1
    one, _, _, _, _ = bench.build_case("cavity512_d3q19_bgk_fp32", backend_config={"in_place": True})
    a = f_0.clone()
    for i in range(2):
        f_0, f_1 = two(f_0, f_1, bc_mask, missing, 1.7, i)
        f_0, f_1 = f_1, f_0
        one(a, a, bc_mask, missing, 1.7, i)
    torch.cuda.synchronize()
    assert torch.equal(a, f_0)
