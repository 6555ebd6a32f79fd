"""IndicesBoundaryMasker on the GPU: bc_mask and missing_mask bit-exact against the oracle (JAX semantics)."""

import numpy as np
import pytest
import torch

from oracle import cases
from tests.util import build_product, to_np

pytestmark = pytest.mark.gpu

CASES = [
    cases.cavity3d(16),
    cases.cavity3d(0, shape=(12, 10, 8), lattice="D3Q27"),
    cases.cavity2d(24),
    cases.cavity2d(0, shape=(20, 14)),
    cases.sphere3d((48, 24, 24)),
    cases.sphere3d((40, 24, 20), lattice="D3Q19", collision="BGK", inlet="pressure", outlet="pressure"),
    cases.sphere3d((36, 24, 24), outlet="donothing", moving_sphere=(0.01, 0.0, 0.0)),
]


@pytest.mark.parametrize("case", CASES, ids=lambda c: c["name"])
def test_masks_bit_exact(case):
    ost, _ = cases.build_oracle(case)
    _, _, bc_o, mm_o = ost.prepare_fields()
    st, bcs = build_product(case)
    f_0, f_1, bc_mask, missing = st.prepare_fields()
    assert bc_mask.dtype == torch.uint8 and missing.dtype == torch.bool
    assert np.array_equal(to_np(bc_mask), bc_o)
    assert np.array_equal(to_np(missing), mm_o)
    assert all(getattr(b, "indices", None) is None for b in bcs)  # indices_boundary_masker.py:131
    # initial fields: f_0 = w_l, f_1 = 0 (nse_stepper.py:121-146)
    w = st.velocity_set.w.astype(np.float32)
    assert np.array_equal(to_np(f_0), np.broadcast_to(w.reshape((-1,) + (1,) * st.velocity_set.d), f_0.shape))
    assert float(f_1.abs().max()) == 0.0


def test_reference_masker_test_sphere():
    """tests/boundary_conditions/mask/test_bc_indices_masker_jax.py:19-79 on the GPU, incl. the forced id 5."""
    import xlb_b200 as xlb
    from tests.util import init_xlb
    from xlb_b200.operator.boundary_condition import FullwayBounceBackBC
    from xlb_b200.operator.boundary_masker import IndicesBoundaryMasker

    for name, shape in [("D2Q9", (50, 50)), ("D3Q19", (50, 50, 50)), ("D3Q27", (50, 50, 50)), ("D2Q9", (4, 4))]:
        vs, _ = init_xlb(name, "FP32FP32")
        g = xlb.grid.grid_factory(shape)
        missing = g.create_field(cardinality=vs.q, dtype=xlb.Precision.BOOL)
        bc_mask = g.create_field(cardinality=1, dtype=xlb.Precision.UINT8)
        nr = shape[0]
        M = np.meshgrid(*([np.arange(nr)] * vs.d))
        idx = np.where(sum((m - nr // 2) ** 2 for m in M) < (nr // 4) ** 2)
        indices = [tuple(idx[i]) for i in range(vs.d)]
        bc = FullwayBounceBackBC(indices=indices)
        bc.id = 5
        bc_mask, missing = IndicesBoundaryMasker()([bc], bc_mask, missing, None)
        assert missing.dtype == torch.bool and bc_mask.dtype == torch.uint8
        assert tuple(bc_mask.shape) == (1,) + shape and tuple(missing.shape) == (vs.q,) + shape
        m = to_np(bc_mask)
        sel = (0,) + tuple(np.array(i) for i in indices)
        assert np.all(m[sel] == 5)
        m[sel] = 0
        assert np.all(m == 0)
