"""The CUDA path against the committed golden fixtures (reference run under oracle/refshim): no oracle code involved."""

import numpy as np
import pytest

from oracle import cases
from tests.golden_util import golden_cases, load
from tests.util import TOL_ONE_STEP, build_product, rel_err, to_dev, to_np

pytestmark = pytest.mark.gpu
CASES = golden_cases()


@pytest.mark.parametrize("case,n_steps", CASES, ids=lambda v: v["name"] if isinstance(v, dict) else str(v))
def test_cuda_matches_reference_fixture(case, n_steps):
    g = load(case)
    st, _ = build_product(case)
    f_0, f_1, bc_mask, missing = st.prepare_fields()
    assert np.array_equal(to_np(bc_mask), g["bc_mask"])
    assert np.array_equal(to_np(missing), g["missing_mask"])
    assert np.array_equal(to_np(f_0), g["f_init"].astype(to_np(f_0).dtype))
    tol = TOL_ONE_STEP[case["policy"]] * (4 if case["collision"] == "KBC" else 1)
    fin = to_dev(cases.random_state(case, seed=0))
    st(fin, f_1, bc_mask, missing, case["omega"], 0)
    assert rel_err(to_np(f_1), g["f_one"]) <= tol
    for i in range(n_steps):
        f_0, f_1 = st(f_0, f_1, bc_mask, missing, case["omega"], i)
        f_0, f_1 = f_1, f_0
    ntol = {"FP32FP32": 1e-5, "FP64FP64": 1e-12, "FP32FP16": 2.0**-8}[case["policy"]]
    assert rel_err(to_np(f_0), g["f_n"]) <= ntol
