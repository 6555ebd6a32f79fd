"""Print the one-step parity error (CUDA vs oracle, relative to max|f|) of every ONE_STEP case of tests/test_gpu_step.py.
Run on a GPU box:  python tools/parity_report.py > gpurun_out/parity_report.txt"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from tests import test_gpu_step as T
from tests.util import TOL_ONE_STEP, rel_err
from xlb_b200 import _lib

for case in T.ONE_STEP:
    got, ref = T._one_step(case, _lib.VARIANT_AUTO)
    err = rel_err(got, ref)
    d = np.abs(got.astype(np.float64) - ref.astype(np.float64))
    idx = np.unravel_index(np.argmax(d), d.shape)
    # rounding context: the FP32 oracle and the CUDA result against the same step evaluated in FP64
    extra = ""
    if case["policy"] == "FP32FP32":
        from oracle import cases as C

        c64 = dict(case, policy="FP64FP64")
        for spec in c64["bcs"]:
            if "profile_array" in spec:
                spec["profile_array"] = spec["profile_array"].astype(np.float64)
        o64, _ = C.build_oracle(c64)
        _, _, bc_o, mm_o = o64.prepare_fields()
        _, truth = o64(C.random_state(case, seed=0).astype(np.float64), None, bc_o, mm_o, case["omega"], 0)
        extra = f" cuda_vs_fp64={rel_err(got, truth):.3e} oracle32_vs_fp64={rel_err(ref, truth):.3e}"
    tol = TOL_ONE_STEP[case["policy"]] * (4 if case["collision"] == "KBC" else 1)
    print(f"{case['name']:100s} err={err:.3e} tol={tol:.1e} worst@{tuple(int(i) for i in idx)} ref={float(ref[idx]):.6g}{extra}", file=sys.stderr)
    print(f"{case['name']:100s} err={err:.3e} tol={tol:.1e} worst@{tuple(int(i) for i in idx)} ref={float(ref[idx]):.6g}{extra}")
