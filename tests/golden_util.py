"""Loader for tests/golden/*.npz (outputs of the unmodified reference run under oracle/refshim; see oracle/gen_golden.py)."""

import json
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def golden_cases():
    from oracle.gen_golden import golden_cases as gc

    return gc()


def load(case):
    z = np.load(os.path.join(GOLDEN, case["name"] + ".npz"))
    shape = tuple(int(v) for v in z["missing_shape"])
    n = int(np.prod(shape))
    missing = np.unpackbits(z["missing_bits"])[:n].reshape(shape).astype(bool)
    out = {
        "bc_mask": z["bc_mask"], "missing_mask": missing, "f_init": z["f_init"], "f_one": z["f_one"], "f_n": z["f_n"],
        "n_steps": int(z["n_steps"]),
    }
    if "force_n" in z.files:
        out["force_n"] = z["force_n"]
    return out


def index():
    with open(os.path.join(GOLDEN, "INDEX.json")) as fh:
        return json.load(fh)
