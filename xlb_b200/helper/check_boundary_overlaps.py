"""Duplicate-index diagnostics (xlb/helper/check_boundary_overlaps.py:5-24), JAX behaviour: warn, and the
last BC in list order wins."""

import numpy as np


def check_bc_overlaps(bclist, dim, compute_backend=None):
    columns = []
    for bc in bclist:
        idx = getattr(bc, "indices", None)
        if idx is None:
            continue
        arr = np.asarray(idx).reshape(dim, -1)
        if np.unique(arr, axis=-1).shape[-1] != arr.shape[-1]:
            print(f"WARNING: there are duplicate indices in {bc.__class__.__name__} and hence the order in bc list matters!")
        columns.append(arr)
    if columns:
        every = np.concatenate(columns, axis=1)
        if np.unique(every, axis=-1).shape[-1] != every.shape[-1]:
            print("WARNING: there are duplicate indices in the boundary condition list and hence the order in this list matters!")
