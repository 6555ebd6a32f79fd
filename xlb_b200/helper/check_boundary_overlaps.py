"""Duplicate-index diagnostics (xlb/helper/check_boundary_overlaps.py:5-24), JAX behaviour: warn, and the
last BC in list order wins."""

import numpy as np


def _keys(arr):
    """One int64 key per index column (row-major over the bounding box of the indices): 1-D ``np.unique`` on keys is
    orders of magnitude faster than ``np.unique(axis=-1)`` on the 1e7-column lists of the 1536^3-class cases."""
    arr = np.asarray(arr, dtype=np.int64)
    lo = arr.min(axis=1, keepdims=True)
    span = (arr.max(axis=1) - lo[:, 0] + 1).astype(np.int64)
    key = np.zeros(arr.shape[1], dtype=np.int64)
    for a in range(arr.shape[0]):
        key = key * span[a] + (arr[a] - lo[a, 0])
    return key


def check_bc_overlaps(bclist, dim, compute_backend=None):
    columns = []
    for bc in bclist:
        idx = getattr(bc, "indices", None)
        if idx is None:
            continue
        arr = np.asarray(idx).reshape(dim, -1)
        if arr.shape[1] and np.unique(_keys(arr)).shape[0] != arr.shape[1]:
            print(f"WARNING: there are duplicate indices in {bc.__class__.__name__} and hence the order in bc list matters!")
        columns.append(arr)
    if columns:
        every = np.concatenate(columns, axis=1)
        if every.shape[1] and np.unique(_keys(every)).shape[0] != every.shape[1]:
            print("WARNING: there are duplicate indices in the boundary condition list and hence the order in this list matters!")
