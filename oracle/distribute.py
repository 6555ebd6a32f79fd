"""Oracle for the x-slab ring exchange (TEST INFRASTRUCTURE -- see oracle/__init__.py).

Restates ``distribute_operator._sharded_operator`` (xlb/distribute/distribute.py:18-46):
each shard streams with a *local* periodic roll, then the two wrapped planes are repaired by
two ring shifts (``lax.ppermute``): populations with c_x=+1 of local plane 0 travel to the
right neighbour's plane 0, populations with c_x=-1 of the local last plane travel to the left
neighbour's last plane.  The ring is periodic (:23-24).
"""

import numpy as np

from . import ops


def sharded_stream(lat, f, n_shards):
    """Stream ``f`` [q, nx, ...] as ``n_shards`` x-slabs + ring repair; must equal ops.stream(f)."""
    shards = np.split(f, n_shards, axis=1)
    res = [ops.stream(lat, s) for s in shards]  # local roll wraps inside the shard
    ri, li = lat.right_indices, lat.left_indices
    left_comm = [r[ri, :1, ...].copy() for r in res]  # :26-29
    right_comm = [r[li, -1:, ...].copy() for r in res]
    n = n_shards
    for i in range(n):
        # rightPerm: i -> (i+1) % n  (:23, 31-35); leftPerm: (i+1) % n -> i  (:24, 37-41)
        res[(i + 1) % n][ri, :1, ...] = left_comm[i]
        res[i][li, -1:, ...] = right_comm[(i + 1) % n]
    return np.concatenate(res, axis=1)
