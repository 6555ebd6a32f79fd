"""Oracle restatement of MomentumTransfer (TEST INFRASTRUCTURE -- see oracle/__init__.py).

``MomentumTransfer.jax_implementation`` (xlb/operator/force/momentum_transfer.py:174-211) with
``FetchPopulations.jax_implementation`` (:63-70) for the stream-then-collide sequence of the dense stepper.
"""

import numpy as np

from . import ops


def momentum_transfer(lat, no_slip_bc, f_0, bc_mask, missing_mask):
    """Net force on the solid tagged with ``no_slip_bc`` by momentum exchange; returns an array of shape (d,)."""
    f_post_collision = f_0  # momentum_transfer.py:66
    f_post_stream = ops.stream(lat, f_post_collision)  # :67
    f_post_stream = no_slip_bc(f_post_collision, f_post_stream, bc_mask, missing_mask)  # :68
    boundary = bc_mask == no_slip_bc.id  # :199
    boundary = np.broadcast_to(boundary, (lat.q,) + boundary.shape[1:])  # :200-201
    is_edge = np.logical_and(boundary, ~missing_mask[0])  # :204
    opp = lat.opp_indices
    phi = f_post_collision[opp] + f_post_stream  # :206
    phi = np.where(np.logical_and(missing_mask, is_edge), phi, phi.dtype.type(0.0))  # :207
    force = np.tensordot(lat.c[:, opp].astype(phi.dtype), phi, axes=(-1, 0))  # :208
    return np.sum(force, axis=tuple(i + 1 for i in range(lat.d)))  # :209
