"""Field allocation for the Navier-Stokes stepper (xlb/helper/nse_fields.py:16-55)."""

from ..default_config import DefaultConfig
from ..grid import grid_factory
from ..precision_policy import Precision


def create_nse_fields(grid_shape=None, grid=None, velocity_set=None, compute_backend=None, precision_policy=None):
    """Returns ``(grid, f_0, f_1, missing_mask, bc_mask)``: two population buffers in the store dtype,
    ``missing_mask`` bool [q, ...] (the JAX convention) and ``bc_mask`` uint8 [1, ...], all zero."""
    velocity_set = velocity_set or DefaultConfig.velocity_set
    compute_backend = compute_backend or DefaultConfig.default_backend
    precision_policy = precision_policy or DefaultConfig.default_precision_policy
    if grid is None:
        if grid_shape is None:
            raise ValueError("grid_shape must be provided when grid is None")
        grid = grid_factory(grid_shape, compute_backend=compute_backend, velocity_set=velocity_set)
    f_0 = grid.create_field(cardinality=velocity_set.q, dtype=precision_policy.store_precision)
    f_1 = grid.create_field(cardinality=velocity_set.q, dtype=precision_policy.store_precision)
    bc_mask = grid.create_field(cardinality=1, dtype=Precision.UINT8)
    missing_mask = grid.create_field(cardinality=velocity_set.q, dtype=Precision.BOOL)
    return grid, f_0, f_1, missing_mask, bc_mask
