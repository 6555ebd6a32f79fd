"""Field allocation for the Navier-Stokes stepper: the call of xlb/helper/nse_fields.py:16-55, same arguments, same
five-tuple ``(grid, f_0, f_1, missing_mask, bc_mask)`` (note the order: the missing mask comes before the id mask)."""

from ..default_config import DefaultConfig
from ..grid import grid_factory
from ..precision_policy import Precision

# (cardinality selector, element type selector) per returned field, in return order after the grid
_FIELDS = (
    ("f_0", lambda vs: vs.q, lambda pol: pol.store_precision),
    ("f_1", lambda vs: vs.q, lambda pol: pol.store_precision),
    ("missing_mask", lambda vs: vs.q, lambda pol: Precision.BOOL),  # bool [q, ...]: the JAX convention, the parity target
    ("bc_mask", lambda vs: 1, lambda pol: Precision.UINT8),
)


def create_nse_fields(grid_shape=None, grid=None, velocity_set=None, compute_backend=None, precision_policy=None):
    """All-zero device fields of one lattice block.  ``grid`` wins over ``grid_shape``; anything left ``None`` falls back
    to ``xlb.init``'s defaults."""
    cfg = DefaultConfig
    vs = cfg.velocity_set if velocity_set is None else velocity_set
    backend = cfg.default_backend if compute_backend is None else compute_backend
    policy = cfg.default_precision_policy if precision_policy is None else precision_policy
    if grid is None:
        if grid_shape is None:
            raise ValueError("grid_shape must be provided when grid is None")
        grid = grid_factory(grid_shape, compute_backend=backend, velocity_set=vs)
    fields = tuple(grid.create_field(cardinality=card(vs), dtype=elem(policy)) for _, card, elem in _FIELDS)
    return (grid,) + fields
