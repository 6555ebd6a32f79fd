"""Equilibrium initialisation (xlb/helper/initializers.py:25-72)."""

from ..operator.equilibrium import QuadraticEquilibrium


def initialize_eq(f, grid, velocity_set, precision_policy, compute_backend, rho=None, u=None):
    """Fill ``f`` with feq(rho, u) through the equilibrium kernel; ``rho`` / ``u`` default to 1 and 0."""
    if rho is None:
        rho = grid.create_field(cardinality=1, fill_value=1.0, dtype=precision_policy.compute_precision)
    if u is None:
        u = grid.create_field(cardinality=velocity_set.d, fill_value=0.0, dtype=precision_policy.compute_precision)
    equilibrium = QuadraticEquilibrium(velocity_set=velocity_set, precision_policy=precision_policy, compute_backend=compute_backend)
    f = equilibrium(rho, u, f)
    del rho, u
    return f
