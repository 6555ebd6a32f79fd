from ..operator import Operator


class Equilibrium(Operator):
    """Base class of equilibrium operators (xlb/operator/equilibrium/equilibrium.py)."""
