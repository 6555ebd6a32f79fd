from .boundary_condition_registry import boundary_condition_registry as boundary_condition_registry
from .boundary_condition import BoundaryCondition as BoundaryCondition, ImplementationStep as ImplementationStep
from .bc_simple import (
    FullwayBounceBackBC as FullwayBounceBackBC,
    HalfwayBounceBackBC as HalfwayBounceBackBC,
    EquilibriumBC as EquilibriumBC,
    DoNothingBC as DoNothingBC,
)
from .bc_zouhe import ZouHeBC as ZouHeBC
from .bc_extrapolation_outflow import ExtrapolationOutflowBC as ExtrapolationOutflowBC
from .bc_regularized import RegularizedBC as RegularizedBC
