from .equilibrium import Equilibrium as Equilibrium
from .quadratic_equilibrium import QuadraticEquilibrium as QuadraticEquilibrium
