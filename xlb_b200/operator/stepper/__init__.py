from .stepper import Stepper as Stepper
from .nse_stepper import IncompressibleNavierStokesStepper as IncompressibleNavierStokesStepper
from .host_stepper import HostStreamedStepper as HostStreamedStepper
from .graph_stepper import GraphedSteps as GraphedSteps
