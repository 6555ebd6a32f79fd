"""xlb_b200 -- B200-native implementation of XLB's per-timestep LBM update behind the ``xlb.operator`` API.

Use it as ``import xlb_b200 as xlb`` and select ``ComputeBackend.CUDA``: the stepper / collision / stream /
equilibrium / macroscopic / boundary-condition constructors and call signatures, the velocity sets, the
precision policies and the ``[q, nx, ny, nz]`` field layout are those of the reference (hsalehipour/XLB).
All arithmetic runs in hand-written sm_100a kernels (``csrc/``) reached through a C ABI
(``include/xlb_b200.h``); there is no CPU fallback.
"""

__version__ = "0.1.0"

from .compute_backend import ComputeBackend as ComputeBackend
from .precision_policy import PrecisionPolicy as PrecisionPolicy, Precision as Precision
from .default_config import init as init, DefaultConfig as DefaultConfig

from . import velocity_set as velocity_set
from . import operator as operator
from .operator import equilibrium as _equilibrium, collision as _collision, stream as _stream  # noqa: F401
from .operator import boundary_condition as _bc, boundary_masker as _bm, macroscopic as _macro, stepper as _stepper, force as _force  # noqa: F401
from . import grid as grid
from . import helper as helper
from . import distribute as distribute
