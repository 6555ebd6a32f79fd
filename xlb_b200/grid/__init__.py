from .grid import grid_factory as grid_factory, Grid as Grid
from .cuda_grid import CudaGrid as CudaGrid
