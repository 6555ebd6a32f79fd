"""Device grid: fields are torch CUDA tensors ``[cardinality, nx, ny(, nz)]``, C-contiguous, z fastest --
the layout of jax_grid.py:38-59 / warp_grid.py:17-35 -- optionally an x-slab of a larger global grid."""

import torch

from ..compute_backend import ComputeBackend
from ..default_config import DefaultConfig, current_device
from ..precision_policy import Precision
from .grid import Grid


class CudaGrid(Grid):
    def __init__(self, shape, backend_config=None):
        self.backend_config = dict(backend_config or {})
        super().__init__(shape, ComputeBackend.CUDA)

    def _initialize_backend(self):
        # slab decomposition (xlb_b200.distribute): this process owns global planes [x0, x0 + nx_local)
        self.global_shape = self.shape
        self.nDevices = 1
        self.rank = 0
        self.x_offset = 0
        self.halo = 0  # ghost planes on each side of the x axis

    @property
    def device(self):
        return current_device()

    def create_field(self, cardinality, dtype=None, fill_value=None):
        """grid.create_field (jax_grid.py:38-59): zeros (or ``fill_value``) of ``Precision`` ``dtype``;
        default dtype is the store precision of the default policy."""
        if dtype is None:
            dtype = DefaultConfig.default_precision_policy.store_precision
        tdt = dtype.torch_dtype if isinstance(dtype, Precision) else dtype
        shape = (int(cardinality),) + self.local_shape_with_halo
        if fill_value:
            t = torch.full(shape, fill_value, dtype=tdt, device=self.device)
        else:
            t = torch.zeros(shape, dtype=tdt, device=self.device)
        return t

    @property
    def local_shape_with_halo(self):
        if self.halo:
            return (self.shape[0] + 2 * self.halo,) + tuple(self.shape[1:])
        return tuple(self.shape)
