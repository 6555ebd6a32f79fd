"""Global configuration (xlb/default_config.py:16-100): ``init`` stores the velocity set, backend and
precision policy every operator falls back to, selects the CUDA device of this process and loads the
kernel library."""

import os
from dataclasses import dataclass

from .compute_backend import ComputeBackend, require_cuda_backend


@dataclass
class DefaultConfig:
    default_precision_policy = None
    velocity_set = None
    default_backend = None
    device = None  # torch.device of this process (one process per GPU)


def _select_device():
    import torch

    if not torch.cuda.is_available():
        return None  # host-side objects (BC descriptors, velocity sets) still work; fields and kernels raise
    idx = int(os.environ.get("XLB_CUDA_DEVICE", os.environ.get("LOCAL_RANK", "0")))
    idx = idx % torch.cuda.device_count()
    torch.cuda.set_device(idx)
    return torch.device("cuda", idx)


def init(velocity_set, default_backend, default_precision_policy):
    """Same call as the reference's ``xlb.init`` (default_config.py:60-100)."""
    require_cuda_backend(default_backend)
    DefaultConfig.velocity_set = velocity_set
    DefaultConfig.default_backend = default_backend
    DefaultConfig.default_precision_policy = default_precision_policy
    DefaultConfig.device = _select_device()
    if DefaultConfig.device is not None:
        from . import _lib

        _lib.lib()  # fail loudly now if the CUDA extension is missing


def default_backend() -> ComputeBackend:
    return DefaultConfig.default_backend


def current_device():
    """The CUDA device fields live on; raises when there is none (no CPU fallback)."""
    import torch

    if DefaultConfig.device is None:
        DefaultConfig.device = _select_device()
    if DefaultConfig.device is None:
        raise RuntimeError("xlb_b200 needs a CUDA device: torch.cuda.is_available() is False and there is no CPU fallback")
    return DefaultConfig.device
