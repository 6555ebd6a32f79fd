"""Compute-backend enum (mirrors xlb/compute_backend.py:8-18 and adds the one backend this package implements)."""

from enum import Enum, auto


class ComputeBackend(Enum):
    """``CUDA`` is the only backend that runs here: hand-written sm_100a kernels behind a C ABI.

    ``JAX`` / ``WARP`` / ``NEON`` stay defined so that scripts which merely *name* them still import;
    selecting one raises ``ValueError`` (there is no multi-backend dispatch and no CPU fallback)."""

    JAX = auto()
    WARP = auto()
    NEON = auto()
    CUDA = auto()


def require_cuda_backend(backend):
    if backend != ComputeBackend.CUDA:
        raise ValueError(f"Compute backend {backend} is not available in this build; use ComputeBackend.CUDA")
