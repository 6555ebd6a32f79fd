"""Operator base class: the drop-in boundary of the reference (xlb/operator/operator.py:26-133).

Implementations are registered per (class name, backend, signature) and ``__call__`` runs the first
registered implementation whose signature binds the arguments -- the dispatch rule user scripts and the
stepper rely on.  Both calling conventions of the reference coexist because of that rule: the JAX-style
functional form (``macroscopic(f) -> rho, u``) and the Warp-style out-parameter form
(``macroscopic(f, rho, u)``).  Only ``ComputeBackend.CUDA`` implementations exist.
"""

import inspect
import traceback
from ctypes import byref

import torch

from .. import _lib
from ..compute_backend import ComputeBackend, require_cuda_backend
from ..default_config import DefaultConfig
from ..precision_policy import abi_code_of


class Operator:
    _backends = {}

    def __init__(self, velocity_set=None, precision_policy=None, compute_backend=None):
        self.velocity_set = velocity_set or DefaultConfig.velocity_set
        self.precision_policy = precision_policy or DefaultConfig.default_precision_policy
        self.compute_backend = compute_backend or DefaultConfig.default_backend
        if self.compute_backend not in ComputeBackend:
            raise ValueError(f"Compute_backend {compute_backend} is not supported")
        require_cuda_backend(self.compute_backend)
        self._construct_cuda()

    # -- registration / dispatch (operator.py:74-133)
    _signatures = {}  # registered function -> inspect.Signature (resolved once, not per call)
    _bound = {}       # (class name, backend, n positional, keyword names) -> the implementation that bound last time

    @classmethod
    def register_backend(cls, backend_name):
        def decorator(func):
            owner = func.__qualname__.split(".")[0]
            sig = inspect.signature(func)
            cls._backends[(owner, backend_name, str(sig))] = func
            cls._signatures[func] = sig
            cls._bound.clear()
            return func

        return decorator

    def __call__(self, *args, callback=None, **kwargs):
        """Runs the first registered implementation of this class and backend whose signature binds the arguments
        (operator.py:89-133).  As in the reference, whatever an implementation raises is re-raised as a plain
        ``Exception`` that names the operator (operator.py:128-133); the original error is chained as ``__cause__``."""
        me = self.__class__.__name__
        shape_key = (me, self.compute_backend, len(args), tuple(sorted(kwargs)))
        method = self._bound.get(shape_key)
        key = error = trace = None
        if method is None:
            candidates = [(k, m) for k, m in self._backends.items() if k[0] == me and k[1] == self.compute_backend]
            if not candidates:
                have = [k for k in self._backends if k[0] == me]
                raise NotImplementedError(f"No implementation found for operator {me} with backend {self.compute_backend}. Available implementations: {have}")
            for key, cand in candidates:
                try:
                    self._signatures[cand].bind(self, *args, **kwargs)
                except TypeError as exc:  # signature does not bind: try the next registered form
                    error, trace = exc, traceback.format_exc()
                    continue
                method = cand
                self._bound[shape_key] = cand  # binding depends only on the argument count / keyword names
                break
        if method is not None:
            try:
                result = method(self, *args, **kwargs)
            except Exception as exc:
                error, trace = exc, traceback.format_exc()
                key = (me, self.compute_backend, str(self._signatures[method]))
            else:
                if callback and callable(callback):
                    callback(result if result is not None else (args, kwargs))
                return result
        raise Exception(f"Error captured for backend with key {key} for operator {me}: {error}\n {trace}") from error

    @property
    def supported_compute_backend(self):
        return list(self._backends.keys())

    def __repr__(self):
        return f"{self.__class__.__name__}()"

    @property
    def compute_dtype(self):
        return self.precision_policy.compute_precision.torch_dtype

    @property
    def store_dtype(self):
        return self.precision_policy.store_precision.torch_dtype

    def get_precision_policy(self):
        return self.precision_policy

    def get_grid(self):
        return self.grid

    # -- CUDA plumbing shared by all operators
    def _construct_cuda(self):
        """Hook run once at construction (the reference's ``_construct_warp``); resolves nothing by default."""

    def _domain(self, field, x_periodic=1):
        """XlbDomain for a field tensor ``[card, nx, ny(, nz)]`` (2-D fields run as nx=1, ny=NX, nz=NY)."""
        vs = self.velocity_set
        spatial = tuple(field.shape[1:])
        if len(spatial) != vs.d:
            raise ValueError(f"field of shape {tuple(field.shape)} does not match a {vs.d}-D velocity set")
        nx, ny, nz = (1,) + spatial if vs.d == 2 else spatial
        dom = _lib.XlbDomain()
        dom.d, dom.q = vs.d, vs.q
        dom.nx, dom.ny, dom.nz = nx, ny, nz
        dom.x_periodic = x_periodic
        dom.compute_dtype = self.precision_policy.compute_precision.abi_code
        dom.store_dtype = self.precision_policy.store_precision.abi_code
        dom.q_stride = nx * ny * nz
        dom.mask_q_stride = nx * ny * nz
        return dom

    @staticmethod
    def _check_field(t, name, card=None):
        if not isinstance(t, torch.Tensor) or not t.is_cuda:
            raise ValueError(f"{name} must be a CUDA torch.Tensor (xlb_b200 has no CPU path)")
        if not t.is_contiguous():
            raise ValueError(f"{name} must be C-contiguous [cardinality, nx, ny, nz]")
        if card is not None and t.shape[0] != card:
            raise ValueError(f"{name} must have cardinality {card}, got {t.shape[0]}")
        return t

    @staticmethod
    def _ptr(t):
        return None if t is None else t.data_ptr()

    @staticmethod
    def _code(t):
        return abi_code_of(t.dtype)


lib = _lib.lib
check = _lib.check
stream_ptr = _lib.current_stream_ptr
__all__ = ["Operator", "lib", "check", "stream_ptr", "byref"]
