"""Precision / PrecisionPolicy enums (xlb/precision_policy.py:13-110) with torch / numpy / C-ABI dtype accessors."""

from enum import Enum, auto

import numpy as np
import torch

# element-type codes of include/xlb_b200.h
XLB_F16, XLB_F32, XLB_F64, XLB_U8 = 0, 1, 2, 3


class Precision(Enum):
    FP64 = auto()
    FP32 = auto()
    FP16 = auto()
    UINT8 = auto()
    BOOL = auto()

    @property
    def torch_dtype(self):
        return {
            Precision.FP64: torch.float64,
            Precision.FP32: torch.float32,
            Precision.FP16: torch.float16,
            Precision.UINT8: torch.uint8,
            Precision.BOOL: torch.bool,
        }[self]

    @property
    def np_dtype(self):
        return {
            Precision.FP64: np.float64,
            Precision.FP32: np.float32,
            Precision.FP16: np.float16,
            Precision.UINT8: np.uint8,
            Precision.BOOL: np.bool_,
        }[self]

    @property
    def wp_dtype(self):
        """Scalar constructor scripts use inside Warp-style per-index profiles (``wp_dtype(index[1])``,
        examples/cfd/flow_past_sphere_3d.py:84-100).  This package evaluates such profiles on the host with NumPy, so the
        NumPy scalar type is returned (it converts scalars and whole index arrays alike)."""
        return self.np_dtype

    @property
    def jax_dtype(self):
        """NumPy dtype standing in for the reference's ``jnp`` dtype accessor (precision_policy.py:39-54)."""
        return self.np_dtype

    @property
    def abi_code(self):
        return {Precision.FP64: XLB_F64, Precision.FP32: XLB_F32, Precision.FP16: XLB_F16, Precision.UINT8: XLB_U8, Precision.BOOL: XLB_U8}[self]


def abi_code_of(torch_dtype):
    try:
        return {torch.float16: XLB_F16, torch.float32: XLB_F32, torch.float64: XLB_F64, torch.uint8: XLB_U8, torch.bool: XLB_U8}[torch_dtype]
    except KeyError:
        raise ValueError(f"dtype {torch_dtype} is not supported by the CUDA backend") from None


class PrecisionPolicy(Enum):
    """``<compute><store>``: arithmetic in the first precision, fields kept in HBM in the second."""

    FP64FP64 = auto()
    FP64FP32 = auto()
    FP64FP16 = auto()
    FP32FP32 = auto()
    FP32FP16 = auto()

    @property
    def compute_precision(self):
        return Precision.FP64 if self.name.startswith("FP64") else Precision.FP32

    @property
    def store_precision(self):
        return {"FP64": Precision.FP64, "FP32": Precision.FP32, "FP16": Precision.FP16}[self.name[4:]]

    def cast_to_compute(self, tensor):
        return tensor.to(self.compute_precision.torch_dtype)

    def cast_to_store(self, tensor):
        return tensor.to(self.store_precision.torch_dtype)
