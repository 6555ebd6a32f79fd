"""ctypes binding of libxlb_b200.so (include/xlb_b200.h).  This is the only way the package computes:
if the library is missing or a call fails, a RuntimeError is raised -- there is no fallback path."""

import ctypes
import os
from ctypes import POINTER, Structure, c_char_p, c_double, c_int, c_int32, c_int64, c_longlong, c_void_p

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("XLB_B200_LIB", os.path.join(_HERE, "_lib", "libxlb_b200.so"))

ABI_VERSION = 4

MASK_STATIC_MAX_CELLS = 16 * 1024 * 1024  # XLB_MASK_STATIC_MAX_CELLS

# enums of include/xlb_b200.h
BGK, KBC, SMAGORINSKY_LES_BGK = 0, 1, 2
BC_NONE, BC_FULLWAY, BC_EQUILIBRIUM, BC_HALFWAY, BC_DONOTHING, BC_ZOUHE_VELOCITY, BC_ZOUHE_PRESSURE, BC_EXTRAPOLATION = range(8)
BC_REGULARIZED_VELOCITY, BC_REGULARIZED_PRESSURE = 8, 9
VARIANT_AUTO, VARIANT_SCALAR, VARIANT_VECTOR, VARIANT_GENERIC, VARIANT_BULK = range(5)
EINVAL, EUNSUPPORTED, EALIGN = -1, -2, -3


class XlbDomain(Structure):
    _fields_ = [
        ("d", c_int32),
        ("q", c_int32),
        ("nx", c_int32),
        ("ny", c_int32),
        ("nz", c_int32),
        ("x_periodic", c_int32),
        ("compute_dtype", c_int32),
        ("store_dtype", c_int32),
        ("q_stride", c_int64),
        ("mask_q_stride", c_int64),
    ]


class XlbCollisionParams(Structure):
    _fields_ = [
        ("smagorinsky_coef", c_double),
        ("has_force", c_int32),
        ("bc_kind_mask", c_int32),
        ("force", c_double * 3),
        ("halfway_fraction", c_double),
        ("bc_cells", c_void_p),
        ("n_bc_cells", c_int64),
        ("missing_bits", c_void_p),
    ]


class XlbStepRange(Structure):
    _fields_ = [("x_begin", c_int32), ("x_stride", c_int32), ("n_planes", c_int32), ("y_begin", c_int32), ("y_stride", c_int32), ("n_rows", c_int32)]


class XlbHaloPush(Structure):
    _fields_ = [("peer_f1", c_void_p * 2), ("src", c_int32 * 2), ("dst", c_int32 * 2), ("peer_q_stride", c_int64 * 2)]


class XlbBcEntry(Structure):
    _fields_ = [
        ("kind", c_int32),
        ("pv_dtype", c_int32),
        ("normal", c_int32 * 3),
        ("pv_mode", c_int32),
        ("pv_stride", c_int64 * 4),
        ("pv", c_void_p),
        ("table", c_double * 27),
    ]


# numpy twin of XlbBcEntry, used to assemble the 256-row table that is uploaded to the device
BC_ENTRY_DTYPE = np.dtype(
    [
        ("kind", np.int32),
        ("pv_dtype", np.int32),
        ("normal", np.int32, (3,)),
        ("pv_mode", np.int32),
        ("pv_stride", np.int64, (4,)),
        ("pv", np.uint64),
        ("table", np.float64, (27,)),
    ],
    align=True,
)

EXPORTS = {
    # name: (restype, argtypes)
    "xlb_last_error": (c_char_p, []),
    "xlb_abi_version": (c_int, []),
    "xlb_bc_entry_size": (c_int, []),
    "xlb_domain_size": (c_int, []),
    "xlb_collision_params_size": (c_int, []),
    "xlb_launch_count": (c_longlong, []),
    "xlb_step": (c_int, [POINTER(XlbDomain), c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_double, c_int, c_int, c_int, c_int, c_void_p]),
    "xlb_step_ex": (c_int, [POINTER(XlbDomain), c_int, POINTER(XlbCollisionParams), c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_double, c_int, c_int, c_int, c_int, c_void_p]),
    "xlb_step_range": (c_int, [POINTER(XlbDomain), c_int, POINTER(XlbCollisionParams), POINTER(XlbStepRange), POINTER(XlbHaloPush), c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_double, c_int, c_int, c_void_p]),
    "xlb_step_inplace": (c_int, [POINTER(XlbDomain), c_int, POINTER(XlbCollisionParams), c_void_p, c_void_p, c_void_p, c_int, c_double, c_int, c_void_p]),
    "xlb_halo_signal": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p]),
    "xlb_halo_wait": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p]),
    "xlb_ipc_export": (c_int, [c_void_p, c_void_p, POINTER(c_int64)]),
    "xlb_ipc_import": (c_int, [c_void_p, c_int64, POINTER(c_void_p), POINTER(c_void_p)]),
    "xlb_ipc_close": (c_int, [c_void_p]),
    "xlb_copy_runs": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_int64, c_int, c_void_p]),
    "xlb_pack_missing": (c_int, [POINTER(XlbDomain), c_void_p, c_void_p, c_void_p]),
    "xlb_list_bc_cells": (c_int, [POINTER(XlbDomain), c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_int64, c_void_p, c_void_p]),
    "xlb_stream": (c_int, [POINTER(XlbDomain), c_int, c_void_p, c_void_p, c_void_p]),
    "xlb_macroscopic": (c_int, [POINTER(XlbDomain), c_int, c_void_p, c_int, c_void_p, c_void_p, c_void_p]),
    "xlb_second_moment": (c_int, [POINTER(XlbDomain), c_int, c_void_p, c_int, c_void_p, c_void_p]),
    "xlb_equilibrium": (c_int, [POINTER(XlbDomain), c_int, c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    "xlb_collide": (c_int, [POINTER(XlbDomain), c_int, c_int, c_void_p, c_void_p, c_void_p, c_double, c_void_p]),
    "xlb_collide_ex": (c_int, [POINTER(XlbDomain), c_int, POINTER(XlbCollisionParams), c_int, c_void_p, c_void_p, c_void_p, c_double, c_void_p]),
    "xlb_apply_bc": (c_int, [POINTER(XlbDomain), c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p]),
    "xlb_momentum_transfer": (c_int, [POINTER(XlbDomain), c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p]),
    "xlb_mask_tag": (c_int, [POINTER(XlbDomain), c_void_p, c_int64, POINTER(c_int32), POINTER(c_int32), c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "xlb_mask_missing": (c_int, [POINTER(XlbDomain), c_void_p, POINTER(c_int32), POINTER(c_int32), c_void_p, c_void_p]),
    "xlb_mask_indices_static": (c_int, [POINTER(XlbDomain), c_void_p, c_int64, POINTER(c_int32), POINTER(c_int32), c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "xlb_mask_mesh_ray": (c_int, [POINTER(XlbDomain), c_void_p, c_int64, POINTER(c_int32), POINTER(c_int32), c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
}

_lib = None


def lib():
    """Load (once) and return the shared library; raise loudly when it is absent or stale."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"CUDA extension not found at {LIB_PATH}: run `python -c 'import __graft_entry__ as g; g.build()'` "
            "(or `python xlb_b200/build.py`).  xlb_b200 has no CPU fallback."
        )
    handle = ctypes.CDLL(LIB_PATH)
    partial = os.environ.get("XLB_B200_TUNING_LIB") == "1"  # tools/sweep.py: single-instance tuning builds may predate an entry point
    for name, (res, args) in EXPORTS.items():
        if partial and not hasattr(handle, name):
            continue
        fn = getattr(handle, name)  # AttributeError = library does not match include/xlb_b200.h
        fn.restype = res
        fn.argtypes = args
    if handle.xlb_abi_version() != ABI_VERSION:
        raise RuntimeError(f"{LIB_PATH}: ABI version {handle.xlb_abi_version()} != expected {ABI_VERSION}; rebuild")
    if handle.xlb_bc_entry_size() != ctypes.sizeof(XlbBcEntry) or handle.xlb_bc_entry_size() != BC_ENTRY_DTYPE.itemsize:
        raise RuntimeError("XlbBcEntry layout mismatch between the library and the Python binding")
    if not partial and handle.xlb_collision_params_size() != ctypes.sizeof(XlbCollisionParams):
        raise RuntimeError("XlbCollisionParams layout mismatch between the library and the Python binding")
    if handle.xlb_domain_size() != ctypes.sizeof(XlbDomain):
        raise RuntimeError("XlbDomain layout mismatch between the library and the Python binding")
    _lib = handle
    return _lib


def check(rc, what):
    """Map a C return code to a Python exception (SURVEY 8(b) error row)."""
    if rc == 0:
        return
    msg = lib().xlb_last_error().decode(errors="replace")
    if rc == EUNSUPPORTED:
        raise NotImplementedError(f"{what}: {msg}")
    if rc in (EINVAL, EALIGN):
        raise ValueError(f"{what}: {msg}")
    raise RuntimeError(f"{what}: {msg} (code {rc})")


def launch_count():
    return int(lib().xlb_launch_count())


def current_stream_ptr():
    import torch

    return c_void_p(torch.cuda.current_stream().cuda_stream)


def int3(values):
    return (c_int32 * 3)(*[int(v) for v in values])
