"""The C-ABI library loads and exports every symbol include/xlb_b200.h declares (no compute calls: CPU only)."""

import ctypes
import os
import re

from xlb_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "xlb_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(xlb_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_exported(lib_built):
    handle = ctypes.CDLL(lib_built)
    names = _declared_symbols()
    assert len(names) >= 10
    for name in names:
        assert hasattr(handle, name), f"{name} declared in include/xlb_b200.h but not exported"


def test_binding_covers_header(lib_built):
    assert sorted(_lib.EXPORTS) == _declared_symbols()


def test_struct_layouts_and_version(lib_built):
    handle = _lib.lib()
    assert handle.xlb_abi_version() == _lib.ABI_VERSION
    assert handle.xlb_bc_entry_size() == ctypes.sizeof(_lib.XlbBcEntry) == _lib.BC_ENTRY_DTYPE.itemsize
    assert handle.xlb_domain_size() == ctypes.sizeof(_lib.XlbDomain)
    assert handle.xlb_launch_count() >= 0


def test_argument_validation_without_gpu(lib_built):
    """Bad arguments are rejected before any CUDA call -> safe on a CPU-only box."""
    handle = _lib.lib()
    dom = _lib.XlbDomain()
    dom.d, dom.q, dom.nx, dom.ny, dom.nz = 3, 20, 4, 4, 4
    rc = handle.xlb_stream(ctypes.byref(dom), 1, None, None, None)
    assert rc == _lib.EUNSUPPORTED
    assert b"velocity set" in handle.xlb_last_error()
    dom.q = 19
    rc = handle.xlb_step(ctypes.byref(dom), 0, None, None, None, None, None, 0, 1.0, 0, 0, 4, 0, None)
    assert rc == _lib.EINVAL


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    import pytest

    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        _lib.lib()
