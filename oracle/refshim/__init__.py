"""refshim -- run the UNMODIFIED reference (``/root/reference/xlb``) on NumPy.  TEST INFRASTRUCTURE ONLY.

The reference's parity target is its JAX backend, but ``jax`` (and ``warp`` / ``neon`` / the visualisation
stack, all imported at module level by ``xlb``) cannot be installed in the build container.  ``install()``
puts two kinds of stand-ins into ``sys.modules``:

* a small functional ``jax`` / ``jax.numpy`` / ``jax.lax`` / ``jax.sharding`` implemented on NumPy, covering
  exactly the primitives the reference's JAX code path calls (array creation, where / sum / tensordot / roll /
  pad, ``x.at[idx].set``, ``jit`` / ``vmap`` as eager Python, ``lax.broadcast_in_dim``, a 1-device mesh) with
  JAX's dtype rules re-stated: python scalars are weak, ``int (x) float -> that float`` (NumPy would give
  float64), and with ``jax_enable_x64`` off every float64 / int64 is canonicalised to 32 bits;
* permissive dummies for ``warp``, ``neon``, ``cupy``, ``trimesh``, ``pyvista``, ``matplotlib`` ... whose
  attributes are only ever touched at import time (decorators, type annotations) on the JAX path.

With that, ``import xlb`` works and ``IncompressibleNavierStokesStepper`` & co. execute the reference's own
Python source line by line; ``oracle/gen_golden.py`` uses it to produce the fixtures in ``tests/golden/``.
What the shim cannot reproduce is XLA's fusion / reduction order -- which is exactly why the reference's
results are only defined to rounding (1e-5 rel FP32, 1e-12 FP64 in BASELINE.json).

Nothing in ``xlb_b200`` imports this; it needs ``/root/reference`` and therefore only runs in the build container.
"""

import functools
import importlib.abc
import importlib.machinery
import sys
import types

import numpy as np

REFERENCE_ROOT = "/root/reference"
_state = {"x64": False}


# ----------------------------------------------------------------------------------------- dtype rules
def _canon(dt):
    """jax.dtypes.canonicalize_dtype: with x64 off, 64-bit types become 32-bit."""
    dt = np.dtype(dt)
    if _state["x64"]:
        return dt
    return {np.dtype(np.float64): np.dtype(np.float32), np.dtype(np.int64): np.dtype(np.int32), np.dtype(np.uint64): np.dtype(np.uint32),
            np.dtype(np.complex128): np.dtype(np.complex64)}.get(dt, dt)


def _default_float():
    return np.dtype(np.float64 if _state["x64"] else np.float32)


def _default_int():
    return np.dtype(np.int64 if _state["x64"] else np.int32)


def _is_weak(x):
    return isinstance(x, (bool, int, float)) and not isinstance(x, np.generic)


def _result_dtype(*xs):
    """JAX type promotion for the cases the reference hits: strong array dtypes promote among themselves
    (bool < int < float by category, widest within the winning category); python scalars are weak."""
    strong = [_canon(np.asarray(x).dtype) for x in xs if not _is_weak(x)]
    weak = [x for x in xs if _is_weak(x)]
    if not strong:
        if any(isinstance(x, float) for x in weak):
            return _default_float()
        if any(isinstance(x, int) and not isinstance(x, bool) for x in weak):
            return _default_int()
        return np.dtype(bool)
    floats = [d for d in strong if d.kind == "f"]
    ints = [d for d in strong if d.kind in "iu"]
    if floats:
        return max(floats, key=lambda d: d.itemsize)
    if any(isinstance(x, float) for x in weak):
        return _default_float()  # int array (x) python float -> default float
    if ints:
        signed = [d for d in ints if d.kind == "i"]
        pick = max(ints, key=lambda d: d.itemsize)
        if signed and pick.kind == "u":
            pick = np.dtype("int%d" % min(64 if _state["x64"] else 32, pick.itemsize * 16))
        return pick
    if any(isinstance(x, int) and not isinstance(x, bool) for x in weak):
        return _default_int()
    return np.dtype(bool)


_ARITH = {np.add, np.subtract, np.multiply, np.true_divide, np.floor_divide, np.power, np.maximum, np.minimum, np.negative,
          np.positive, np.absolute, np.square, np.sqrt, np.remainder, np.exp, np.log}
_COMPARE = {np.equal, np.not_equal, np.less, np.less_equal, np.greater, np.greater_equal}


def _base(x):
    return x.view(np.ndarray) if isinstance(x, np.ndarray) else x


def _wrap(x):
    if isinstance(x, np.ndarray):
        dt = _canon(x.dtype)
        if dt != x.dtype:
            x = x.astype(dt)
        return x.view(JArray)
    if isinstance(x, np.generic):
        return np.asarray(x, dtype=_canon(x.dtype)).view(JArray)
    if isinstance(x, tuple):
        return tuple(_wrap(v) for v in x)
    return x


class _AtIndex:
    def __init__(self, arr, idx):
        self.arr, self.idx = arr, idx

    def _idx(self):
        idx = self.idx
        if isinstance(idx, tuple):
            return tuple(_base(i) for i in idx)
        return _base(idx)

    def set(self, value):
        out = np.array(_base(self.arr), copy=True)
        out[self._idx()] = np.asarray(_base(value)).astype(out.dtype) if isinstance(value, np.ndarray) else value
        return _wrap(out)

    def add(self, value):
        out = np.array(_base(self.arr), copy=True)
        np.add.at(out, self._idx(), value)
        return _wrap(out)

    def get(self):
        return _wrap(_base(self.arr)[self._idx()])


class _At:
    def __init__(self, arr):
        self.arr = arr

    def __getitem__(self, idx):
        return _AtIndex(self.arr, idx)


class JArray(np.ndarray):
    """ndarray with JAX's promotion rules on ufuncs, ``.at[...]`` functional updates and immutability by convention."""

    __array_priority__ = 100

    @property
    def at(self):
        return _At(self)

    def block_until_ready(self):
        return self

    def __array_ufunc__(self, ufunc, method, *inputs, out=None, **kwargs):
        if method != "__call__" or out is not None:
            res = getattr(ufunc, method)(*[_base(i) for i in inputs], **kwargs)
            return _wrap(res)
        ins = inputs
        if ufunc in _ARITH or ufunc in _COMPARE:
            dt = _result_dtype(*inputs)
            if ufunc is np.true_divide and dt.kind != "f":
                dt = _default_float()
            ins = [np.asarray(_base(i), dtype=dt) for i in inputs]
        else:
            ins = [_base(i) for i in inputs]
        return _wrap(ufunc(*ins, **kwargs))

    def __array_function__(self, func, types_, args, kwargs):
        def unwrap(a):
            if isinstance(a, JArray):
                return a.view(np.ndarray)
            if isinstance(a, (list, tuple)):
                return type(a)(unwrap(v) for v in a)
            return a

        return _wrap(func(*unwrap(args), **{k: unwrap(v) for k, v in kwargs.items()}))

    def astype(self, dtype, *a, **k):
        return _wrap(self.view(np.ndarray).astype(_canon(dtype), *a, **k))

    def copy(self, *a, **k):
        return _wrap(self.view(np.ndarray).copy(*a, **k))

    def __getitem__(self, idx):
        if isinstance(idx, tuple):
            idx = tuple(_base(i) for i in idx)
        else:
            idx = _base(idx)
        return _wrap(self.view(np.ndarray)[idx])

    def __iter__(self):
        base = self.view(np.ndarray)
        for i in range(base.shape[0]):
            yield _wrap(base[i])

    def __hash__(self):
        return id(self)


# ----------------------------------------------------------------------------------------- jax.numpy
def _asarray(x, dtype=None):
    if dtype is not None:
        return _wrap(np.asarray(_base(x) if isinstance(x, np.ndarray) else x, dtype=_canon(dtype)))
    if isinstance(x, np.ndarray):
        return _wrap(_base(x))
    arr = np.asarray(x)
    if arr.dtype.kind == "f":
        arr = arr.astype(_default_float()) if not isinstance(x, np.generic) else arr
    elif arr.dtype.kind == "i" and not isinstance(x, np.generic):
        arr = arr.astype(_default_int())
    return _wrap(arr)


def _scalar_in(fn):
    """jnp.<fn>(python scalar) -> weakly typed default-float result; modelled as a 0-d array of the default float
    dtype (identical to the weak type whenever the other operands are not wider than the default float, which
    holds for every policy once ``jax_enable_x64`` follows the reference's own rule)."""

    @functools.wraps(fn)
    def inner(x, *a, **k):
        if _is_weak(x):
            return _wrap(np.asarray(fn(_default_float().type(x), *a, **k)))
        x = _asarray(x)
        if x.dtype.kind != "f":
            x = x.astype(_default_float())
        return _wrap(fn(_base(x), *a, **k))

    return inner


def _make_jnp():
    m = types.ModuleType("jax.numpy")
    m.ndarray = JArray
    m.array = lambda x, dtype=None, copy=True: _asarray(np.array(_base(x) if isinstance(x, np.ndarray) else x), dtype)
    m.asarray = _asarray
    for name in ("float16", "float32", "float64", "int8", "int16", "int32", "int64", "uint8", "uint16", "uint32", "bool_", "inf", "pi", "newaxis"):
        setattr(m, name, getattr(np, name))
    m.zeros = lambda shape, dtype=None: _wrap(np.zeros(shape, dtype=_canon(dtype) if dtype is not None else _default_float()))
    m.ones = lambda shape, dtype=None: _wrap(np.ones(shape, dtype=_canon(dtype) if dtype is not None else _default_float()))
    m.full = lambda shape, fill_value, dtype=None: _wrap(
        np.full(shape, fill_value, dtype=_canon(dtype) if dtype is not None else _result_dtype(fill_value))
    )
    m.zeros_like = lambda x, dtype=None: _wrap(np.zeros_like(_base(_asarray(x)), dtype=dtype))
    m.ones_like = lambda x, dtype=None: _wrap(np.ones_like(_base(_asarray(x)), dtype=dtype))
    m.arange = lambda *a, dtype=None: _asarray(np.arange(*a), dtype)

    def where(cond, x=None, y=None):
        if x is None:
            return tuple(_wrap(i) for i in np.where(_base(cond)))
        dt = _result_dtype(x, y)
        return _wrap(np.where(_base(_asarray(cond)), np.asarray(_base(x), dtype=dt), np.asarray(_base(y), dtype=dt)))

    m.where = where

    def _sum(x, axis=None, keepdims=False, dtype=None):
        x = _asarray(x)
        if x.dtype == bool:
            x = x.astype(_default_int())
        if axis is not None and not isinstance(axis, (int, tuple, list, np.integer)):
            axis = tuple(axis)  # momentum_transfer.py:209 passes a generator expression as ``axis``
        return _wrap(np.sum(_base(x), axis=axis, keepdims=keepdims, dtype=dtype))

    m.sum = _sum

    def tensordot(a, b, axes=2):
        dt = _result_dtype(a, b)
        return _wrap(np.tensordot(np.asarray(_base(a), dtype=dt), np.asarray(_base(b), dtype=dt), axes=axes))

    m.tensordot = tensordot
    m.dot = lambda a, b: tensordot(a, b, axes=1)

    def roll(x, shift, axis=None):
        if isinstance(shift, (tuple, list)):
            shift = tuple(int(s) for s in shift)
        else:
            shift = int(shift)
        return _wrap(np.roll(_base(x), shift, axis=axis))

    m.roll = roll
    m.pad = lambda x, pad_width, mode="constant", constant_values=0: _wrap(np.pad(_base(x), pad_width, mode=mode, constant_values=constant_values))
    m.logical_and = lambda a, b: _wrap(np.logical_and(_base(a), _base(b)))
    m.logical_or = lambda a, b: _wrap(np.logical_or(_base(a), _base(b)))
    m.logical_not = lambda a: _wrap(np.logical_not(_base(a)))
    m.broadcast_to = lambda x, shape: _wrap(np.broadcast_to(_base(_asarray(x)), shape))
    m.square = lambda x: _asarray(x) * _asarray(x)
    m.sqrt = _scalar_in(np.sqrt)
    m.abs = lambda x: _wrap(np.abs(_base(_asarray(x))))
    m.maximum = lambda a, b: np.maximum(_asarray(a), b)
    m.minimum = lambda a, b: np.minimum(_asarray(a), b)
    m.stack = lambda xs, axis=0: _wrap(np.stack([_base(_asarray(x)) for x in xs], axis=axis))
    m.concatenate = lambda xs, axis=0: _wrap(np.concatenate([_base(_asarray(x)) for x in xs], axis=axis))
    m.all = lambda x, axis=None: _wrap(np.all(_base(x), axis=axis))
    m.any = lambda x, axis=None: _wrap(np.any(_base(x), axis=axis))
    m.allclose = lambda a, b, rtol=1e-5, atol=1e-8: bool(np.allclose(_base(a), _base(b), rtol=rtol, atol=atol))
    m.array_equal = lambda a, b: bool(np.array_equal(_base(a), _base(b)))
    m.isnan = lambda x: _wrap(np.isnan(_base(x)))
    m.transpose = lambda x, axes=None: _wrap(np.transpose(_base(x), axes))
    m.reshape = lambda x, shape: _wrap(np.reshape(_base(x), shape))
    m.expand_dims = lambda x, axis: _wrap(np.expand_dims(_base(x), axis))
    m.meshgrid = lambda *xs, indexing="xy": [_wrap(g) for g in np.meshgrid(*[_base(_asarray(x)) for x in xs], indexing=indexing)]
    m.linalg = types.SimpleNamespace(norm=lambda x, axis=None: _wrap(np.linalg.norm(_base(x), axis=axis)))

    def __getattr__(name):
        raise AttributeError(f"refshim jax.numpy has no attribute {name!r}; add it to oracle/refshim if the reference path needs it")

    m.__getattr__ = __getattr__
    return m


# ----------------------------------------------------------------------------------------- jax core
class _Device:
    platform = "cpu"
    device_kind = "refshim-numpy"
    id = 0

    def __repr__(self):
        return "RefshimCpuDevice(id=0)"


_DEVICE = _Device()


def _jit(fun=None, **kwargs):
    if fun is None:
        return lambda f: f
    return fun


def _vmap(fun, in_axes=0, out_axes=0):
    def mapped(*args):
        axes = in_axes if isinstance(in_axes, (tuple, list)) else (in_axes,) * len(args)
        n = next(np.shape(a)[ax] for a, ax in zip(args, axes) if ax is not None)
        outs = []
        for i in range(n):
            sl = [a if ax is None else _wrap(np.take(_base(_asarray(a)), i, axis=ax)) for a, ax in zip(args, axes)]
            outs.append(fun(*sl))
        return _wrap(np.stack([_base(_asarray(o)) for o in outs], axis=out_axes))

    return mapped


def _broadcast_in_dim(operand, shape, broadcast_dimensions):
    operand = _base(_asarray(operand))
    tmp = [1] * len(shape)
    for src, dst in enumerate(broadcast_dimensions):
        tmp[dst] = operand.shape[src]
    return _wrap(np.broadcast_to(operand.reshape(tmp), shape))


class _Config:
    def update(self, key, value):
        if key == "jax_enable_x64":
            _state["x64"] = bool(value)

    @property
    def jax_enable_x64(self):
        return _state["x64"]


class _PartitionSpec(tuple):
    def __new__(cls, *names):
        return super().__new__(cls, names)


class _Mesh:
    def __init__(self, devices, axis_names):
        self.devices = np.asarray(devices, dtype=object)
        self.axis_names = tuple(axis_names)


class _NamedSharding:
    def __init__(self, mesh, spec):
        self.mesh, self.spec = mesh, spec

    def addressable_devices_indices_map(self, shape):
        return {_DEVICE: tuple(slice(None) for _ in shape)}


def _make_jax():
    jax = types.ModuleType("jax")
    jax.__path__ = []
    jnp = _make_jnp()
    lax = types.ModuleType("jax.lax")
    lax.broadcast_in_dim = _broadcast_in_dim

    def _no_collective(*a, **k):
        raise NotImplementedError("refshim: multi-device collectives (lax.ppermute / shard_map) are not emulated; 1 device only")

    lax.ppermute = _no_collective
    sharding = types.ModuleType("jax.sharding")
    sharding.PartitionSpec = _PartitionSpec
    sharding.Mesh = _Mesh
    sharding.NamedSharding = _NamedSharding
    experimental = types.ModuleType("jax.experimental")
    experimental.__path__ = []
    mesh_utils = types.ModuleType("jax.experimental.mesh_utils")
    mesh_utils.create_device_mesh = lambda shape, devices=None: np.array([_DEVICE], dtype=object).reshape(shape)
    experimental.mesh_utils = mesh_utils
    shard_map_mod = types.ModuleType("jax.experimental.shard_map")
    shard_map_mod.shard_map = _no_collective
    experimental.shard_map = shard_map_mod
    image = types.ModuleType("jax.image")
    image.resize = _no_collective
    dlpack = types.ModuleType("jax.dlpack")

    jax.numpy, jax.lax, jax.sharding, jax.experimental, jax.image, jax.dlpack = jnp, lax, sharding, experimental, image, dlpack
    jax.jit, jax.vmap, jax.shard_map = _jit, _vmap, _no_collective
    jax.Array = JArray
    jax.config = _Config()
    jax.devices = lambda backend=None: [_DEVICE] if backend in (None, "cpu") else []
    jax.local_devices = lambda: [_DEVICE]
    jax.device_count = lambda: 1
    jax.local_device_count = lambda: 1
    jax.process_index = lambda: 0
    jax.default_backend = lambda: "cpu"
    jax.device_put = lambda x, device=None: _asarray(x)
    jax.device_get = lambda x: np.asarray(_base(x))
    jax.make_array_from_single_device_arrays = lambda shape, sharding, arrays: arrays[0]
    jax.default_device = _DEVICE
    jax.block_until_ready = lambda x: x
    jax.__version__ = "0.0-refshim"
    mods = {
        "jax": jax, "jax.numpy": jnp, "jax.lax": lax, "jax.sharding": sharding, "jax.experimental": experimental,
        "jax.experimental.mesh_utils": mesh_utils, "jax.experimental.shard_map": shard_map_mod, "jax.image": image, "jax.dlpack": dlpack,
    }
    return mods


# ----------------------------------------------------------------------------------------- permissive dummies
class _Dummy:
    """Absorbs anything done to it at import time: attribute access, calls, decoration, indexing, arithmetic."""

    def __init__(self, name="dummy"):
        object.__setattr__(self, "_name", name)

    def __getattr__(self, item):
        if item.startswith("__") and item.endswith("__"):
            raise AttributeError(item)
        return _Dummy(f"{self._name}.{item}")

    def __call__(self, *args, **kwargs):
        if len(args) == 1 and not kwargs and callable(args[0]) and not isinstance(args[0], _Dummy) and isinstance(args[0], (types.FunctionType, type)):
            return args[0]  # used as a bare decorator
        return _Dummy(f"{self._name}()")

    def __getitem__(self, item):
        return _Dummy(f"{self._name}[]")

    def __iter__(self):
        return iter(())

    def __mro_entries__(self, bases):
        return (object,)

    def __repr__(self):
        return f"<refshim dummy {self._name}>"

    def _binop(self, other):
        return _Dummy(self._name)

    __add__ = __radd__ = __sub__ = __rsub__ = __mul__ = __rmul__ = __truediv__ = __rtruediv__ = __or__ = __ror__ = _binop
    __eq__ = lambda self, other: False  # noqa: E731
    __hash__ = lambda self: id(self)  # noqa: E731
    __bool__ = lambda self: False  # noqa: E731


class _DummyModule(types.ModuleType):
    def __init__(self, name):
        super().__init__(name)
        self.__path__ = []

    def __getattr__(self, item):
        if item.startswith("__") and item.endswith("__"):
            raise AttributeError(item)
        return _Dummy(f"{self.__name__}.{item}")


DUMMY_PACKAGES = {"warp", "neon", "py_neon", "cupy", "cupyx", "trimesh", "pyvista", "matplotlib", "nvtx", "kvikio", "h5py", "pxr", "stl", "phantomgaze",
                  "vtk", "orbax", "PIL", "mpi4py", "numba"}


class _DummyFinder(importlib.abc.MetaPathFinder, importlib.abc.Loader):
    def find_spec(self, fullname, path=None, target=None):
        if fullname.split(".")[0] in DUMMY_PACKAGES:
            return importlib.machinery.ModuleSpec(fullname, self, is_package=True)
        return None

    def create_module(self, spec):
        return _DummyModule(spec.name)

    def exec_module(self, module):
        pass


_installed = False


def install(reference_root=REFERENCE_ROOT):
    """Make ``import xlb`` (the reference) work on NumPy.  Idempotent."""
    global _installed
    if _installed:
        return
    for name in ("jax", "warp"):
        if name in sys.modules and not getattr(sys.modules[name], "__version__", "").endswith("refshim") and name == "jax":
            raise RuntimeError("a real jax is already imported; refshim is for containers without it")
    sys.modules.update(_make_jax())
    sys.meta_path.insert(0, _DummyFinder())
    if reference_root not in sys.path:
        sys.path.insert(0, reference_root)
    _installed = True


def set_x64(flag):
    _state["x64"] = bool(flag)
