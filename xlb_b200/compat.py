"""Run scripts written against the reference package name unchanged.

A script of the reference starts with ``import xlb`` / ``from xlb.operator.stepper import ...``.  ``install_as_xlb()`` makes
those names resolve to this package -- ``xlb`` *is* ``xlb_b200`` afterwards, ``xlb.operator.stepper`` *is*
``xlb_b200.operator.stepper`` (the same module objects, so the operator registry and the BC id registry exist once) -- so
that the only edit a script needs is the backend it selects (``ComputeBackend.CUDA``):

    import xlb_b200.compat; xlb_b200.compat.install_as_xlb()
    import runpy; runpy.run_path("examples/cfd/lid_driven_cavity_2d.py", run_name="__main__")

Sub-packages of the reference that are outside this package's scope (``xlb.utils``: VTK / image / USD output) are NOT
provided; importing them raises ``ModuleNotFoundError`` exactly as for any missing module, and a caller that wants to
run a script's post-processing supplies its own ``xlb.utils``.
"""

import importlib
import importlib.abc
import importlib.util
import sys

_ALIAS = "xlb"
_TARGET = __name__.rsplit(".", 1)[0]  # "xlb_b200"


class _AliasLoader(importlib.abc.Loader):
    def __init__(self, real_name):
        self.real_name = real_name

    def create_module(self, spec):
        return importlib.import_module(self.real_name)  # the existing module object: no second copy

    def exec_module(self, module):
        pass  # already executed under its real name


class _AliasFinder(importlib.abc.MetaPathFinder):
    def find_spec(self, fullname, path=None, target=None):
        if fullname != _ALIAS and not fullname.startswith(_ALIAS + "."):
            return None
        real = _TARGET + fullname[len(_ALIAS):]
        try:
            found = importlib.util.find_spec(real)
        except (ImportError, ValueError):
            found = None
        if found is None:
            return None  # e.g. xlb.utils: not part of this package -> the normal ModuleNotFoundError
        return importlib.util.spec_from_loader(fullname, _AliasLoader(real), is_package=found.submodule_search_locations is not None)


_finder = None


def install_as_xlb():
    """Idempotent.  Refuses to shadow a real ``xlb`` that is already imported."""
    global _finder
    existing = sys.modules.get(_ALIAS)
    target = importlib.import_module(_TARGET)
    if existing is not None and existing is not target:
        raise RuntimeError("a different 'xlb' package is already imported; cannot alias xlb -> xlb_b200 in this process")
    if _finder is None:
        _finder = _AliasFinder()
        sys.meta_path.insert(0, _finder)
    sys.modules[_ALIAS] = target
    return target


def uninstall():
    global _finder
    if _finder is not None:
        sys.meta_path.remove(_finder)
        _finder = None
    for name in [n for n in sys.modules if n == _ALIAS or n.startswith(_ALIAS + ".")]:
        del sys.modules[name]
