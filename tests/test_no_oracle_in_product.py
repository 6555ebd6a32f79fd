"""The product never imports the oracle and has no CPU fallback (task rule 3)."""

import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _py_files(top):
    for base, _, names in os.walk(top):
        if "__pycache__" in base:
            continue
        for n in names:
            if n.endswith(".py"):
                yield os.path.join(base, n)


def test_product_does_not_import_oracle():
    pat = re.compile(r"^\s*(from|import)\s+oracle\b", re.M)
    for path in _py_files(os.path.join(ROOT, "xlb_b200")):
        assert not pat.search(open(path).read()), f"{path} imports the oracle"


def test_product_does_not_import_jax_warp_triton():
    pat = re.compile(r"^\s*(from|import)\s+(jax|warp|triton|neon)\b", re.M)
    for path in _py_files(os.path.join(ROOT, "xlb_b200")):
        assert not pat.search(open(path).read()), f"{path} imports a foreign backend"


def test_bench_uses_oracle_only_in_cpu_legs():
    src = open(os.path.join(ROOT, "bench.py")).read()
    for m in re.finditer(r"^\s*(from|import)\s+oracle\b.*$", src, re.M):
        # every oracle import must sit inside a function whose name marks it as a CPU-baseline leg
        head = src[: m.start()]
        fn = re.findall(r"^def\s+(\w+)", head, re.M)[-1]
        assert fn in ("cpu_baseline", "run_reference"), f"bench.py imports oracle inside {fn}()"
