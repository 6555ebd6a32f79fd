"""Builds libxlb_b200.so in-tree with nvcc for sm_100a (no JIT cache; the .so travels with the repo).

One object per (lattice, precision policy) instance of csrc/xlb_step_inst.cu plus csrc/xlb_capi.cu,
compiled in parallel, then linked into ``xlb_b200/_lib/libxlb_b200.so``.
"""

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT_DIR = os.path.join(HERE, "_lib")
OBJ_DIR = os.path.join(HERE, "_lib", "obj")
LIB_PATH = os.path.join(OUT_DIR, "libxlb_b200.so")

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr", "-fmad=false", "-Xptxas", "-v"]

POLICIES = {  # tag -> (compute type, store type)
    "f32f32": ("float", "float"),
    "f32f16": ("float", "__half"),
    "f64f64": ("double", "double"),
    "f64f32": ("double", "float"),
    "f64f16": ("double", "__half"),
}
LATTICES = (9, 19, 27)
# per-policy tuning flags (B200 sweeps, profiles/r2k_sweep_diet.log, r2l_sweep_diet_bgk.log): the MUFU + Newton reciprocal is ahead
# for FP32 / FP64 storage (KBC +4 %, BGK +0.6 %) but costs the packed-half2 kernel of FP16 storage 3.7 %, which keeps __frcp_rn
POLICY_FLAGS = {"f32f16": ["-DXLB_TUNE_DIET_RCP=0"]}


def _nvcc():
    return shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"


def _sources_digest(extra=""):
    h = hashlib.sha256()
    for root in (CSRC, os.path.join(HERE, "..", "include")):
        for name in sorted(os.listdir(root)):
            if name.endswith((".cu", ".cuh", ".h")):
                with open(os.path.join(root, name), "rb") as fh:
                    h.update(name.encode())
                    h.update(fh.read())
    h.update(extra.encode())
    return h.hexdigest()


def _run(cmd, log, stamp=None, obj_digest=None):
    """Run one compile / link job.  With ``stamp``: skip it when the object exists and was built from the same sources and
    command line (so that changing one instance's flags does not recompile the other fifteen)."""
    if stamp is not None and os.path.exists(stamp) and os.path.exists(cmd[-1]) and open(stamp).read().strip() == obj_digest:
        return ""
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    with open(log, "w") as fh:
        fh.write(" ".join(cmd) + "\n" + r.stdout)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + r.stdout[-6000:])
    if stamp is not None:
        with open(stamp, "w") as fh:
            fh.write(obj_digest)
    return r.stdout


def build(force=False, verbose=False, extra_flags=(), lib_path=LIB_PATH, only=None, jobs=None):
    """Compile and link; returns the library path.  ``only`` restricts step instances to a set of tags
    like {"q19_f32f32"} (the other entry points then report XLB_EUNSUPPORTED) -- used by tuning sweeps."""
    os.makedirs(OBJ_DIR, exist_ok=True)
    src_digest = _sources_digest()
    digest = _sources_digest(" ".join(extra_flags) + str(sorted(only) if only else "") + str(sorted(POLICY_FLAGS.items())))
    stamp = lib_path + ".stamp"
    if not force and os.path.exists(lib_path) and os.path.exists(stamp) and open(stamp).read().strip() == digest:
        return lib_path
    nvcc = _nvcc()
    jobs_list = []
    objs = []
    tagbase = hashlib.md5(lib_path.encode()).hexdigest()[:6]
    for q in LATTICES:
        for tag, (tc, ts) in POLICIES.items():
            name = f"q{q}_{tag}"
            obj = os.path.join(OBJ_DIR, f"step_{name}_{tagbase}.o")
            flags = [f"-DXLB_Q={q}", f"-DXLB_TC={tc}", f"-DXLB_TS={ts}", f"-DXLB_TAG={name}", *POLICY_FLAGS.get(tag, [])]
            if only is not None and name not in only:
                flags.append("-DXLB_STUB=1")
            cmd = [nvcc, *ARCH, *COMMON, *flags, *extra_flags, "-c", os.path.join(CSRC, "xlb_step_inst.cu"), "-o", obj]  # extra_flags last: sweeps override
            jobs_list.append((cmd, obj + ".log", obj + ".stamp", hashlib.sha256((src_digest + " ".join(cmd)).encode()).hexdigest()))
            objs.append(obj)
    capi = os.path.join(OBJ_DIR, f"capi_{tagbase}.o")
    capi_cmd = [nvcc, *ARCH, *COMMON, "-c", os.path.join(CSRC, "xlb_capi.cu"), "-o", capi]
    jobs_list.append((capi_cmd, capi + ".log", capi + ".stamp", hashlib.sha256((src_digest + " ".join(capi_cmd)).encode()).hexdigest()))
    objs.append(capi)
    if force:
        jobs_list = [(cmd, log, None, None) for cmd, log, _, _ in jobs_list]
    workers = jobs or min(len(jobs_list), os.cpu_count() or 4)
    with ThreadPoolExecutor(max_workers=workers) as ex:
        outs = list(ex.map(lambda j: _run(*j), jobs_list))
    if verbose:
        for o in outs:
            sys.stdout.write(o)
    _run([nvcc, *ARCH, "-shared", "-o", lib_path, *objs, "-lcudart"], lib_path + ".link.log")
    with open(stamp, "w") as fh:
        fh.write(digest)
    return lib_path


if __name__ == "__main__":
    p = build(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(p)
