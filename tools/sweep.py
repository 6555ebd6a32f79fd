#!/usr/bin/env python
"""Tuning sweep of the fused step kernel.

    python tools/sweep.py build q19_f32f32 "V=4,BLOCK=128,MINB=1" "V=2,BLOCK=256,MINB=2" ...   (here, no GPU)
    python tools/sweep.py run cavity512_d3q19_bgk_fp32 [--steps 20]                             (on the GPU box)

``build`` compiles one library per variant into xlb_b200/_lib/sweep/ (only the named (lattice, policy) instance is
real, the others are stubs); ``run`` times every library found there with bench.py in a subprocess
(XLB_B200_LIB selects the library) and prints one line per variant.
"""

import glob
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
SWEEP_DIR = os.path.join(ROOT, "xlb_b200", "_lib", "sweep")


def build(instance, variants):
    from xlb_b200 import build as b

    os.makedirs(SWEEP_DIR, exist_ok=True)
    for var in variants:
        flags = [f"-DXLB_TUNE_{kv.split('=')[0]}={kv.split('=')[1]}" for kv in var.split(",") if kv]
        name = instance + "__" + var.replace(",", "_").replace("=", "")
        path = os.path.join(SWEEP_DIR, f"lib_{name}.so")
        b.build(force=True, extra_flags=flags, lib_path=path, only={instance})
        print("built", path)


def run(workload, steps, only="", variant=0, shape=None):
    rows = []
    for lib in sorted(glob.glob(os.path.join(SWEEP_DIR, f"lib_{only}*.so"))):
        env = dict(os.environ, XLB_B200_LIB=lib, XLB_B200_TUNING_LIB="1")
        r = subprocess.run(
            [sys.executable, os.path.join(ROOT, "bench.py"), "--workload", workload, "--steps", str(steps), "--warmup", "3", "--no-e2e", "--no-cpu-baseline", "--variant", str(variant)] + (["--shape", shape] if shape else []) + os.environ.get("XLB_SWEEP_BENCH_ARGS", "").split(),
            env=env, capture_output=True, text=True,
        )
        line = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
        if r.returncode != 0 or not line:
            print(os.path.basename(lib), "FAILED", r.stderr[-300:].replace("\n", " | "))
            continue
        j = json.loads(line[-1])
        rows.append((j["value"], os.path.basename(lib)))
        print(f"{os.path.basename(lib):60s} {j['value']:10.1f} MLUPS  frac={j['roofline']['frac']:.3f} kernel_ms={j['roofline']['kernel_ms']:.3f}", flush=True)
    rows.sort(reverse=True)
    print("BEST", rows[:3])


if __name__ == "__main__":
    if sys.argv[1] == "build":
        build(sys.argv[2], sys.argv[3:])
    else:
        steps = 20
        if "--steps" in sys.argv:
            steps = int(sys.argv[sys.argv.index("--steps") + 1])
        only = sys.argv[sys.argv.index("--only") + 1] if "--only" in sys.argv else ""
        variant = int(sys.argv[sys.argv.index("--variant") + 1]) if "--variant" in sys.argv else 0
        shape = sys.argv[sys.argv.index("--shape") + 1] if "--shape" in sys.argv else None
        run(sys.argv[2], steps, only, variant, shape)
