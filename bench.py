#!/usr/bin/env python
"""bench.py -- MLUPS of the fused LBM step (BASELINE.json metric) on N B200s of one node.

    python bench.py --gpus 1 --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --steps K --warmup W    # the reference's CPU path (oracle port), host cores
    torchrun ... bench.py --gpus N ...                       # N ranks, x-slab decomposition, NCCL halo ring

A "step" is one pass of the hot path (``IncompressibleNavierStokesStepper.__call__``: pull-stream + BCs +
macroscopic + equilibrium + collide) over the whole lattice.  One JSON line is printed by rank 0.

Keys beyond the base contract: ``roofline`` (algorithmic bytes / measured kernel time vs the measured HBM
copy peak of MEASURED_PEAKS.json), ``cpu_baseline`` (the NumPy oracle port on this box's host cores, bounded
sample), ``e2e`` (same metric with HOST buffers: pinned host -> device copy of the populations, the step,
device -> host copy of the result, all inside the timed region).
"""

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
_STDOUT = sys.stdout
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "MLUPS"
UNIT = "MLUPS (1e6 lattice-cell updates / s)"

WORKLOADS = {
    # name: (lattice, collision, policy, shape, omega, case builder name)
    "cavity512_d3q19_bgk_fp32": ("D3Q19", "BGK", "FP32FP32", (512, 512, 512), 1.0, "cavity"),
    "cavity512_d3q19_bgk_fp64": ("D3Q19", "BGK", "FP64FP64", (512, 512, 512), 1.0, "cavity"),
    "cavity128_d3q19_bgk_fp32": ("D3Q19", "BGK", "FP32FP32", (128, 128, 128), 1.0, "cavity"),
    "cavity256_d3q19_bgk_fp32": ("D3Q19", "BGK", "FP32FP32", (256, 256, 256), 1.0, "cavity"),
    "sphere384_d3q27_kbc_fp32": ("D3Q27", "KBC", "FP32FP32", (384, 128, 128), 1.6, "sphere"),
    "sphere768_d3q27_kbc_fp32": ("D3Q27", "KBC", "FP32FP32", (768, 256, 256), 1.6, "sphere"),
    "sphere384x256x512_d3q27_kbc_fp32": ("D3Q27", "KBC", "FP32FP32", (384, 256, 512), 1.6, "sphere"),
    "cavity384x256x512_d3q27_bgk_fp32": ("D3Q27", "BGK", "FP32FP32", (384, 256, 512), 1.7, "cavity"),
    "cavity768x256x256_d3q27_kbc_fp32": ("D3Q27", "KBC", "FP32FP32", (768, 256, 256), 1.6, "cavity"),
    "cavity768_d3q19_bgk_fp16": ("D3Q19", "BGK", "FP32FP16", (512, 768, 768), 1.0, "cavity"),
    "cavity512_d3q19_bgk_fp16": ("D3Q19", "BGK", "FP32FP16", (512, 512, 512), 1.0, "cavity"),
    "cavity1024_d3q19_bgk_fp16": ("D3Q19", "BGK", "FP32FP16", (1024, 1024, 1024), 1.0, "cavity"),
    # 1024^3 in FP32: 81.6 GB per population buffer -- one B200 holds it only with --in-place (one buffer, the AA pattern)
    "cavity1024_d3q19_bgk_fp32": ("D3Q19", "BGK", "FP32FP32", (1024, 1024, 1024), 1.0, "cavity"),
    "cavity640_d3q19_bgk_fp16": ("D3Q19", "BGK", "FP32FP16", (640, 640, 640), 1.0, "cavity"),
    # BASELINE configs[4] (windtunnel_3d.py with a sphere instead of the STL car): 1536 x 1536 cross-section,
    # 192 planes per GPU (weak scaling: 1536^3 on 8 GPUs); omega from Re = 50 000 like windtunnel_3d.py:42-45
    "windtunnel192x1536x1536_d3q27_kbc_fp32": ("D3Q27", "KBC", "FP32FP32", (192, 1536, 1536), 1.0 / (3.0 * (0.02 * 1535 / 50000.0) + 0.5), "windtunnel"),
}
STORE_BYTES = {"FP32FP32": 4, "FP64FP64": 8, "FP32FP16": 2, "FP64FP32": 4, "FP64FP16": 2}
DTYPE_NAME = {"FP32FP32": "f32", "FP64FP64": "f64", "FP32FP16": "f32 (f16 storage)", "FP64FP32": "f64 (f32 storage)", "FP64FP16": "f64 (f16 storage)"}


def b_alg(q, policy):
    """Algorithmic bytes per cell update: q reads + q writes in the store dtype + 1 bc_mask byte."""
    return 2 * q * STORE_BYTES[policy] + 1


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# --------------------------------------------------------------------------- clocks sampling
def ncu_traffic(workload, variant):
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch of this workload's kernel, from the committed
    ``ncu --set full`` capture listed in profiles/ncu_traffic.json (bytes per launch), else None."""
    if variant != 0:
        return None
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as fh:
            return json.load(fh).get(workload, {}).get("dram_bytes_per_launch")
    except (OSError, ValueError):
        return None


class ClockSampler:
    QUERY = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index=0):
        self.index = index
        self.rows = []
        self._stop = threading.Event()
        self._spawned = threading.Event()
        self._t = None

    def _run(self):
        while not self._stop.is_set():
            try:
                proc = subprocess.Popen(
                    ["nvidia-smi", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits", "-i", str(self.index)],
                    stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True,
                )
                self._spawned.set()
                out = proc.communicate(timeout=5)[0].strip()
                if out:
                    self.rows.append([c.strip() for c in out.split(",")])
            except Exception:
                self._spawned.set()
            self._stop.wait(0.2)

    def start(self):
        """Returns once the first nvidia-smi process exists: creating it (a fork of this process, with the GIL held) must not
        fall between the start event and the first launch of the timed region; the query itself then runs during the region."""
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()
        self._spawned.wait(timeout=5)

    def stop(self):
        self._stop.set()
        if self._t:
            self._t.join(timeout=6)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
            except Exception:
                continue
            for n, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {
            "sm_mhz": sm[len(sm) // 2] if sm else None,
            "sm_max_mhz": max(mx) if mx else None,
            "reasons": sorted(reasons),
            "samples": len(sm),
        }


# --------------------------------------------------------------------------- case construction (product objects)
def build_case(workload, slab=None, shape=None, variant=0, radius=None, backend_config=None):
    """Product-side objects of a workload.  BC index sets are generated per slab (global coords) so that the
    1536^3-class cases never materialise a global index list on one rank."""
    import numpy as np

    import xlb_b200 as xlb
    from xlb_b200.compute_backend import ComputeBackend
    from xlb_b200.operator.boundary_condition import (
        EquilibriumBC, ExtrapolationOutflowBC, FullwayBounceBackBC, HalfwayBounceBackBC, ZouHeBC,
    )
    from xlb_b200.operator.stepper import IncompressibleNavierStokesStepper
    from xlb_b200.precision_policy import PrecisionPolicy

    lattice, collision, policy, shape0, omega, kind = WORKLOADS[workload]
    shape = tuple(shape) if shape is not None else shape0
    pol = PrecisionPolicy[policy]
    vs = getattr(xlb.velocity_set, lattice)(precision_policy=pol, compute_backend=ComputeBackend.CUDA)
    xlb.init(velocity_set=vs, default_backend=ComputeBackend.CUDA, default_precision_policy=pol)
    if slab is not None:
        grid = slab.make_grid(shape)
    else:
        grid = xlb.grid.grid_factory(shape, compute_backend=ComputeBackend.CUDA, backend_config={"slab": False})  # one whole block, also under torchrun
    nx, ny, nz = shape
    box = grid.bounding_box_indices()
    box_ne = grid.bounding_box_indices(remove_edges=True)

    def union(names):
        idx = np.concatenate([np.array(box[n]) for n in names], axis=1)
        return np.unique(idx, axis=-1)

    if kind == "cavity":  # examples/performance/mlups_3d.py:194-204
        bcs = [
            EquilibriumBC(rho=1.0, u=(0.02, 0.0, 0.0), indices=box_ne["top"]),
            FullwayBounceBackBC(indices=union(["bottom", "left", "right", "front", "back"])),
        ]
    else:  # examples/cfd/flow_past_sphere_3d.py:38-113 with Zou-He inlet (BASELINE config 3)
        r = ny // 12
        cx, cy, cz = nx // 6, ny // 2, nz // 2
        u_in = 0.04
        if kind == "windtunnel":  # windtunnel_3d.py:78-106 with a sphere obstacle
            r = min(ny // 8, nx // 4)
            cx = nx // 3
            u_in = 0.02
        if radius is not None:
            r = int(radius)
        xs = np.arange(max(cx - r, 0), min(cx + r + 1, nx))
        ys = np.arange(max(cy - r, 0), min(cy + r + 1, ny))
        zs = np.arange(max(cz - r, 0), min(cz + r + 1, nz))
        X, Y, Z = np.meshgrid(xs, ys, zs, indexing="ij")
        sel = (X - cx) ** 2 + (Y - cy) ** 2 + (Z - cz) ** 2 < r**2
        sphere = np.stack([X[sel], Y[sel], Z[sel]])
        bcs = [
            FullwayBounceBackBC(indices=union(["bottom", "top", "front", "back"])),
            ZouHeBC("velocity", prescribed_value=(u_in, 0.0, 0.0), indices=box_ne["left"]),
            ExtrapolationOutflowBC(indices=box_ne["right"]),
            HalfwayBounceBackBC(indices=sphere),
        ]
    cfg = {"kernel_variant": variant}
    cfg.update(backend_config or {})
    stepper = IncompressibleNavierStokesStepper(grid=grid, boundary_conditions=bcs, collision_type=collision, backend_config=cfg)
    return stepper, vs, pol, shape, omega


# relative cost of a cell of a large half-way obstacle (all solid cells are tagged; they take the half-way path of the fused kernel):
# a 192 x 1536 x 1536 block that holds half of the r = 192 sphere (3.3 % of its cells) runs 18 % slower than one without
# (profiles/r2i_bigsphere.log vs r2m_bench_default.json) -> 1 + 0.18 / 0.033
HALFWAY_CELL_COST = 6.5


def obstacle_plane_cost(kind, shape, radius=None):
    """Relative cost of every global x plane of a sphere / wind-tunnel workload (same geometry as build_case)."""
    import numpy as np

    nx, ny, nz = shape
    r, cx = ny // 12, nx // 6
    if kind == "windtunnel":
        r, cx = min(ny // 8, nx // 4), nx // 3
    if radius is not None:
        r = int(radius)
    t = np.arange(nx) - cx
    solid = np.pi * np.clip(r * r - t * t, 0, None)  # cells of the sphere's cross-section in plane x
    return ny * nz + (HALFWAY_CELL_COST - 1.0) * solid


# --------------------------------------------------------------------------- CPU legs (the only users of oracle/)
def host_threads():
    """Host threads this process may use (torchrun exports OMP_NUM_THREADS=1, which is not the box's core count)."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def probe_reference_jax():
    """SURVEY 8(c) "extra safety" / BASELINE.md 3(2a): is the real reference (xlb on JAX / XLA:CPU) runnable on this box?
    Returns (module or None, one-line reason).  ``baseline/_ref`` is where an offline install of the reference would live."""
    ref_dir = os.path.join(ROOT, "baseline", "_ref")
    if os.path.isdir(ref_dir) and ref_dir not in sys.path:
        sys.path.append(ref_dir)
    try:
        import jax  # noqa: F401
    except Exception as exc:  # ModuleNotFoundError on this image (no jax wheel in /opt/wheelhouse)
        return None, f"not importable ({type(exc).__name__}: {exc})"
    try:
        import xlb as ref_xlb  # the unmodified reference package, if it was installed into baseline/_ref
    except Exception as exc:
        return None, f"jax imports but the reference package does not ({type(exc).__name__}: {exc})"
    return ref_xlb, "importable"


def reference_jax_cavity(ref_xlb, n, omega, budget_s, max_steps):
    """The reference's own mlups_3d.py loop (examples/performance/mlups_3d.py:177-262) on its JAX backend, CPU devices,
    with block_until_ready for honest timing.  Only reached when ``probe_reference_jax`` found jax + xlb on this box."""
    import jax
    import numpy as np
    from xlb.compute_backend import ComputeBackend
    from xlb.grid import grid_factory
    from xlb.operator.boundary_condition import EquilibriumBC, FullwayBounceBackBC
    from xlb.operator.stepper import IncompressibleNavierStokesStepper
    from xlb.precision_policy import PrecisionPolicy

    vs = ref_xlb.velocity_set.D3Q19(precision_policy=PrecisionPolicy.FP32FP32, compute_backend=ComputeBackend.JAX)
    ref_xlb.init(velocity_set=vs, default_backend=ComputeBackend.JAX, default_precision_policy=PrecisionPolicy.FP32FP32)
    grid = grid_factory((n, n, n), compute_backend=ComputeBackend.JAX)
    box = grid.bounding_box_indices()
    box_ne = grid.bounding_box_indices(remove_edges=True)
    walls = [box["bottom"][i] + box["left"][i] + box["right"][i] + box["front"][i] + box["back"][i] for i in range(3)]
    walls = np.unique(np.array(walls), axis=-1).tolist()
    bcs = [EquilibriumBC(rho=1.0, u=(0.02, 0.0, 0.0), indices=box_ne["top"]), FullwayBounceBackBC(indices=walls)]
    st = IncompressibleNavierStokesStepper(grid=grid, boundary_conditions=bcs, collision_type="BGK")
    f_0, f_1, bc_mask, missing = st.prepare_fields()
    f_0, f_1 = st(f_0, f_1, bc_mask, missing, omega, 0)
    f_0, f_1 = f_1, f_0
    jax.block_until_ready(f_0)
    t0 = time.perf_counter()
    steps = 0
    while steps < max_steps and (time.perf_counter() - t0) < budget_s:
        f_0, f_1 = st(f_0, f_1, bc_mask, missing, omega, steps + 1)
        f_0, f_1 = f_1, f_0
        jax.block_until_ready(f_0)
        steps += 1
    return steps, time.perf_counter() - t0, jax.device_count()


def cpu_baseline(workload, budget_s=20.0, max_steps=8):
    """The reference's CPU path for this workload, timed on this box's host cores on a bounded sample: the same BCs /
    lattice / policy at 128^3 (BASELINE configs[0] when the workload is the cavity).  Order of preference: the real
    reference on JAX / XLA:CPU (kind "reference") when jax and the reference package import here; else the oracle port of
    nse_stepper.py:238-282 (kind "port") -- for the cavity the multi-threaded torch-CPU transcription of the oracle on all
    host threads (what XLA:CPU would use), for the other workloads the single-threaded NumPy oracle."""
    import numpy as np
    import torch

    from oracle import cases

    lattice, collision, policy, shape, omega, kind = WORKLOADS[workload]
    ref_xlb, jax_note = probe_reference_jax()
    if ref_xlb is not None and kind == "cavity" and lattice == "D3Q19" and collision == "BGK" and policy == "FP32FP32":
        steps, dt, ndev = reference_jax_cavity(ref_xlb, 128, omega, budget_s, max_steps)
        return {"value": 128**3 * steps / dt / 1e6, "unit": UNIT, "cores": host_threads(), "kind": "reference", "jax": jax_note,
                "sample": f"{steps} steps of the reference's mlups_3d.py loop at 128^3 on JAX / XLA:CPU ({ndev} device(s))", "seconds": dt,
                "steps": steps, "sample_shape": [128, 128, 128]}
    if kind == "cavity":
        case = cases.cavity3d(128, policy=policy, lattice=lattice, collision=collision, omega=omega)
    else:
        case = cases.sphere3d((96, 64, 64), policy=policy, lattice=lattice, collision=collision, omega=omega)
    st, _ = cases.build_oracle(case)
    f0, f1, bc, mm = st.prepare_fields()
    f0 = f0.astype(st.sdt)
    ncell = int(np.prod(case["shape"]))
    cores, how = 1, "NumPy oracle, single thread"
    step = st
    if kind == "cavity" and collision == "BGK":
        from oracle.torch_cpu import TorchCavityStepper, to_torch

        torch.set_num_threads(host_threads())  # torchrun sets OMP_NUM_THREADS=1 for its workers
        step = TorchCavityStepper(st)
        f0, bc = to_torch(f0), to_torch(bc)
        cores, how = torch.get_num_threads(), f"torch-CPU transcription of the oracle, {torch.get_num_threads()} threads"
    f0, f1 = step(f0, f1, bc, mm, omega, 0)  # warm-up (page faults, caches)
    f0, f1 = f1, f0
    t0 = time.perf_counter()
    steps = 0
    while steps < max_steps and (time.perf_counter() - t0) < budget_s:
        f0, f1 = step(f0, f1, bc, mm, omega, steps + 1)
        f0, f1 = f1, f0
        steps += 1
    dt = time.perf_counter() - t0
    return {
        "value": ncell * steps / dt / 1e6,
        "unit": UNIT,
        "cores": cores,
        "kind": "port",
        "jax": jax_note,
        "sample": f"{steps} steps of {case['name']} ({ncell} cells), {how}; oracle port of nse_stepper.py:238-282; host has {os.cpu_count()} cores",
        "seconds": dt,
        "steps": steps,
        "sample_shape": list(case["shape"]),
    }


CPU_KEYS = ("value", "unit", "cores", "kind", "sample", "sample_shape", "jax")


def run_reference(args):
    """--impl reference: the reference's CPU path on this box's host cores (the real JAX backend when it imports here,
    else the oracle port), on a bounded sample of the workload.  ``config.shape`` is the shape that actually ran;
    ``config.sample_of`` names the workload's full shape."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    workload = args.workload
    lattice, collision, policy, shape, omega, kind = WORKLOADS[workload]
    base = cpu_baseline(workload, budget_s=max(20.0, 6.0 * args.steps), max_steps=max(args.steps, 1))
    line = {
        "impl": "reference",
        "metric": METRIC,
        "value": base["value"],
        "unit": UNIT,
        "n_gpus": args.gpus,
        "steps": args.steps,
        "warmup": args.warmup,
        "ms_per_step": base["seconds"] / max(base["steps"], 1) * 1e3,
        "higher_is_better": True,
        "scaling": "weak",
        "vs_baseline": None,
        "dtype": DTYPE_NAME[policy],
        "data": "synthetic",
        "config": {"workload": workload, "lattice": lattice, "collision": collision, "policy": policy, "shape": base["sample_shape"],
                   "sample_of": list(shape), "omega": omega, "host_threads": base["cores"]},
        "cpu_baseline": {k: base[k] for k in CPU_KEYS},
        "e2e": {"value": base["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    _STDOUT.write(json.dumps(line) + "\n")
    _STDOUT.flush()
    return 0


# --------------------------------------------------------------------------- native arm
def time_steps(stepper, fields, omega, steps, warmup, graph=0, t0=0):
    """``warmup`` untimed steps, then ``steps`` timed ones between two CUDA events on the launching stream.
    Returns (ms per step, kernels launched per step, fields).  ``graph`` > 0 replays GraphedSteps of that many steps."""
    import torch

    from xlb_b200 import _lib

    f_0, f_1, bc_mask, missing = fields
    t = t0
    for _ in range(warmup):
        f_0, f_1 = stepper(f_0, f_1, bc_mask, missing, omega, t)
        f_0, f_1 = f_1, f_0
        t += 1
    graphed = None
    if graph:
        from xlb_b200.operator.stepper import GraphedSteps

        steps = max(graph, steps // graph * graph)
        c0 = _lib.launch_count()
        graphed = GraphedSteps(stepper, f_0, f_1, bc_mask, missing, omega, steps=graph)
        lps = (_lib.launch_count() - c0) / (graph + 2)
        for _ in range(3):
            graphed()
    torch.cuda.synchronize()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    c0 = _lib.launch_count()
    ev0.record()
    if graphed is not None:
        for _ in range(steps // graph):
            graphed()
    else:
        for _ in range(steps):
            f_0, f_1 = stepper(f_0, f_1, bc_mask, missing, omega, t)
            f_0, f_1 = f_1, f_0
            t += 1
    ev1.record()
    torch.cuda.synchronize()
    if graphed is None:
        lps = (_lib.launch_count() - c0) / steps
    return ev0.elapsed_time(ev1) / steps, lps, (f_0, f_1, bc_mask, missing)


# the other BASELINE configs, timed in the same run as the headline: name -> (workload, shape or None, steps, graph)
EXTRA_CONFIGS = [
    ("C1_cavity128_d3q19_bgk_fp32_eager", "cavity128_d3q19_bgk_fp32", None, 400, 0),
    ("C1_cavity128_d3q19_bgk_fp32_graph20", "cavity128_d3q19_bgk_fp32", None, 400, 20),
    ("C2_cavity512_d3q19_bgk_fp64", "cavity512_d3q19_bgk_fp64", None, 20, 0),
    ("C3_sphere384x128x128_d3q27_kbc_fp32", "sphere384_d3q27_kbc_fp32", None, 200, 0),
    ("C3_sphere384x128x128_d3q27_kbc_fp32_graph20", "sphere384_d3q27_kbc_fp32", None, 200, 20),
    ("sphere768x256x256_d3q27_kbc_fp32", "sphere768_d3q27_kbc_fp32", None, 40, 0),
    ("cavity384x256x512_d3q27_bgk_fp32", "cavity384x256x512_d3q27_bgk_fp32", None, 40, 0),
    ("C4block_cavity512x768x768_d3q19_bgk_fp16", "cavity768_d3q19_bgk_fp16", None, 20, 0),
    ("C5block_windtunnel192x1536x1536_d3q27_kbc_fp32", "windtunnel192x1536x1536_d3q27_kbc_fp32", None, 10, 0),
]


def run_extra_configs(peak, budget_s=150.0):
    """MLUPS / kernel time / roofline fraction of the other single-GPU BASELINE configurations (device-timed like the
    headline: CUDA events around the step loop, inputs larger than L2 except the 128^3 case, which says so)."""
    import torch

    out = {}
    t_start = time.perf_counter()
    for name, workload, shape, steps, graph in EXTRA_CONFIGS:
        if time.perf_counter() - t_start > budget_s:
            out[name] = {"skipped": "time budget of the configs block used up"}
            continue
        lattice, collision, policy, shape0, omega, kind = WORKLOADS[workload]
        try:
            stepper, vs, pol, shp, omega = build_case(workload, shape=shape)
            fields = stepper.prepare_fields()
            ms, lps, fields = time_steps(stepper, fields, omega, steps, 5, graph=graph)
            ncell = shp[0] * shp[1] * shp[2]
            bpc = b_alg(vs.q, policy)
            gbs = bpc * ncell / (ms * 1e-3) / 1e9
            out[name] = {
                "workload": workload, "shape": list(shp), "lattice": lattice, "collision": collision, "policy": policy,
                "mlups": ncell / (ms * 1e-3) / 1e6, "kernel_ms": ms, "achieved_gbs": gbs, "frac": gbs / peak, "bytes_per_cell": bpc,
                "launches_per_step": lps, "steps": steps, **({"graph_steps": graph} if graph else {}),
                "finite": bool(torch.isfinite(fields[0].float().sum()).item()),
                **({"l2": "working set (2 x 159 MB) is of the order of the 126 MB L2"} if ncell <= 128**3 else {}),
            }
            del stepper, fields
        except Exception as exc:  # a config that cannot run (e.g. memory) must not take the headline line down
            out[name] = {"error": f"{type(exc).__name__}: {exc}"[:300]}
        torch.cuda.empty_cache()
    return out


def nrank_bitwise_check(workload, slab, world, steps=5):
    """N-rank == 1-rank, bitwise (SURVEY 8(e) correctness check), inside the bench so that the driver's scaling run carries
    the proof: a small global block of the same workload family is stepped ``steps`` times on the slab stepper (all ranks,
    halo ring, overlap schedule) and, redundantly on every rank, as one block; the gathered slabs must equal it exactly."""
    import torch

    from xlb_b200.distribute import distribute

    lattice, collision, policy, shape0, omega, kind = WORKLOADS[workload]
    from xlb_b200.operator.boundary_condition.boundary_condition_registry import boundary_condition_registry

    gshape = (6 * world + 1, 40, 128)  # uneven slabs (remainder plane on rank 0)
    first_id = boundary_condition_registry.next_id
    single, vs, _, _, omega = build_case(workload, shape=gshape)
    s_0, s_1, s_bc, s_mm = single.prepare_fields()
    gen = torch.Generator(device="cpu")
    gen.manual_seed(1234)
    noise = torch.rand((vs.q,) + gshape, generator=gen, dtype=torch.float32)
    w = torch.tensor(vs.w, dtype=torch.float32).reshape(-1, 1, 1, 1)
    init = (w * (1.0 + 0.01 * (2.0 * noise - 1.0))).to(s_0.dtype).cuda()
    s_0.copy_(init)
    boundary_condition_registry.next_id = first_id  # same BC ids in both steppers, so that the masks compare equal as they are
    part, _, _, _, _ = build_case(workload, slab=slab, shape=gshape)
    dstep = distribute(part, part.grid, vs)
    p_0, p_1, p_bc, p_mm = dstep.prepare_fields()
    x0, n = slab.local_range(gshape[0])
    p_0[:, 1 : n + 1].copy_(init[:, x0 : x0 + n])
    dstep.halo.exchange(p_0)
    torch.cuda.synchronize()
    for t in range(steps):
        s_0, s_1 = single(s_0, s_1, s_bc, s_mm, omega, t)
        s_0, s_1 = s_1, s_0
        p_0, p_1 = dstep(p_0, p_1, p_bc, p_mm, omega, t)
        p_0, p_1 = p_1, p_0
    torch.cuda.synchronize()
    same = bool(torch.equal(dstep.gather(p_0), s_0)) and bool(torch.equal(dstep.gather(p_bc), s_bc))
    mode = "peer push" if dstep.peer is not None else "nccl send/recv"
    dstep.close()
    flag = torch.tensor([1 if same else 0], device="cuda")
    import torch.distributed as dist

    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    return bool(flag.item()), {"global_shape": list(gshape), "steps": steps, "workload_family": workload, "halo": mode}


def run_native(args):
    import torch
    import torch.distributed as dist

    from xlb_b200 import _lib

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device (xlb_b200 has no CPU fallback)")
    torch.cuda.set_device(local_rank)
    slab = None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        from xlb_b200.distribute import SlabDecomposition

        slab = SlabDecomposition(rank=rank, world=world)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world > 1:
            tt = torch.tensor([ms], device="cuda", dtype=torch.float64)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            return float(tt.item())
        return ms

    def timed_run(workload, shape, strong, steps, warmup):
        """Build, warm up, time exactly ``steps`` steps (barrier + synchronize on both sides, CUDA events, max over ranks)."""
        lattice, collision, policy, shape0, omega, kind = WORKLOADS[workload]
        shape = tuple(shape) if shape else shape0
        if world > 1 and not strong:
            shape = (shape[0] * world,) + tuple(shape[1:])  # weak scaling: the N=1 block per GPU, stacked along x
        run_slab = slab
        planes = None
        if slab is not None and kind != "cavity" and not args.no_balance:
            # obstacle workloads: slabs of equal COST, not of equal plane count (every rank waits for the slowest one each step)
            from xlb_b200.distribute import SlabDecomposition, balanced_planes

            planes = balanced_planes(obstacle_plane_cost(kind, shape, args.radius), world)
            run_slab = SlabDecomposition(rank=rank, world=world, planes=planes)
        cfg = {"in_place": True} if args.in_place else {}
        if args.bc_cell_list != "auto":
            cfg["bc_cell_list"] = args.bc_cell_list == "on"
        stepper, vs, pol, shape, omega = build_case(workload, slab=run_slab, shape=shape, variant=args.variant, radius=args.radius, backend_config=cfg or None)
        if slab is not None:
            from xlb_b200.distribute import distribute

            stepper = distribute(stepper, stepper.grid, vs, overlap=not args.no_overlap, graph=not args.no_dist_graph, halo=args.halo)
        f_0, f_1, bc_mask, missing = stepper.prepare_fields()
        t = 0
        for _ in range(warmup):
            f_0, f_1 = stepper(f_0, f_1, bc_mask, missing, omega, t)
            f_0, f_1 = f_1, f_0
            t += 1
        barrier()
        return dict(stepper=stepper, vs=vs, policy=policy, shape=shape, omega=omega, fields=(f_0, f_1, bc_mask, missing), t=t,
                    lattice=lattice, collision=collision, ncell=shape[0] * shape[1] * shape[2], planes=planes)

    workload = args.workload
    shape_arg = tuple(int(v) for v in args.shape.split(",")) if args.shape else None
    warmup = max(args.warmup, 3)
    run = timed_run(workload, shape_arg, args.strong, args.steps, warmup)
    stepper, vs, policy, shape, omega, ncell = run["stepper"], run["vs"], run["policy"], run["shape"], run["omega"], run["ncell"]
    lattice, collision = run["lattice"], run["collision"]
    f_0, f_1, bc_mask, missing = run["fields"]
    t = run["t"]
    slab_planes = run.get("planes")  # ``run`` is cleared before the strong-scaling leg

    # ---- optional: replay the step loop as CUDA graphs of --graph steps each (launch-bound small lattices)
    graphed, graph_lps = None, None
    if args.graph and world == 1:
        from xlb_b200.operator.stepper import GraphedSteps

        if args.steps % args.graph:
            raise SystemExit("--steps must be a multiple of --graph")
        c0 = _lib.launch_count()
        graphed = GraphedSteps(stepper, f_0, f_1, bc_mask, missing, omega, steps=args.graph)
        graph_lps = (_lib.launch_count() - c0) / (args.graph + 2)  # kernels per step (2 eager warm-up steps + the capture)
        for _ in range(3):
            graphed()
        torch.cuda.synchronize()

    # ---- timed region: exactly K steps, CUDA events on the launching stream
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    launches0 = _lib.launch_count()
    ev0 = torch.cuda.Event(enable_timing=True)
    ev1 = torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record()
    if graphed is not None:
        for _ in range(args.steps // args.graph):
            graphed()
        t += args.steps
    else:
        for _ in range(args.steps):
            f_0, f_1 = stepper(f_0, f_1, bc_mask, missing, omega, t)
            f_0, f_1 = f_1, f_0
            t += 1
    ev1.record()
    barrier()
    ms = max_over_ranks(ev0.elapsed_time(ev1))
    launches = _lib.launch_count() - launches0
    if graphed is not None:
        launches = int(round(graph_lps * args.steps))  # kernels replayed from the graphs (no host-side launch calls)
    elif world > 1 and getattr(stepper, "launches_per_step", None):
        launches = int(round(stepper.launches_per_step * args.steps))  # replayed from the distributed step's CUDA graph
    clocks = sampler.stop() if rank == 0 else None
    mlups = ncell * args.steps / (ms * 1e-3) / 1e6
    finite = bool(torch.isfinite(f_0.float().sum()).item())

    # ---- roofline of the dominant (only) kernel: per-launch CUDA-event time on the launching stream
    peak, peak_src = measured_peaks()
    roofline = None
    if world == 1:
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(min(args.steps, 10))]
        torch.cuda.synchronize()
        for a, b in evs:
            a.record()
            if graphed is not None:
                graphed()
            else:
                f_0, f_1 = stepper(f_0, f_1, bc_mask, missing, omega, t)
                f_0, f_1 = f_1, f_0
            b.record()
            t += 1
        torch.cuda.synchronize()
        per = sorted(a.elapsed_time(b) for a, b in evs)
        k_ms = sum(per) / len(per) / (args.graph if graphed is not None else 1)
        bytes_per_launch = b_alg(vs.q, policy) * ncell
        achieved = bytes_per_launch / (k_ms * 1e-3) / 1e9
        traffic = ncu_traffic(workload, args.variant)
        roofline = {
            "bound": "hbm",
            "kernel": {0: "xlb_step (auto variant)", 2: "step_vector_kernel", 3: "step_generic_kernel", 4: "step_bulk_kernel"}.get(args.variant, "xlb_step"),
            "achieved": achieved,
            "peak": peak,
            "unit": "GB/s",
            "frac": achieved / peak,
            "traffic": traffic,
            "traffic_source": "committed ncu --set full capture (profiles/ncu_traffic.json), not re-measured in this run" if traffic else None,
            "frac_of_nominal_8000": achieved / 8000.0,
            "bytes_per_launch": bytes_per_launch,
            "bytes_per_cell": b_alg(vs.q, policy),
            "cells_per_launch": ncell,
            "kernel_ms": k_ms,
            "launches_per_step": launches / max(1, args.steps),  # 2 = fused kernel + dense launch over the BC cell list
            "peak_source": peak_src,
        }

    if world > 1:
        # per-GPU roofline of the whole distributed step (edge-plane launch + interior launch + halo ring):
        # algorithmic bytes of one rank's slab over the max-over-ranks step time
        per_rank = b_alg(vs.q, policy) * (ncell // world)
        achieved = per_rank / (ms / args.steps * 1e-3) / 1e9
        roofline = {
            "bound": "hbm", "kernel": "per-GPU slab step (edge-plane launch + interior launch, halo exchange overlapped)",
            "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
            "frac_of_nominal_8000": achieved / 8000.0, "bytes_per_launch": per_rank, "bytes_per_cell": b_alg(vs.q, policy),
            "cells_per_launch": ncell // world, "kernel_ms": ms / args.steps, "peak_source": peak_src,
        }

    # ---- e2e: HOST buffers in, HOST buffers out (pinned), copies inside the timed region
    e2e = e2e_resident = None
    if world == 1 and not args.no_e2e and not args.in_place:
        e2e = run_e2e(stepper, f_0, f_1, bc_mask, missing, omega, t, ncell, max(1, min(args.steps, args.e2e_steps)))
        e2e_resident = run_resident_api(stepper, f_0, f_1, bc_mask, missing, omega, t, ncell, min(args.steps, 20))

    # ---- the rest of the BASELINE metric on the same clock
    extra = {}
    halo_mode = None
    if world > 1:
        halo_mode = "peer push fused into the edge-plane launch (CUDA IPC over NVLink)" if stepper.peer is not None else "NCCL send/recv on a communication stream"
        if stepper.peer is not None and stepper.peer.timed_out():
            raise RuntimeError("a halo wait timed out during the timed region")
        stepper.close()
    del f_0, f_1, bc_mask, missing, stepper, graphed
    run.clear()
    torch.cuda.empty_cache()
    if world == 1 and not args.no_configs:
        extra["configs"] = run_extra_configs(peak)
    if world > 1 and not args.no_configs:
        ok, info = nrank_bitwise_check(workload, slab, world)
        extra["nrank_bitwise"] = ok
        extra["nrank_bitwise_case"] = info
        torch.cuda.empty_cache()
        if not args.strong:
            # BASELINE configs[3]: 1024^3 D3Q19 BGK FP32FP16, strong scaling (global shape fixed, slabs shrink with N)
            sw = "cavity1024_d3q19_bgk_fp16"
            srun = timed_run(sw, None, True, args.steps, warmup)
            st2, (g_0, g_1, g_bc, g_mm), t2 = srun["stepper"], srun["fields"], srun["t"]
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            barrier()
            e0.record()
            for _ in range(args.steps):
                g_0, g_1 = st2(g_0, g_1, g_bc, g_mm, srun["omega"], t2)
                g_0, g_1 = g_1, g_0
                t2 += 1
            e1.record()
            barrier()
            sms = max_over_ranks(e0.elapsed_time(e1))
            per_rank = b_alg(19, "FP32FP16") * (srun["ncell"] // world)
            extra["strong"] = {
                "workload": sw, "shape": list(srun["shape"]), "scaling": "strong", "n_gpus": world, "steps": args.steps,
                "mlups": srun["ncell"] * args.steps / (sms * 1e-3) / 1e6, "ms_per_step": sms / args.steps,
                "per_gpu_frac": per_rank / (sms / args.steps * 1e-3) / 1e9 / peak,
                "finite": bool(torch.isfinite(g_0.float().sum()).item()),
            }
            st2.close()
            del st2, g_0, g_1, g_bc, g_mm
            srun.clear()
            torch.cuda.empty_cache()

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0
    line = {
        "metric": METRIC,
        "value": mlups,
        "unit": UNIT,
        "n_gpus": world,
        "steps": args.steps,
        "warmup": warmup,
        "ms_per_step": ms / args.steps,
        "higher_is_better": True,
        "scaling": "strong" if (world > 1 and args.strong) else "weak",
        "vs_baseline": None,
        "dtype": DTYPE_NAME[policy],
        "data": "synthetic",
        "config": {
            "workload": workload, "lattice": lattice, "collision": collision, "policy": policy, "shape": list(shape), "omega": omega,
            "decomposition": f"x-slab x{world}" if world > 1 else "single block",
            **({"slab_planes": slab_planes, "slab_balance": "planes per rank from a per-plane cost: cells + %.1f x half-way obstacle cells" % (HALFWAY_CELL_COST - 1.0)}
               if slab_planes else {}),
            **({"halo": halo_mode} if halo_mode else {}),
            **({"graph_steps": args.graph} if args.graph and world == 1 else {}),
            **({"buffers": "one population buffer, in-place AA pattern (xlb_step_inplace)"} if args.in_place else {}),
            "l2": "inputs larger than L2 (%s of %.1f GB%s)" % ("one population buffer" if args.in_place else "two population buffers",
                                                                vs.q * ncell * STORE_BYTES[policy] / world / 1e9, "" if args.in_place else " each"),
        },
        "gpu_launches": int(launches),
        "finite": finite,
        "clocks": clocks,
    }
    if roofline:
        line["roofline"] = roofline
    if e2e:
        line["e2e"] = e2e
        line["e2e_resident_api"] = e2e_resident
    line.update(extra)
    if world == 1 and not args.no_cpu_baseline:
        base = cpu_baseline(workload)
        line["cpu_baseline"] = {k: base[k] for k in CPU_KEYS}
    _STDOUT.write(json.dumps(line) + "\n")
    _STDOUT.flush()
    if world > 1:
        dist.destroy_process_group()
    return 0


def run_e2e(stepper, f_0, f_1, bc_mask, missing, omega, t, ncell, steps, chunk_planes=32):
    """Same metric with HOST buffers through the public host-facing API (``HostStreamedStepper``): every step moves the
    populations pinned-host -> device, runs ``xlb_step`` and moves the result device -> pinned-host, as an x-chunk
    pipeline over three streams (masks / BC tables are static and stay resident).  A device->host read of one
    result element closes each step."""
    import torch

    from xlb_b200.operator.stepper import HostStreamedStepper

    from xlb_b200.helper.host_memory import pinned_empty

    h_in, placement = pinned_empty(f_0.shape, f_0.dtype)
    h_out, _ = pinned_empty(f_1.shape, f_1.dtype)
    h_in.copy_(f_0)
    torch.cuda.synchronize()
    hs = HostStreamedStepper(stepper, chunk_planes=chunk_planes)
    ev0 = torch.cuda.Event(enable_timing=True)
    ev1 = torch.cuda.Event(enable_timing=True)
    h_in, h_out = hs(h_in, h_out, bc_mask, missing, omega, t)  # one untimed pass
    h_in, h_out = h_out, h_in
    torch.cuda.synchronize()
    ev0.record()
    probe = 0.0
    for i in range(steps):
        h_in, h_out = hs(h_in, h_out, bc_mask, missing, omega, t + 1 + i)
        h_in, h_out = h_out, h_in
        torch.cuda.current_stream().synchronize()
        probe += float(h_in[0, 0, 0, 0])  # the step's result is on the host
    ev1.record()
    torch.cuda.synchronize()
    ms = ev0.elapsed_time(ev1)
    nbytes = f_0.numel() * f_0.element_size()
    halo = 2 * len(hs._state["chunks"]) * f_0.shape[0] * f_0.shape[2] * f_0.shape[3] * f_0.element_size()
    return {
        "value": ncell * steps / (ms * 1e-3) / 1e6,
        "unit": UNIT,
        "h2d_bytes_per_step": int(nbytes + halo),
        "d2h_bytes_per_step": int(nbytes),
        "steps": steps,
        "ms_per_step": ms / steps,
        "host_memory": placement,
        "note": f"HostStreamedStepper: pinned-host populations, {len(hs._state['chunks'])} x-chunks of {hs._state['P']} planes pipelined over h2d / compute / d2h streams; finite={probe == probe}",
    }


def run_resident_api(stepper, f_0, f_1, bc_mask, missing, omega, t, ncell, steps):
    """The call a user of the reference API makes (device-resident fields, ``stepper(...)`` + swap in a Python loop),
    with the per-step scalars going host -> device and one result element read back device -> host every step."""
    import torch

    torch.cuda.synchronize()
    t0 = time.perf_counter()
    probe = 0.0
    for i in range(steps):
        f_0, f_1 = stepper(f_0, f_1, bc_mask, missing, omega, t + i)
        f_0, f_1 = f_1, f_0
        probe += float(f_0[0, 0, 0, 0])  # blocking 4-byte read of this step's result
    dt = time.perf_counter() - t0
    return {"value": ncell * steps / dt / 1e6, "unit": UNIT, "steps": steps, "ms_per_step": dt / steps * 1e3,
            "note": "device-resident fields through the Python API, blocking D2H read of one result element per step (wall clock)"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--workload", default=None, choices=sorted(WORKLOADS))
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--in-place", action="store_true", help="one population buffer (in-place AA pattern) instead of two; single GPU, cavity / periodic workloads")
    ap.add_argument("--e2e-steps", type=int, default=4)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-overlap", action="store_true")
    ap.add_argument("--bc-cell-list", default="auto", choices=["auto", "on", "off"], help="BC cell list (second, dense launch over the Zou-He / outflow cells): stepper default, forced on, forced off")
    ap.add_argument("--no-balance", action="store_true", help="N>1, obstacle workloads: equal plane counts per rank instead of equal cost")
    ap.add_argument("--halo", default="auto", choices=["auto", "peer", "nccl"], help="N>1: halo exchange fused into the edge launch over peer memory, or NCCL send/recv")
    ap.add_argument("--no-dist-graph", action="store_true", help="N>1: issue the distributed step eagerly instead of replaying its CUDA graph")
    ap.add_argument("--no-configs", action="store_true", help="skip the extra BASELINE configs block (N=1) / the bitwise check and the strong-scaling line (N>1)")
    ap.add_argument("--variant", type=int, default=0, help="kernel variant of xlb_step (0 auto, 2 vector, 3 generic, 4 bulk-staged)")
    ap.add_argument("--radius", type=int, default=None, help="override the obstacle radius of the sphere / wind-tunnel workloads (experiments only)")
    ap.add_argument("--shape", default=None, help="override the workload's shape, e.g. 48,1536,1536 (experiments only; config.shape reports it)")
    ap.add_argument("--graph", type=int, default=0, help="replay the loop as CUDA graphs of this many (even) steps each (GraphedSteps); 0 = eager launches")
    ap.add_argument("--strong", action="store_true", help="N>1: keep the global shape of the workload (strong scaling) instead of one block per GPU")
    args = ap.parse_args()
    if args.workload is None:
        # BASELINE configs[1] (512^3 D3Q19 BGK FP32FP32) per GPU; N>1 stacks N such blocks along x (weak scaling).
        # BASELINE configs[3] is `--workload cavity1024_d3q19_bgk_fp16 --strong`.
        args.workload = "cavity512_d3q19_bgk_fp32"
    # the BC registry mirrors the reference's "registered bc ..." prints; keep stdout to the ONE JSON line
    sys.stdout = sys.stderr
    if args.impl == "reference":
        return run_reference(args)
    return run_native(args)


if __name__ == "__main__":
    sys.exit(main())
