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
    "cavity1024_d3q19_bgk_fp16": ("D3Q19", "BGK", "FP32FP16", (1024, 1024, 1024), 1.0, "cavity"),
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
        self._t = None

    def _run(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(
                    ["nvidia-smi", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits", "-i", str(self.index)],
                    capture_output=True, text=True, timeout=5,
                ).stdout.strip()
                if out:
                    self.rows.append([c.strip() for c in out.split(",")])
            except Exception:
                pass
            self._stop.wait(0.2)

    def start(self):
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()

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
def build_case(workload, slab=None, shape=None, variant=0, radius=None):
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
        grid = xlb.grid.grid_factory(shape, compute_backend=ComputeBackend.CUDA)
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
    stepper = IncompressibleNavierStokesStepper(grid=grid, boundary_conditions=bcs, collision_type=collision, backend_config={"kernel_variant": variant})
    return stepper, vs, pol, shape, omega


# --------------------------------------------------------------------------- CPU legs (the only users of oracle/)
def cpu_baseline(workload, budget_s=20.0, max_steps=8):
    """The oracle port of the reference's JAX step (nse_stepper.py:238-282), timed on this box's host cores on a bounded
    sample: the same BCs / lattice / policy at 128^3 (BASELINE config 1 when the workload is the cavity).  The cavity
    runs the multi-threaded torch-CPU transcription of the oracle (all host threads, like XLA:CPU would use); the
    other workloads run the single-threaded NumPy oracle."""
    import numpy as np
    import torch

    from oracle import cases

    lattice, collision, policy, shape, omega, kind = WORKLOADS[workload]
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
        "sample": f"{steps} steps of {case['name']} ({ncell} cells), {how}; oracle port of nse_stepper.py:238-282; host has {os.cpu_count()} cores",
        "seconds": dt,
        "steps": steps,
    }


def run_reference(args):
    """--impl reference: the reference's CPU path (oracle port; the reference itself needs jax, absent here)."""
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
        "config": {"workload": workload, "lattice": lattice, "collision": collision, "policy": policy, "shape": list(shape), "omega": omega},
        "cpu_baseline": {k: base[k] for k in ("value", "unit", "cores", "kind", "sample")},
        "e2e": {"value": base["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    _STDOUT.write(json.dumps(line) + "\n")
    _STDOUT.flush()
    return 0


# --------------------------------------------------------------------------- native arm
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

    workload = args.workload
    lattice, collision, policy, shape, omega, kind = WORKLOADS[workload]
    if args.shape:
        shape = tuple(int(v) for v in args.shape.split(","))
    if world > 1 and not args.strong:
        shape = (shape[0] * world,) + tuple(shape[1:])  # weak scaling: the N=1 block per GPU, stacked along x
    stepper, vs, pol, shape, omega = build_case(workload, slab=slab, shape=shape, variant=args.variant, radius=args.radius)
    if slab is not None:
        from xlb_b200.distribute import distribute

        stepper = distribute(stepper, stepper.grid, vs, overlap=not args.no_overlap)
    f_0, f_1, bc_mask, missing = stepper.prepare_fields()
    ncell = shape[0] * shape[1] * shape[2]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- warm-up
    t = 0
    for _ in range(max(args.warmup, 3)):
        f_0, f_1 = stepper(f_0, f_1, bc_mask, missing, omega, t)
        f_0, f_1 = f_1, f_0
        t += 1
    barrier()

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
    ms = ev0.elapsed_time(ev1)
    launches = _lib.launch_count() - launches0
    if graphed is not None:
        launches = int(round(graph_lps * args.steps))  # kernels replayed from the graphs (no host-side launch calls)
    if world > 1:
        tt = torch.tensor([ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms = float(tt.item())
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
        roofline = {
            "bound": "hbm",
            "kernel": {0: "xlb_step (auto variant)", 2: "step_vector_kernel", 3: "step_generic_kernel", 4: "step_bulk_kernel"}.get(args.variant, "xlb_step"),
            "achieved": achieved,
            "peak": peak,
            "unit": "GB/s",
            "frac": achieved / peak,
            "traffic": ncu_traffic(workload, args.variant),
            "frac_of_nominal_8000": achieved / 8000.0,
            "bytes_per_launch": bytes_per_launch,
            "bytes_per_cell": b_alg(vs.q, policy),
            "cells_per_launch": ncell,
            "kernel_ms": k_ms,
            "launches_per_step": launches / max(1, args.steps),  # 2 = fused kernel + dense launch over the BC cell list
            "peak_source": peak_src,
        }

    if world > 1:
        # per-GPU roofline of the whole distributed step (two edge-plane launches + interior launch + halo ring):
        # algorithmic bytes of one rank's slab over the max-over-ranks step time
        per_rank = b_alg(vs.q, policy) * (ncell // world)
        achieved = per_rank / (ms / args.steps * 1e-3) / 1e9
        roofline = {
            "bound": "hbm", "kernel": "per-GPU slab step (2 edge-plane launches + interior launch, halo exchange overlapped)",
            "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
            "frac_of_nominal_8000": achieved / 8000.0, "bytes_per_launch": per_rank, "bytes_per_cell": b_alg(vs.q, policy),
            "cells_per_launch": ncell // world, "kernel_ms": ms / args.steps, "peak_source": peak_src,
        }

    # ---- e2e: HOST buffers in, HOST buffers out (pinned), copies inside the timed region
    e2e = None
    if world == 1 and not args.no_e2e:
        e2e = run_e2e(stepper, f_0, f_1, bc_mask, missing, omega, t, ncell, max(1, min(args.steps, args.e2e_steps)))

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
        "warmup": max(args.warmup, 3),
        "ms_per_step": ms / args.steps,
        "higher_is_better": True,
        "scaling": "strong" if (world > 1 and args.strong) else "weak",
        "vs_baseline": None,
        "dtype": DTYPE_NAME[policy],
        "data": "synthetic",
        "config": {
            "workload": workload, "lattice": lattice, "collision": collision, "policy": policy, "shape": list(shape), "omega": omega,
            "decomposition": f"x-slab x{world}" if world > 1 else "single block",
            **({"graph_steps": args.graph} if graphed is not None else {}),
            "l2": "inputs larger than L2 (two population buffers of %.1f GB each)" % (vs.q * ncell * STORE_BYTES[policy] / world / 1e9),
        },
        "gpu_launches": int(launches),
        "finite": finite,
        "clocks": clocks,
    }
    if roofline:
        line["roofline"] = roofline
    if e2e:
        line["e2e"] = e2e
        line["e2e_resident_api"] = run_resident_api(stepper, f_0, f_1, bc_mask, missing, omega, t, ncell, min(args.steps, 20))
    if world == 1 and not args.no_cpu_baseline:
        base = cpu_baseline(workload)
        line["cpu_baseline"] = {k: base[k] for k in ("value", "unit", "cores", "kind", "sample")}
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

    h_in = torch.empty(f_0.shape, dtype=f_0.dtype, pin_memory=True)
    h_out = torch.empty(f_1.shape, dtype=f_1.dtype, pin_memory=True)
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
    ap.add_argument("--e2e-steps", type=int, default=4)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-overlap", action="store_true")
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
