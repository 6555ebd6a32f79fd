"""Host-side pieces of bench.py that need no GPU: the algorithmic-bytes table of the roofline, the clock sampler, and the
reference arm's behaviour on the ranks that do not run it."""

import os
import subprocess
import sys
import time

import bench

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_algorithmic_bytes_per_cell_update():
    # B_alg = 2 * q * s + 1 (q populations read and written once, one bc_mask byte)
    assert bench.b_alg(19, "FP32FP32") == 153 and bench.b_alg(19, "FP64FP64") == 305 and bench.b_alg(19, "FP32FP16") == 77
    assert bench.b_alg(27, "FP32FP32") == 217 and bench.b_alg(27, "FP32FP16") == 109 and bench.b_alg(9, "FP32FP32") == 73


def test_clock_sampler_starts_before_the_timed_region_and_survives_a_missing_nvidia_smi():
    s = bench.ClockSampler(0)
    t0 = time.time()
    s.start()  # returns once the first query process exists (or could not be created): nothing is forked inside the timed region
    assert time.time() - t0 < 5.0
    out = s.stop()
    assert set(out) == {"sm_mhz", "sm_max_mhz", "reasons", "samples"} and isinstance(out["reasons"], list)


def test_reference_arm_runs_on_rank_zero_only():
    env = dict(os.environ, RANK="1", LOCAL_RANK="1", WORLD_SIZE="2")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0"],
                       env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and not [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
