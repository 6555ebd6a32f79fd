#!/usr/bin/env python
"""Print the handful of ncu metrics the profiles/ summaries quote from a .ncu-rep (reads it with `ncu -i ... --page raw --csv`).

    python tools/ncu_summary.py gpurun_out/x.ncu-rep
"""
import csv
import io
import subprocess
import sys

KEYS = (
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__inst_executed.sum", "smsp__inst_executed_op_local_ld.sum", "smsp__inst_executed_op_local_st.sum",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
    "sm__inst_executed_pipe_fma.sum", "sm__inst_executed_pipe_alu.sum", "sm__inst_executed_pipe_lsu.sum", "sm__inst_executed_pipe_fmaheavy.sum",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
)


def main(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    head, units = rows[0], rows[1]
    for vals in rows[2:]:
        rec = dict(zip(head, vals))
        print("==", rec.get("Kernel Name", "?")[:100])
        for k, u in zip(head, units):
            v = rec[k]
            stall = "issue_stalled" in k and k.endswith("_per_warp_active.pct")
            if k in KEYS or stall:
                try:
                    if stall and float(v.replace(",", "")) < 2.0:
                        continue
                except ValueError:
                    pass
                print(f"  {k:90s} {v} {u}")


if __name__ == "__main__":
    main(sys.argv[1])
