#!/bin/bash
# Single-GPU evidence run after the BC-cell-list change: GPU tests, bench lines, shape experiments for the wide-row
# D3Q27 KBC case, and ncu --set full captures of the FP16-storage and wind-tunnel kernels.
#   gpurun --timeout 1500 -- 'bash tools/gpu_round_r1q.sh r1q'
tag=${1:-r1q}
out=gpurun_out
mkdir -p $out
b() { timeout 600 python bench.py --no-e2e --no-cpu-baseline "$@" 2>> $out/${tag}_bench.err | grep '^{' >> $out/${tag}_bench.jsonl; }
timeout 900 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest=$?"; tail -n 3 $out/${tag}_pytest_gpu.log
b --steps 100 --warmup 5
b --workload sphere384_d3q27_kbc_fp32 --steps 200 --warmup 10
b --workload sphere768_d3q27_kbc_fp32 --steps 100 --warmup 5
b --workload cavity768_d3q19_bgk_fp16 --steps 50 --warmup 5
b --workload windtunnel192x1536x1536_d3q27_kbc_fp32 --steps 30 --warmup 3
b --workload windtunnel192x1536x1536_d3q27_kbc_fp32 --steps 30 --warmup 3 --variant 2
b --workload windtunnel192x1536x1536_d3q27_kbc_fp32 --steps 30 --warmup 3 --shape 1728,512,512
b --workload windtunnel192x1536x1536_d3q27_kbc_fp32 --steps 30 --warmup 3 --shape 768,768,768
b --workload windtunnel192x1536x1536_d3q27_kbc_fp32 --steps 30 --warmup 3 --shape 48,1536,1536
cut -c1-420 $out/${tag}_bench.jsonl
# ncu (numbers printed under ncu are not bench values)
timeout 600 ncu --set full --clock-control none --import-source on -k regex:step_vector -s 3 -c 1 -o $out/${tag}_ncu_f16 -f \
  python bench.py --no-e2e --no-cpu-baseline --workload cavity768_d3q19_bgk_fp16 --steps 2 --warmup 2 > $out/${tag}_ncu_f16.log 2>&1; echo "ncu_f16=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:step_bulk -s 3 -c 1 -o $out/${tag}_ncu_wt -f \
  python bench.py --no-e2e --no-cpu-baseline --workload windtunnel192x1536x1536_d3q27_kbc_fp32 --shape 48,1536,1536 --steps 2 --warmup 2 > $out/${tag}_ncu_wt.log 2>&1; echo "ncu_wt=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:step_cells -s 3 -c 1 -o $out/${tag}_ncu_cells -f \
  python bench.py --no-e2e --no-cpu-baseline --workload windtunnel192x1536x1536_d3q27_kbc_fp32 --shape 48,1536,1536 --steps 2 --warmup 2 > $out/${tag}_ncu_cells.log 2>&1; echo "ncu_cells=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $out/${tag}_launches_wt.csv \
  python bench.py --no-e2e --no-cpu-baseline --workload windtunnel192x1536x1536_d3q27_kbc_fp32 --shape 48,1536,1536 --steps 4 --warmup 3 > $out/${tag}_launches_wt.log 2>&1; echo "launches=$?"
ls -la $out | tail -n 20
