#!/bin/bash
# Single-GPU run after the generic-engine trim: the whole GPU test suite, the workloads that use the per-cell engine,
# ncu on the list kernel, the default bench line and the ncu launch list of the default bench command.
#   gpurun --timeout 1200 -- 'bash tools/gpu_round_r1w.sh r1w'
tag=${1:-r1w}
out=gpurun_out
mkdir -p $out
b() { timeout 600 python bench.py --no-e2e --no-cpu-baseline "$@" 2>> $out/${tag}_bench.err | grep '^{' >> $out/${tag}_bench.jsonl; }
timeout 1200 python -m pytest tests -m gpu -q > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest=$?"; tail -n 6 $out/${tag}_pytest_gpu.log
b --workload sphere384_d3q27_kbc_fp32 --steps 200 --warmup 10
b --workload sphere768_d3q27_kbc_fp32 --steps 100 --warmup 5
b --workload windtunnel192x1536x1536_d3q27_kbc_fp32 --steps 30 --warmup 3
b --workload windtunnel192x1536x1536_d3q27_kbc_fp32 --steps 30 --warmup 3 --variant 3 --shape 24,1536,1536
python - <<PY
import json
for l in open('$out/${tag}_bench.jsonl'):
    d=json.loads(l); print(d['config']['workload'], d['config']['shape'], d['roofline']['kernel'], round(d['value']), round(d['ms_per_step'],3), round(d['roofline']['frac'],3), d['gpu_launches'])
PY
timeout 600 python bench.py > $out/${tag}_bench_default.json 2> $out/${tag}_bench_default.err; echo "default=$?"; cut -c1-160 $out/${tag}_bench_default.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:step_ -c 40 --csv --log-file $out/${tag}_launches_default_bench.csv \
  python bench.py --steps 20 --warmup 3 --no-e2e --no-cpu-baseline > $out/${tag}_launches_default_bench.log 2>&1; echo "launches=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:step_cells -s 3 -c 1 -o $out/${tag}_ncu_cells -f \
  python bench.py --no-e2e --no-cpu-baseline --workload windtunnel192x1536x1536_d3q27_kbc_fp32 --shape 48,1536,1536 --steps 2 --warmup 2 > $out/${tag}_ncu_cells.log 2>&1; echo "ncu_cells=$?"
ls -la $out | tail -n 10
