#!/bin/bash
# Single-GPU confirmation run: the whole GPU test suite, FP16 / wind-tunnel / sphere bench lines with the final defaults,
# the default bench line, and the ncu launch list of the default bench command.
#   gpurun --timeout 1500 -- 'bash tools/gpu_round_r1u.sh r1u'
tag=${1:-r1u}
out=gpurun_out
mkdir -p $out
b() { timeout 600 python bench.py --no-e2e --no-cpu-baseline "$@" 2>> $out/${tag}_bench.err | grep '^{' >> $out/${tag}_bench.jsonl; }
timeout 1200 python -m pytest tests -m gpu -q > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest=$?"; tail -n 8 $out/${tag}_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $out/${tag}_smoke.log 2>&1; echo "smoke=$?"; tail -n 2 $out/${tag}_smoke.log
b --workload cavity768_d3q19_bgk_fp16 --steps 50 --warmup 5
b --workload cavity1024_d3q19_bgk_fp16 --steps 30 --warmup 5
b --workload cavity1024_d3q19_bgk_fp16 --steps 30 --warmup 5 --variant 4
b --workload sphere384_d3q27_kbc_fp32 --steps 200 --warmup 10
b --workload sphere768_d3q27_kbc_fp32 --steps 100 --warmup 5
b --workload windtunnel192x1536x1536_d3q27_kbc_fp32 --steps 30 --warmup 3
python - <<PY
import json
for l in open('$out/${tag}_bench.jsonl'):
    d=json.loads(l); print(d['config']['workload'], d['config']['shape'], d['roofline']['kernel'], round(d['value']), round(d['ms_per_step'],3), round(d['roofline']['frac'],3), d['gpu_launches'])
PY
timeout 600 python bench.py > $out/${tag}_bench_default.json 2> $out/${tag}_bench_default.err; echo "default=$?"; cut -c1-160 $out/${tag}_bench_default.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $out/${tag}_bench_reference.json 2> $out/${tag}_bench_reference.err; echo "reference=$?"; cut -c1-200 $out/${tag}_bench_reference.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:xlb -c 60 --csv --log-file $out/${tag}_launches_default_bench.csv \
  python bench.py --steps 20 --warmup 3 --no-e2e --no-cpu-baseline > $out/${tag}_launches_default_bench.log 2>&1; echo "launches=$?"
ls -la $out | tail -n 12
