#!/bin/bash
tag=${1:-r2g}
out=gpurun_out
mkdir -p $out
timeout 1200 python -m pytest tests -m gpu -q -x > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest=$?"; tail -n 4 $out/${tag}_pytest_gpu.log
timeout 900 python bench.py --steps 20 --warmup 5 > $out/${tag}_bench_default.json 2> $out/${tag}_bench_default.err; echo "bench=$?"
python - <<PY
import json
d=json.load(open('$out/${tag}_bench_default.json'))
print(round(d['value']), d['roofline']['frac'], 'e2e', round(d['e2e']['value']), d['e2e']['ms_per_step'])
for k,v in d['configs'].items():
    print(k, {kk:(round(vv,4) if isinstance(vv,float) else vv) for kk,vv in v.items() if kk in ('mlups','kernel_ms','frac','launches_per_step','error','skipped')})
PY
timeout 300 ncu --set full --clock-control none --import-source on -k regex:step_bulk -s 6 -c 1 -o $out/${tag}_ncu_c3_sphere384 -f python bench.py --workload sphere384_d3q27_kbc_fp32 --steps 4 --warmup 3 --no-e2e --no-cpu-baseline --no-configs > /dev/null 2>&1; echo "ncu_c3=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:step_vector_h2 -s 6 -c 1 -o $out/${tag}_ncu_f16 -f python bench.py --workload cavity768_d3q19_bgk_fp16 --steps 4 --warmup 3 --no-e2e --no-cpu-baseline --no-configs > /dev/null 2>&1; echo "ncu_f16=$?"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $out/${tag}_launches_default_bench.csv python bench.py --steps 4 --warmup 3 --no-e2e --no-cpu-baseline --no-configs > /dev/null 2>&1; echo "ncu_list=$?"
