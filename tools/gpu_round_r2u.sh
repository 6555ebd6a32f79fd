#!/bin/bash
# r2u: full GPU suite + default bench with the final library (loads-first K1c, in-place K1d, K1b claim order + multi-row tiles)
tag=${1:-r2u}
out=gpurun_out
mkdir -p $out
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv,noheader > $out/${tag}_gpu.txt
timeout 1500 python -m pytest tests -m gpu -q -x > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest=$?"; tail -n 6 $out/${tag}_pytest_gpu.log
timeout 900 python bench.py --steps 20 --warmup 5 > $out/${tag}_bench_default.json 2> $out/${tag}_bench_default.err; echo "bench=$?"
python - <<PY
import json
d=json.load(open('$out/${tag}_bench_default.json'))
print(round(d['value']), d['roofline']['frac'], 'e2e', round(d['e2e']['value']), d['e2e']['ms_per_step'])
for k,v in d['configs'].items():
    print(k, {kk:(round(vv,4) if isinstance(vv,float) else vv) for kk,vv in v.items() if kk in ('mlups','kernel_ms','frac','launches_per_step','error','skipped')})
PY
b() { timeout 600 python bench.py --no-e2e --no-cpu-baseline --no-configs "$@" 2>> $out/${tag}_bench.err | grep '^{' >> $out/${tag}_bench_more.jsonl; }
b --workload cavity128_d3q19_bgk_fp32 --steps 400 --warmup 6 --variant 4
b --workload cavity128_d3q19_bgk_fp32 --steps 400 --warmup 6 --variant 4 --graph 20
b --workload cavity128_d3q19_bgk_fp32 --steps 400 --warmup 6
b --workload cavity512_d3q19_bgk_fp32 --steps 40 --warmup 6 --variant 4
b --workload cavity384x256x512_d3q27_bgk_fp32 --steps 40 --warmup 6 --variant 4
python - <<PY
import json
for l in open('$out/${tag}_bench_more.jsonl'):
    d=json.loads(l); print(d['config']['workload'], d['config']['shape'], d['config'].get('graph_steps',''), round(d['value']), round(d['ms_per_step'],4), round(d['roofline']['frac'],3))
PY
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_launches_default_bench.csv \
  python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-configs > $out/${tag}_ncu1.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:step_bulk -s 4 -c 1 -f -o $out/${tag}_ncu_c3 \
  python bench.py --workload sphere384_d3q27_kbc_fp32 --steps 4 --warmup 3 --no-e2e --no-cpu-baseline --no-configs > $out/${tag}_ncu2.log 2>&1
ls -la $out/${tag}_ncu_c3.ncu-rep
