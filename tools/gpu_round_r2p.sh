#!/bin/bash
# r2p: full GPU suite + default bench + in-place benches + ncu of the FP16 kernel after the phased load order
tag=${1:-r2p}
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
b --workload cavity768_d3q19_bgk_fp16 --steps 20 --warmup 6
b --workload cavity1024_d3q19_bgk_fp16 --steps 10 --warmup 4
b --workload cavity512_d3q19_bgk_fp32 --steps 40 --warmup 6 --in-place
b --workload cavity384x256x512_d3q27_bgk_fp32 --steps 40 --warmup 6 --in-place
b --workload cavity768x256x256_d3q27_kbc_fp32 --steps 40 --warmup 6 --in-place
b --workload cavity1024_d3q19_bgk_fp32 --steps 10 --warmup 4 --in-place
python - <<PY
import json
for l in open('$out/${tag}_bench_more.jsonl'):
    d=json.loads(l); print(d['config']['workload'], d['config']['shape'], d['config'].get('buffers','two buffers')[:22], round(d['value']), round(d['ms_per_step'],3), round(d['roofline']['frac'],3))
PY
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_launches_default_bench.csv \
  python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-configs > $out/${tag}_ncu1.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:step_vector_h2 -s 4 -c 1 -f -o $out/${tag}_ncu_f16 \
  python bench.py --workload cavity768_d3q19_bgk_fp16 --steps 4 --warmup 3 --no-e2e --no-cpu-baseline --no-configs > $out/${tag}_ncu2.log 2>&1
ls -la $out/${tag}_ncu_f16.ncu-rep
