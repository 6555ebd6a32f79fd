#!/bin/bash
tag=${1:-r2m}
out=gpurun_out
mkdir -p $out
timeout 1500 python -m pytest tests -m gpu -q -x > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest=$?"; tail -n 6 $out/${tag}_pytest_gpu.log
b() { timeout 600 python bench.py --no-e2e --no-cpu-baseline --no-configs "$@" 2>> $out/${tag}_bench.err | grep '^{' >> $out/${tag}_bench_inplace.jsonl; }
b --workload cavity512_d3q19_bgk_fp32 --steps 40 --warmup 6
b --workload cavity512_d3q19_bgk_fp32 --steps 40 --warmup 6 --in-place
b --workload cavity384x256x512_d3q27_bgk_fp32 --steps 40 --warmup 6 --in-place
b --workload cavity768x256x256_d3q27_kbc_fp32 --steps 40 --warmup 6 --in-place
b --workload cavity512_d3q19_bgk_fp64 --steps 20 --warmup 6 --in-place
b --workload cavity1024_d3q19_bgk_fp32 --steps 10 --warmup 4 --in-place
python - <<PY
import json
for l in open('$out/${tag}_bench_inplace.jsonl'):
    d=json.loads(l); print(d['config']['workload'], d['config']['shape'], d['config'].get('buffers','two buffers')[:22], round(d['value']), round(d['ms_per_step'],3), round(d['roofline']['frac'],3))
PY
timeout 900 python bench.py --steps 20 --warmup 5 > $out/${tag}_bench_default.json 2> $out/${tag}_bench_default.err; echo "bench=$?"
python - <<PY
import json
d=json.load(open('$out/${tag}_bench_default.json'))
print(round(d['value']), d['roofline']['frac'], 'e2e', round(d['e2e']['value']), d['e2e']['ms_per_step'])
for k,v in d['configs'].items():
    print(k, {kk:(round(vv,4) if isinstance(vv,float) else vv) for kk,vv in v.items() if kk in ('mlups','kernel_ms','frac','launches_per_step','error','skipped')})
PY
