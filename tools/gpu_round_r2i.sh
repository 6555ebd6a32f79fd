#!/bin/bash
tag=${1:-r2i}
out=gpurun_out
mkdir -p $out
timeout 1200 python -m pytest tests -m gpu -q -x > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest=$?"; tail -n 4 $out/${tag}_pytest_gpu.log
for v in 0 2; do timeout 400 python bench.py --workload windtunnel192x1536x1536_d3q27_kbc_fp32 --radius 192 --variant $v --steps 10 --warmup 3 --no-e2e --no-cpu-baseline --no-configs 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('windtunnel block, r=192 sphere, variant',$v, round(d['value']), round(d['roofline']['frac'],3), d['roofline']['launches_per_step'])" | tee -a $out/${tag}_bigsphere.log; done
timeout 900 python bench.py --steps 20 --warmup 5 > $out/${tag}_bench_default.json 2> $out/${tag}_bench_default.err; echo "bench=$?"
python - <<PY
import json
d=json.load(open('$out/${tag}_bench_default.json'))
print(round(d['value']), d['roofline']['frac'], 'e2e', round(d['e2e']['value']), d['e2e']['ms_per_step'], d['e2e'].get('host_memory'))
for k,v in d['configs'].items():
    print(k, {kk:(round(vv,4) if isinstance(vv,float) else vv) for kk,vv in v.items() if kk in ('mlups','kernel_ms','frac','launches_per_step','error','skipped')})
PY
