#!/bin/bash
# r3b: the round's closing run (bulk pipeline without the backwards walk): full GPU suite, default bench, reference arm, launch list, smoke
tag=${1:-r3b}
out=gpurun_out
mkdir -p $out
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv,noheader > $out/${tag}_gpu.txt
timeout 1500 python -m pytest tests -m gpu -q -x > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest=$?"; tail -n 6 $out/${tag}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $out/${tag}_smoke.log 2>&1; echo "smoke=$?"; tail -n 2 $out/${tag}_smoke.log
timeout 900 python bench.py > $out/${tag}_bench_default.json 2> $out/${tag}_bench_default.err; echo "bench=$?"
python - <<PY
import json
d=json.load(open('$out/${tag}_bench_default.json'))
print(round(d['value']), d['steps'], round(d['ms_per_step'],4), d['roofline']['frac'], round(d['roofline']['kernel_ms'],4), 'e2e', round(d['e2e']['value']), d['e2e']['ms_per_step'], d['clocks'])
for k,v in d['configs'].items():
    print(k, {kk:(round(vv,4) if isinstance(vv,float) else vv) for kk,vv in v.items() if kk in ('mlups','kernel_ms','frac','launches_per_step','error','skipped')})
PY
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $out/${tag}_bench_reference.json 2> $out/${tag}_bench_reference.err; echo "ref=$?"; cut -c1-400 $out/${tag}_bench_reference.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_launches_default_bench.csv \
  python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-configs > $out/${tag}_ncu1.log 2>&1
