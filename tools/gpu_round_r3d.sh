#!/bin/bash
# r3d: the binary that ships (header comment added after r3b; same code): GPU suite, smoke, default bench
tag=${1:-r3d}
out=gpurun_out
mkdir -p $out
timeout 900 python -m pytest tests -m gpu -q -x > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest=$?"; tail -n 2 $out/${tag}_pytest_gpu.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $out/${tag}_smoke.log 2>&1; echo "smoke=$?"
timeout 600 python bench.py > $out/${tag}_bench_default.json 2> $out/${tag}_bench_default.err; echo "bench=$?"
python - <<PY
import json
d=json.load(open('$out/${tag}_bench_default.json'))
print(round(d['value']), d['steps'], round(d['ms_per_step'],4), d['roofline']['frac'], 'e2e', round(d['e2e']['value']), d['clocks'])
for k,v in d['configs'].items():
    print(k, round(v.get('mlups',0)), round(v.get('frac',0),3))
PY
