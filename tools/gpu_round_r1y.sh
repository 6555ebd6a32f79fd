#!/bin/bash
# Final single-GPU confirmation: whole GPU suite + smoke with the committed library, BC-cell-list threshold check on the small sphere.
tag=${1:-r1y}
out=gpurun_out
mkdir -p $out
b() { timeout 600 python bench.py --no-e2e --no-cpu-baseline "$@" 2>> $out/${tag}_bench.err | grep '^{' >> $out/${tag}_bench.jsonl; }
timeout 1200 python -m pytest tests -m gpu -q > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest=$?"; tail -n 6 $out/${tag}_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $out/${tag}_smoke.log 2>&1; echo "smoke=$?"; tail -n 2 $out/${tag}_smoke.log
b --workload sphere384_d3q27_kbc_fp32 --steps 300 --warmup 10
XLB_B200_BC_CELL_LIST=1 b --workload sphere384_d3q27_kbc_fp32 --steps 300 --warmup 10
XLB_B200_BC_CELL_LIST=0 b --workload sphere768_d3q27_kbc_fp32 --steps 100 --warmup 5
b --workload sphere768_d3q27_kbc_fp32 --steps 100 --warmup 5
b --workload cavity768_d3q19_bgk_fp16 --steps 50 --warmup 5
python - <<PY
import json
for l in open('$out/${tag}_bench.jsonl'):
    d=json.loads(l); print(d['config']['workload'], d['config']['shape'], round(d['value']), round(d['ms_per_step'],4), round(d['roofline']['frac'],3), d['roofline']['launches_per_step'])
PY
