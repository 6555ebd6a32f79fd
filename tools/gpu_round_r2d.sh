#!/bin/bash
# 2-GPU call after the peer-stride fix and the dynamic tile scheduler: single-GPU suite on GPU 0 (scheduler), multi-GPU bitwise
# tests, the bench at N=2 (weak + strong + nrank_bitwise) and the single-GPU bench with the configs block.
tag=${1:-r2d}
n=${2:-2}
out=gpurun_out
mkdir -p $out
CUDA_VISIBLE_DEVICES=0 timeout 1200 python -m pytest tests -m gpu -q -x --deselect tests/test_gpu_multi.py > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest1=$?"; tail -n 6 $out/${tag}_pytest_gpu.log
bash tools/mgpu_round_r2.sh $tag $n
CUDA_VISIBLE_DEVICES=0 timeout 900 python bench.py --steps 20 --warmup 5 > $out/${tag}_bench_default.json 2> $out/${tag}_bench_default.err; echo "bench=$?"
python - <<PY
import json
d=json.load(open('$out/${tag}_bench_default.json'))
print(round(d['value']), d['roofline']['frac'])
for k,v in d['configs'].items():
    print(k, {kk:(round(vv,4) if isinstance(vv,float) else vv) for kk,vv in v.items() if kk in ('mlups','kernel_ms','frac','launches_per_step','error','skipped')})
PY
