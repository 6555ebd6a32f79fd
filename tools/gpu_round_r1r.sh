#!/bin/bash
# Single-GPU run after the packed-half2 FP16 kernel and the mesh masker: GPU tests, FP16 launch-bound sweep, bench lines,
# footprint experiments (same kernel, larger grids), ncu capture of the FP16 kernel.
#   gpurun --timeout 1500 -- 'bash tools/gpu_round_r1r.sh r1r'
tag=${1:-r1r}
out=gpurun_out
mkdir -p $out
b() { timeout 600 python bench.py --no-e2e --no-cpu-baseline "$@" 2>> $out/${tag}_bench.err | grep '^{' >> $out/${tag}_bench.jsonl; }
timeout 900 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest=$?"; tail -n 5 $out/${tag}_pytest_gpu.log
timeout 600 python tools/sweep.py run cavity768_d3q19_bgk_fp16 --steps 30 --variant 2 > $out/${tag}_sweep_f16_h2.log 2>&1; cat $out/${tag}_sweep_f16_h2.log
b --workload cavity768_d3q19_bgk_fp16 --steps 50 --warmup 5
b --workload cavity1024_d3q19_bgk_fp16 --steps 30 --warmup 5
b --workload cavity768_d3q19_bgk_fp16 --steps 100 --warmup 5 --shape 512,512,512
b --steps 100 --warmup 5
b --steps 50 --warmup 5 --shape 1024,512,512
b --steps 30 --warmup 5 --shape 768,768,768
b --workload sphere384_d3q27_kbc_fp32 --steps 200 --warmup 10
b --workload sphere768_d3q27_kbc_fp32 --steps 100 --warmup 5
b --workload windtunnel192x1536x1536_d3q27_kbc_fp32 --steps 30 --warmup 3
cut -c1-130 $out/${tag}_bench.jsonl; python - <<'PY'
import json
for l in open('gpurun_out/'+'%s'%"r1r"+'_bench.jsonl'):
    d=json.loads(l); print(d['config']['workload'], d['config']['shape'], round(d['value']), d['ms_per_step'], round(d['roofline']['frac'],3))
PY
timeout 600 ncu --set full --clock-control none --import-source on -k regex:step_vector -s 3 -c 1 -o $out/${tag}_ncu_f16 -f \
  python bench.py --no-e2e --no-cpu-baseline --workload cavity768_d3q19_bgk_fp16 --steps 2 --warmup 2 > $out/${tag}_ncu_f16.log 2>&1; echo "ncu_f16=$?"
ls -la $out | tail -n 12
