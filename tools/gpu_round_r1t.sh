#!/bin/bash
# Single-GPU run: GPU tests, FP16 sweeps (vector packed-half2 widened variants; bulk-staged packed-half2 consumer), bench lines
# of the default library, ncu captures of the two FP16 kernels.
#   gpurun --timeout 1500 -- 'bash tools/gpu_round_r1t.sh r1t'
tag=${1:-r1t}
out=gpurun_out
mkdir -p $out
b() { timeout 600 python bench.py --no-e2e --no-cpu-baseline "$@" 2>> $out/${tag}_bench.err | grep '^{' >> $out/${tag}_bench.jsonl; }
timeout 900 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest=$?"; tail -n 5 $out/${tag}_pytest_gpu.log
timeout 300 python tools/sweep.py run cavity768_d3q19_bgk_fp16 --steps 30 --variant 2 --only q19_f32f16__V > $out/${tag}_sweep_f16_vec.log 2>&1; cat $out/${tag}_sweep_f16_vec.log
timeout 400 python tools/sweep.py run cavity768_d3q19_bgk_fp16 --steps 30 --variant 4 --only q19_f32f16__HC --shape 512,768,1024 > $out/${tag}_sweep_f16_bulk_nz1024.log 2>&1; cat $out/${tag}_sweep_f16_bulk_nz1024.log
timeout 300 python tools/sweep.py run cavity768_d3q19_bgk_fp16 --steps 30 --variant 4 --only q19_f32f16__HC2 > $out/${tag}_sweep_f16_bulk_nz768.log 2>&1; cat $out/${tag}_sweep_f16_bulk_nz768.log
b --workload cavity768_d3q19_bgk_fp16 --steps 50 --warmup 5
b --workload cavity768_d3q19_bgk_fp16 --steps 50 --warmup 5 --variant 4
b --workload cavity768_d3q19_bgk_fp16 --steps 50 --warmup 5 --shape 512,768,1024
b --workload cavity768_d3q19_bgk_fp16 --steps 50 --warmup 5 --shape 512,768,1024 --variant 4
b --workload cavity1024_d3q19_bgk_fp16 --steps 30 --warmup 5
b --workload cavity1024_d3q19_bgk_fp16 --steps 30 --warmup 5 --variant 4
b --workload sphere384_d3q27_kbc_fp32 --steps 200 --warmup 10
b --workload windtunnel192x1536x1536_d3q27_kbc_fp32 --steps 30 --warmup 3
python - <<PY
import json
for l in open('$out/${tag}_bench.jsonl'):
    d=json.loads(l); print(d['config']['workload'], d['config']['shape'], d['roofline']['kernel'], round(d['value']), round(d['ms_per_step'],3), round(d['roofline']['frac'],3), d['gpu_launches'])
PY
timeout 600 python bench.py --steps 200 --warmup 5 > $out/${tag}_bench_default.json 2> $out/${tag}_bench_default.err; echo "default=$?"; cut -c1-200 $out/${tag}_bench_default.json
timeout 600 ncu --set full --clock-control none --import-source on -k regex:step_vector -s 3 -c 1 -o $out/${tag}_ncu_f16_vec -f \
  python bench.py --no-e2e --no-cpu-baseline --workload cavity768_d3q19_bgk_fp16 --shape 512,768,1024 --steps 2 --warmup 2 > $out/${tag}_ncu_f16_vec.log 2>&1; echo "ncu_vec=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:step_bulk -s 3 -c 1 -o $out/${tag}_ncu_f16_bulk -f \
  python bench.py --no-e2e --no-cpu-baseline --workload cavity768_d3q19_bgk_fp16 --shape 512,768,1024 --variant 4 --steps 2 --warmup 2 > $out/${tag}_ncu_f16_bulk.log 2>&1; echo "ncu_bulk=$?"
ls -la $out | tail -n 14
