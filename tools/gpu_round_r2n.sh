#!/bin/bash
# r2n: in-place (AA) kernels after the slot-order / recompute / packed-BC-state rework: launch-bound sweep, tests, ncu capture
tag=${1:-r2n}
out=gpurun_out
mkdir -p $out
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv,noheader > $out/${tag}_gpu.txt
XLB_SWEEP_BENCH_ARGS="--in-place" timeout 900 python tools/sweep.py run cavity512_d3q19_bgk_fp32 --steps 40 --only q19_f32f32__AA > $out/${tag}_sweep_aa.log 2>&1
cat $out/${tag}_sweep_aa.log
timeout 900 python -m pytest tests/test_gpu_aa.py -q -x > $out/${tag}_pytest_aa.log 2>&1; echo "pytest_aa=$?"; tail -n 4 $out/${tag}_pytest_aa.log
b() { timeout 600 python bench.py --no-e2e --no-cpu-baseline --no-configs "$@" 2>> $out/${tag}_bench.err | grep '^{' >> $out/${tag}_bench_inplace.jsonl; }
b --workload cavity512_d3q19_bgk_fp32 --steps 40 --warmup 6
b --workload cavity512_d3q19_bgk_fp32 --steps 40 --warmup 6 --in-place
b --workload cavity384x256x512_d3q27_bgk_fp32 --steps 40 --warmup 6 --in-place
b --workload cavity768x256x256_d3q27_kbc_fp32 --steps 40 --warmup 6 --in-place
b --workload cavity512_d3q19_bgk_fp64 --steps 20 --warmup 6 --in-place
b --workload cavity768_d3q19_bgk_fp16 --steps 20 --warmup 6 --in-place
python - <<PY
import json
for l in open('$out/${tag}_bench_inplace.jsonl'):
    d=json.loads(l); print(d['config']['workload'], d['config']['shape'], d['config'].get('buffers','two buffers')[:22], round(d['value']), round(d['ms_per_step'],3), round(d['roofline']['frac'],3))
PY
# per-launch durations of P and L steps, then one full capture of each
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:step_aa -c 12 --csv --log-file $out/${tag}_launches_aa.csv \
  python bench.py --workload cavity512_d3q19_bgk_fp32 --in-place --steps 4 --warmup 3 --no-e2e --no-cpu-baseline --no-configs > $out/${tag}_ncu1.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:step_aa -s 4 -c 2 -f -o $out/${tag}_aa_full \
  python bench.py --workload cavity512_d3q19_bgk_fp32 --in-place --steps 4 --warmup 3 --no-e2e --no-cpu-baseline --no-configs > $out/${tag}_ncu2.log 2>&1
ls -la $out/${tag}_aa_full.ncu-rep
