#!/bin/bash
# r2r: configs[2] (384x128x128 D3Q27 KBC sphere): tile claim order (last plane first) and BC cell list on / off
tag=${1:-r2r}
out=gpurun_out
mkdir -p $out
{
for rep in 1 2; do
echo "== sphere384_d3q27_kbc_fp32 (q27_f32f32) rep $rep"; timeout 600 python tools/sweep.py run sphere384_d3q27_kbc_fp32 --steps 200 --only q27_f32f32__BULK
done
echo "== sphere384 with the BC cell list forced on"; XLB_SWEEP_BENCH_ARGS="--bc-cell-list on" timeout 600 python tools/sweep.py run sphere384_d3q27_kbc_fp32 --steps 200 --only q27_f32f32__BULK
echo "== sphere768_d3q27_kbc_fp32"; timeout 600 python tools/sweep.py run sphere768_d3q27_kbc_fp32 --steps 40 --only q27_f32f32__BULK
echo "== sphere768 with the BC cell list forced off"; XLB_SWEEP_BENCH_ARGS="--bc-cell-list off" timeout 600 python tools/sweep.py run sphere768_d3q27_kbc_fp32 --steps 40 --only q27_f32f32__BULK
} > $out/${tag}_sweep_c3.log 2>&1
cat $out/${tag}_sweep_c3.log
