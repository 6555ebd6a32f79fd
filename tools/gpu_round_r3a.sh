#!/bin/bash
# r3a: backwards plane order on odd steps in the BULK pipeline only (K1 / K1c / K1d keep it): configs[2] and the wide lattices
tag=${1:-r3a}
out=gpurun_out
mkdir -p $out
{
for rep in 1 2; do
echo "== sphere384_d3q27_kbc_fp32 rep $rep"; timeout 600 python tools/sweep.py run sphere384_d3q27_kbc_fp32 --steps 200 --only q27_f32f32__BULK_REVERSE
done
echo "== sphere384 --graph 20"; XLB_SWEEP_BENCH_ARGS="--graph 20" timeout 600 python tools/sweep.py run sphere384_d3q27_kbc_fp32 --steps 200 --only q27_f32f32__BULK_REVERSE
echo "== sphere768_d3q27_kbc_fp32"; timeout 600 python tools/sweep.py run sphere768_d3q27_kbc_fp32 --steps 40 --only q27_f32f32__BULK_REVERSE
echo "== cavity384x256x512_d3q27_bgk_fp32"; timeout 600 python tools/sweep.py run cavity384x256x512_d3q27_bgk_fp32 --steps 40 --only q27_f32f32__BULK_REVERSE
} > $out/${tag}_sweep_bulk_reverse.log 2>&1
cat $out/${tag}_sweep_bulk_reverse.log
