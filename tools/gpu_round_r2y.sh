#!/bin/bash
# r2y: planes walked backwards on odd time steps (what the previous step wrote last is still in L2)
tag=${1:-r2y}
out=gpurun_out
mkdir -p $out
{
for rep in 1 2; do
echo "== cavity128_d3q19_bgk_fp32 rep $rep"; timeout 600 python tools/sweep.py run cavity128_d3q19_bgk_fp32 --steps 400 --only q19_f32f32__REVERSE
done
echo "== cavity128 --graph 20"; XLB_SWEEP_BENCH_ARGS="--graph 20" timeout 600 python tools/sweep.py run cavity128_d3q19_bgk_fp32 --steps 400 --only q19_f32f32__REVERSE
echo "== cavity128 shape 256,256,256"; timeout 600 python tools/sweep.py run cavity128_d3q19_bgk_fp32 --steps 100 --only q19_f32f32__REVERSE --shape 256,256,256
echo "== cavity512_d3q19_bgk_fp32"; timeout 600 python tools/sweep.py run cavity512_d3q19_bgk_fp32 --steps 40 --only q19_f32f32__REVERSE
echo "== cavity128 in place"; XLB_SWEEP_BENCH_ARGS="--in-place" timeout 600 python tools/sweep.py run cavity128_d3q19_bgk_fp32 --steps 400 --only q19_f32f32__REVERSE
} > $out/${tag}_sweep_reverse.log 2>&1
cat $out/${tag}_sweep_reverse.log
