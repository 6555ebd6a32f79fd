#!/bin/bash
# r2z: L2 eviction policy on the population stores (evict_last) and loads (evict_first)
tag=${1:-r2z}
out=gpurun_out
mkdir -p $out
{
for rep in 1; do
echo "== cavity128_d3q19_bgk_fp32 rep $rep"; timeout 600 python tools/sweep.py run cavity128_d3q19_bgk_fp32 --steps 400 --only q19_f32f32__L2HINT
done
echo "== cavity128 --graph 20"; XLB_SWEEP_BENCH_ARGS="--graph 20" timeout 600 python tools/sweep.py run cavity128_d3q19_bgk_fp32 --steps 400 --only q19_f32f32__L2HINT
echo "== cavity128 shape 256,256,256"; timeout 600 python tools/sweep.py run cavity128_d3q19_bgk_fp32 --steps 100 --only q19_f32f32__L2HINT --shape 256,256,256
echo "== cavity512_d3q19_bgk_fp32"; timeout 600 python tools/sweep.py run cavity512_d3q19_bgk_fp32 --steps 40 --only q19_f32f32__L2HINT
} > $out/${tag}_sweep_l2hint.log 2>&1
cat $out/${tag}_sweep_l2hint.log
