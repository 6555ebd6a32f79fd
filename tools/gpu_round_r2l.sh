#!/bin/bash
out=gpurun_out; tag=${1:-r2l}
run() { echo "== $1 ($2)" >> $out/${tag}_sweep_diet_bgk.log; timeout 600 python tools/sweep.py run $1 --steps $3 --only $2 >> $out/${tag}_sweep_diet_bgk.log 2>&1; }
for rep in 1 2; do
run cavity768_d3q19_bgk_fp16 q19_f32f16 30
run cavity512_d3q19_bgk_fp32 q19_f32f32 30
done
grep -v BEST $out/${tag}_sweep_diet_bgk.log
