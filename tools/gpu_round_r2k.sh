#!/bin/bash
# same-box A/B of the arithmetic diet (XLB_TUNE_DIET: rcp.approx + Newton, Horner equilibrium pair, pair-sum moments)
out=gpurun_out; tag=${1:-r2k}
run() { echo "== $1 ($2)" >> $out/${tag}_sweep_diet.log; timeout 600 python tools/sweep.py run $1 --steps $3 --only $2 >> $out/${tag}_sweep_diet.log 2>&1; }
for rep in 1 2; do
run cavity768_d3q19_bgk_fp16 q19_f32f16 30
run sphere384_d3q27_kbc_fp32 q27_f32f32 200
run sphere768_d3q27_kbc_fp32 q27_f32f32 50
run cavity512_d3q19_bgk_fp32 q19_f32f32 30
done
grep -v BEST $out/${tag}_sweep_diet.log
