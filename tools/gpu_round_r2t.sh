#!/bin/bash
# r2t: multi-row tiles of the bulk pipeline on 128-wide rows (configs[2]); K1 load-order mode 2
tag=${1:-r2t}
out=gpurun_out
mkdir -p $out
{
for rep in 1 2; do
echo "== sphere384_d3q27_kbc_fp32 (q27_f32f32) rep $rep"; timeout 600 python tools/sweep.py run sphere384_d3q27_kbc_fp32 --steps 200 --only q27_f32f32__MULTIROW
done
echo "== sphere768 shape 768,512,128"; timeout 600 python tools/sweep.py run sphere768_d3q27_kbc_fp32 --steps 40 --only q27_f32f32__MULTIROW --shape 768,512,128
echo "== sphere768_d3q27_kbc_fp32 (unchanged path)"; timeout 600 python tools/sweep.py run sphere768_d3q27_kbc_fp32 --steps 40 --only q27_f32f32__MULTIROW
echo "== cavity512_d3q19_bgk_fp32 (q19_f32f32)"; timeout 600 python tools/sweep.py run cavity512_d3q19_bgk_fp32 --steps 40 --only q19_f32f32__K1
echo "== cavity128_d3q19_bgk_fp32 (q19_f32f32)"; timeout 600 python tools/sweep.py run cavity128_d3q19_bgk_fp32 --steps 400 --only q19_f32f32__K1
} > $out/${tag}_sweep.log 2>&1
cat $out/${tag}_sweep.log
# correctness of the multi-row path with the tuning library (only q27_f32f32 is real in it): the D3Q27 FP32 bulk cases
XLB_B200_LIB=xlb_b200/_lib/sweep/lib_q27_f32f32__MULTIROW1.so XLB_B200_TUNING_LIB=1 timeout 600 python -m pytest tests/test_gpu_step.py -q -k "test_bulk_variant and (sphere3d_28x16x128 or sphere3d_28x15x128 or periodic_3x4x128 or periodic_2x2x128)" > $out/${tag}_pytest_multirow.log 2>&1; echo "pytest=$?"; tail -n 5 $out/${tag}_pytest_multirow.log
