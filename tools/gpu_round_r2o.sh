#!/bin/bash
# r2o: A/B of the phased load order (ids -> all population loads -> classification -> shifts) in K1c (FP16) and K1 (FP32)
tag=${1:-r2o}
out=gpurun_out
mkdir -p $out
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv,noheader > $out/${tag}_gpu.txt
{
echo "== cavity768_d3q19_bgk_fp16 (q19_f32f16)"; timeout 600 python tools/sweep.py run cavity768_d3q19_bgk_fp16 --steps 20 --only q19_f32f16__H2
echo "== cavity512_d3q19_bgk_fp32 (q19_f32f32)"; timeout 600 python tools/sweep.py run cavity512_d3q19_bgk_fp32 --steps 40 --only q19_f32f32__K1
echo "== cavity128_d3q19_bgk_fp32 (q19_f32f32)"; timeout 600 python tools/sweep.py run cavity128_d3q19_bgk_fp32 --steps 400 --only q19_f32f32__K1
echo "== cavity768_d3q19_bgk_fp16 again (order effects)"; timeout 600 python tools/sweep.py run cavity768_d3q19_bgk_fp16 --steps 20 --only q19_f32f16__H2
} > $out/${tag}_sweep_phased.log 2>&1
cat $out/${tag}_sweep_phased.log
