#!/bin/bash
# 8-GPU evidence after the BC cell list / packed-half2 kernels: bitwise test at 8 ranks, BASELINE configs[4]-like wind tunnel
# (1536^3 D3Q27 KBC, weak) and configs[3] (1024^3 FP32FP16, strong) at N=8.
#   gpurun --gpus 8 --timeout 420 -- 'bash tools/mgpu_round2.sh r1v'
tag=${1:-r1v}
out=gpurun_out
mkdir -p $out
run() { n=$1; shift; timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n "$@"; }
timeout 200 python -m pytest tests/test_gpu_multi.py -x -q -k "8" > $out/${tag}_mgpu8_pytest.log 2>&1; echo "pytest8=$?"; tail -n 3 $out/${tag}_mgpu8_pytest.log
run 8 --workload windtunnel192x1536x1536_d3q27_kbc_fp32 --steps 30 --warmup 3 > $out/${tag}_windtunnel_n8.json 2> $out/${tag}_windtunnel_n8.err; echo "wt8=$?"; cut -c1-260 $out/${tag}_windtunnel_n8.json
run 8 --workload cavity1024_d3q19_bgk_fp16 --strong --steps 100 --warmup 5 > $out/${tag}_c4_strong_n8.json 2> $out/${tag}_c4_strong_n8.err; echo "c4=$?"; cut -c1-260 $out/${tag}_c4_strong_n8.json
tail -n 3 $out/${tag}_windtunnel_n8.err
