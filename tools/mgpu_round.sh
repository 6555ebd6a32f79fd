#!/bin/bash
# Multi-GPU evidence run (one 8-GPU box): bitwise test at 8 ranks, weak-scaling curve of the headline workload,
# BASELINE configs[3] (1024^3 FP32FP16 strong) and configs[4]-like wind tunnel (1536^3 D3Q27 KBC weak) at N=8.
#   gpurun --gpus 8 -- 'bash tools/mgpu_round.sh r1'
tag=${1:-r1}
out=gpurun_out
run() { n=$1; shift; timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n "$@"; }
nvidia-smi -L | head -n 8 > $out/${tag}_gpus.txt
timeout 900 python -m pytest tests/test_gpu_multi.py -x -q -k "8" > $out/${tag}_mgpu8_pytest.log 2>&1; echo "pytest8=$?"
run 8 --steps 100 --warmup 5 > $out/${tag}_weak_n8.json 2> $out/${tag}_weak_n8.err; echo "weak8=$?"
run 8 --workload windtunnel192x1536x1536_d3q27_kbc_fp32 --steps 30 --warmup 3 > $out/${tag}_windtunnel_n8.json 2> $out/${tag}_windtunnel_n8.err; echo "wt8=$?"
run 8 --workload cavity1024_d3q19_bgk_fp16 --strong --steps 100 --warmup 5 > $out/${tag}_c4_strong_n8.json 2> $out/${tag}_c4_strong_n8.err; echo "c4=$?"
run 4 --steps 100 --warmup 5 > $out/${tag}_weak_n4.json 2> $out/${tag}_weak_n4.err; echo "weak4=$?"
run 2 --steps 100 --warmup 5 > $out/${tag}_weak_n2.json 2> $out/${tag}_weak_n2.err; echo "weak2=$?"
timeout 600 python bench.py --steps 100 --warmup 5 --no-e2e --no-cpu-baseline > $out/${tag}_weak_n1.json 2> $out/${tag}_weak_n1.err; echo "weak1=$?"
run 8 --steps 100 --warmup 5 --no-overlap > $out/${tag}_weak_n8_nooverlap.json 2> $out/${tag}_weak_n8_nooverlap.err; echo "weak8no=$?"
tail -n 3 $out/${tag}_mgpu8_pytest.log
cat $out/${tag}_weak_n8.json | cut -c1-300
