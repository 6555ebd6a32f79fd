#!/bin/bash
# Round 2 multi-GPU evidence on N GPUs of one box (N = $2, default 2): bitwise tests (peer push / graph / NCCL schedules, 2-D
# slabs, the reference's distributed example), then the bench at N with the fused peer push (default), without the CUDA graph,
# and with the NCCL schedule, plus the strong-scaling line the default run carries.
#   gpurun --gpus 2 --timeout 900 -- 'bash tools/mgpu_round_r2.sh r2c 2'
tag=${1:-r2c}
n=${2:-2}
out=gpurun_out
mkdir -p $out
nvidia-smi --query-gpu=name --format=csv,noheader | sort | uniq -c > $out/${tag}_gpus.txt; nvidia-smi topo -m >> $out/${tag}_gpus.txt 2>&1
run() { timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n "$@"; }
timeout 900 python -m pytest tests/test_gpu_multi.py -x -q -k "$n" > $out/${tag}_mgpu${n}_pytest.log 2>&1; echo "pytest=$?"; tail -n 12 $out/${tag}_mgpu${n}_pytest.log
run --steps 50 --warmup 5 > $out/${tag}_weak_n${n}_peer_graph.json 2> $out/${tag}_weak_n${n}_peer_graph.err; echo "peer_graph=$?"; cut -c1-200 $out/${tag}_weak_n${n}_peer_graph.json; tail -n 4 $out/${tag}_weak_n${n}_peer_graph.err
run --steps 50 --warmup 5 --no-dist-graph --no-configs > $out/${tag}_weak_n${n}_peer_eager.json 2> $out/${tag}_weak_n${n}_peer_eager.err; echo "peer_eager=$?"; cut -c1-200 $out/${tag}_weak_n${n}_peer_eager.json
run --steps 50 --warmup 5 --halo nccl --no-configs > $out/${tag}_weak_n${n}_nccl.json 2> $out/${tag}_weak_n${n}_nccl.err; echo "nccl=$?"; cut -c1-200 $out/${tag}_weak_n${n}_nccl.json
timeout 300 python bench.py --steps 50 --warmup 5 --no-e2e --no-cpu-baseline --no-configs > $out/${tag}_weak_n1.json 2> /dev/null; echo "n1=$?"; cut -c1-160 $out/${tag}_weak_n1.json
python - <<PY
import json,glob
for f in sorted(glob.glob('$out/${tag}_weak_n*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e:
        print(f,'unreadable',e); continue
    print(f.split('/')[-1], round(d['value']), round(d['ms_per_step'],4), d['config'].get('halo'), d.get('nrank_bitwise'), (d.get('strong') or {}).get('mlups'), (d.get('strong') or {}).get('per_gpu_frac'))
PY
