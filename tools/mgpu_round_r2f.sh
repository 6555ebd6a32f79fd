#!/bin/bash
# 8-GPU evidence for round 2: bitwise tests at 8 and 4 ranks, the default bench at N = 8 / 4 (weak headline + nrank_bitwise +
# configs[3] strong), and the configs[4] wind tunnel (1536^3 D3Q27 KBC, weak) at N = 8.
#   gpurun --gpus 8 --timeout 1500 -- 'bash tools/mgpu_round_r2f.sh r2f'
tag=${1:-r2f}
out=gpurun_out
mkdir -p $out
nvidia-smi --query-gpu=name --format=csv,noheader | sort | uniq -c > $out/${tag}_gpus.txt
run() { n=$1; shift; timeout 700 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n "$@"; }
timeout 900 python -m pytest tests/test_gpu_multi.py -x -q -k "8 or 4" > $out/${tag}_mgpu_pytest.log 2>&1; echo "pytest=$?"; tail -n 6 $out/${tag}_mgpu_pytest.log
run 8 --steps 50 --warmup 5 > $out/${tag}_weak_n8.json 2> $out/${tag}_weak_n8.err; echo "n8=$?"; tail -n 3 $out/${tag}_weak_n8.err | cut -c1-200
run 4 --steps 50 --warmup 5 > $out/${tag}_weak_n4.json 2> $out/${tag}_weak_n4.err; echo "n4=$?"
run 8 --workload windtunnel192x1536x1536_d3q27_kbc_fp32 --steps 20 --warmup 3 --no-configs > $out/${tag}_windtunnel_n8.json 2> $out/${tag}_windtunnel_n8.err; echo "wt8=$?"; tail -n 3 $out/${tag}_windtunnel_n8.err | cut -c1-200
CUDA_VISIBLE_DEVICES=0 timeout 300 python bench.py --steps 50 --warmup 5 --no-e2e --no-cpu-baseline --no-configs > $out/${tag}_weak_n1.json 2> /dev/null; echo "n1=$?"
CUDA_VISIBLE_DEVICES=0 timeout 300 python bench.py --workload windtunnel192x1536x1536_d3q27_kbc_fp32 --steps 10 --warmup 3 --no-e2e --no-cpu-baseline --no-configs > $out/${tag}_windtunnel_n1.json 2> /dev/null; echo "wt1=$?"
python - <<PY
import json,glob
for f in sorted(glob.glob('$out/${tag}_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e:
        print(f,'unreadable',e); continue
    st=d.get('strong') or {}
    print(f.split('/')[-1], 'N',d['n_gpus'], round(d['value']), 'ms',round(d['ms_per_step'],4), 'frac',round(d['roofline']['frac'],3), d['config'].get('halo','')[:9], 'bitwise',d.get('nrank_bitwise'), 'strong',round(st.get('mlups',0)), round(st.get('per_gpu_frac',0),3))
PY
