#!/bin/bash
# 8-GPU evidence after the FP16 load reorder and the cost-balanced slabs: default bench at N = 8 (weak headline + nrank_bitwise +
# configs[3] strong), the configs[4] wind tunnel (1536^3 D3Q27 KBC, weak, balanced planes), bitwise tests at 8 ranks.
#   gpurun --gpus 8 --timeout 900 -- 'bash tools/mgpu_round_r2q.sh r2q'
tag=${1:-r2q}
out=gpurun_out
mkdir -p $out
nvidia-smi --query-gpu=name --format=csv,noheader | sort | uniq -c > $out/${tag}_gpus.txt
run() { n=$1; shift; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n "$@"; }
run 8 --steps 50 --warmup 5 > $out/${tag}_weak_n8.json 2> $out/${tag}_weak_n8.err; echo "n8=$?"; tail -n 3 $out/${tag}_weak_n8.err | cut -c1-200
run 8 --workload windtunnel192x1536x1536_d3q27_kbc_fp32 --steps 20 --warmup 3 --no-configs > $out/${tag}_windtunnel_n8.json 2> $out/${tag}_windtunnel_n8.err; echo "wt8=$?"; tail -n 3 $out/${tag}_windtunnel_n8.err | cut -c1-200
timeout 400 python -m pytest tests/test_gpu_multi.py -x -q -k "8" > $out/${tag}_mgpu_pytest.log 2>&1; echo "pytest=$?"; tail -n 6 $out/${tag}_mgpu_pytest.log
python - <<PY
import json,glob
for f in sorted(glob.glob('$out/${tag}_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e:
        print(f,'unreadable',e); continue
    st=d.get('strong') or {}
    print(f.split('/')[-1], 'N',d['n_gpus'], round(d['value']), 'ms',round(d['ms_per_step'],4), 'frac',round(d['roofline']['frac'],3), d['config'].get('halo','')[:9], d['config'].get('slab_planes'), 'bitwise',d.get('nrank_bitwise'), 'strong',round(st.get('mlups',0)), round(st.get('per_gpu_frac',0),3))
PY
