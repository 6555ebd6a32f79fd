#!/bin/bash
# r2x: 2-GPU check of the final library: bitwise tests at 2 ranks (+ the reference's distributed example), default bench at N = 2
tag=${1:-r2x}
out=gpurun_out
mkdir -p $out
nvidia-smi --query-gpu=name --format=csv,noheader | sort | uniq -c > $out/${tag}_gpus.txt
timeout 600 python -m pytest tests/test_gpu_multi.py -x -q > $out/${tag}_mgpu2_pytest.log 2>&1; echo "pytest=$?"; tail -n 4 $out/${tag}_mgpu2_pytest.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 50 --warmup 5 > $out/${tag}_weak_n2.json 2> $out/${tag}_weak_n2.err; echo "n2=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --workload sphere768_d3q27_kbc_fp32 --steps 40 --warmup 5 --no-configs > $out/${tag}_sphere_n2.json 2> $out/${tag}_sphere_n2.err; echo "s2=$?"
python - <<PY
import json,glob
for f in sorted(glob.glob('$out/${tag}_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e:
        print(f,'unreadable',e); continue
    st=d.get('strong') or {}
    print(f.split('/')[-1], 'N',d['n_gpus'], round(d['value']), 'ms',round(d['ms_per_step'],4), 'frac',round(d['roofline']['frac'],3), d['config'].get('slab_planes'), 'bitwise',d.get('nrank_bitwise'), 'strong',round(st.get('mlups',0)), round(st.get('per_gpu_frac',0),3))
PY
