#!/bin/bash
# r3c: 2-GPU check of the closing library (backward plane order on odd steps in the edge / interior launches and their graphs)
tag=${1:-r3c}
out=gpurun_out
mkdir -p $out
timeout 400 python -m pytest tests/test_gpu_multi.py -x -q > $out/${tag}_mgpu2_pytest.log 2>&1; echo "pytest=$?"; tail -n 3 $out/${tag}_mgpu2_pytest.log
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 50 --warmup 5 > $out/${tag}_weak_n2.json 2> $out/${tag}_weak_n2.err; echo "n2=$?"
python - <<PY
import json
d=json.loads(open('$out/${tag}_weak_n2.json').read().strip().splitlines()[-1]); st=d.get('strong') or {}
print('N',d['n_gpus'], round(d['value']), 'ms',round(d['ms_per_step'],4), 'frac',round(d['roofline']['frac'],3), 'bitwise',d.get('nrank_bitwise'), 'strong',round(st.get('mlups',0)), round(st.get('per_gpu_frac',0),3))
PY
