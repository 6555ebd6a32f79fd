#!/bin/bash
# 2-GPU call: tuning sweeps on GPU 0 (bulk kernel: static / dynamic tile scheduling / 3 stages with the merged halo copies;
# FP16 kernel: decode loop, row pointers), then the multi-GPU tests and the N=2 bench with the bitwise check.
tag=${1:-r2e}
out=gpurun_out
mkdir -p $out
export CUDA_VISIBLE_DEVICES=0
for w in sphere384_d3q27_kbc_fp32 sphere768_d3q27_kbc_fp32 windtunnel192x1536x1536_d3q27_kbc_fp32; do
  steps=100; [ $w = windtunnel192x1536x1536_d3q27_kbc_fp32 ] && steps=10
  echo "== $w" >> $out/${tag}_sweep_q27.log
  timeout 600 python tools/sweep.py run $w --steps $steps --only q27 >> $out/${tag}_sweep_q27.log 2>&1
done
cat $out/${tag}_sweep_q27.log
echo "== cavity768_d3q19_bgk_fp16" > $out/${tag}_sweep_f16.log
timeout 600 python tools/sweep.py run cavity768_d3q19_bgk_fp16 --steps 30 --only q19 >> $out/${tag}_sweep_f16.log 2>&1
timeout 300 python bench.py --workload cavity768_d3q19_bgk_fp16 --steps 30 --warmup 3 --no-e2e --no-cpu-baseline --no-configs 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('main lib (old decode)', round(d['value']), round(d['roofline']['frac'],3))" >> $out/${tag}_sweep_f16.log
cat $out/${tag}_sweep_f16.log
unset CUDA_VISIBLE_DEVICES
timeout 900 python -m pytest tests/test_gpu_multi.py -x -q -k "2" > $out/${tag}_mgpu2_pytest.log 2>&1; echo "pytest=$?"; tail -n 5 $out/${tag}_mgpu2_pytest.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 50 --warmup 5 > $out/${tag}_weak_n2.json 2> $out/${tag}_weak_n2.err; echo "bench2=$?"
python - <<PY
import json
d=json.loads(open('$out/${tag}_weak_n2.json').read().strip().splitlines()[-1])
print(round(d['value']), round(d['ms_per_step'],4), d['config'].get('halo'), 'nrank_bitwise', d.get('nrank_bitwise'), d.get('strong'))
PY
