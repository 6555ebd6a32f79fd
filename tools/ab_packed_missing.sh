for rep in 1 2; do for pm in 1 0; do for w in sphere384_d3q27_kbc_fp32 sphere768_d3q27_kbc_fp32; do
XLB_B200_PACKED_MISSING=$pm timeout 300 python bench.py --workload $w --steps 200 --warmup 10 --no-e2e --no-cpu-baseline --no-configs 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('packed=$pm', '$w', round(d['value']), round(d['roofline']['frac'],4))"
done; done; done
for pm in 1 0; do XLB_B200_PACKED_MISSING=$pm timeout 400 python bench.py --workload windtunnel192x1536x1536_d3q27_kbc_fp32 --radius 192 --steps 10 --warmup 3 --no-e2e --no-cpu-baseline --no-configs 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('packed=$pm bigsphere', round(d['value']), round(d['roofline']['frac'],4))"; done
