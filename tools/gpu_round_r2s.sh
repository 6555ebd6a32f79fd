#!/bin/bash
# r2s: is the 0.81-0.89 of the nz = 128 configs (C1 128^3, C3 384x128x128) a matter of ROW LENGTH (512-byte rows) or of lattice size?
tag=${1:-r2s}
out=gpurun_out
mkdir -p $out
b() { timeout 300 python bench.py --no-e2e --no-cpu-baseline --no-configs "$@" 2>> $out/${tag}_bench.err | grep '^{' >> $out/${tag}_rows.jsonl; }
b --workload cavity512_d3q19_bgk_fp32 --steps 40 --warmup 5 --shape 512,512,128
b --workload cavity512_d3q19_bgk_fp32 --steps 40 --warmup 5 --shape 512,128,512
b --workload cavity512_d3q19_bgk_fp32 --steps 40 --warmup 5 --shape 512,256,256
b --workload cavity512_d3q19_bgk_fp32 --steps 40 --warmup 5 --shape 512,1024,64
b --workload cavity128_d3q19_bgk_fp32 --steps 400 --warmup 5 --shape 64,128,256
b --workload cavity128_d3q19_bgk_fp32 --steps 400 --warmup 5 --shape 128,128,128
b --workload cavity128_d3q19_bgk_fp32 --steps 400 --warmup 5 --shape 256,128,64
b --workload sphere768_d3q27_kbc_fp32 --steps 40 --warmup 5 --shape 768,512,128
b --workload sphere768_d3q27_kbc_fp32 --steps 40 --warmup 5 --shape 768,256,256
b --workload sphere768_d3q27_kbc_fp32 --steps 40 --warmup 5 --shape 768,128,512
b --workload sphere384_d3q27_kbc_fp32 --steps 200 --warmup 5 --shape 384,64,256
b --workload sphere384_d3q27_kbc_fp32 --steps 200 --warmup 5 --shape 384,128,128
b --workload sphere384_d3q27_kbc_fp32 --steps 200 --warmup 5 --shape 192,128,256
python - <<PY
import json
for l in open('$out/${tag}_rows.jsonl'):
    d=json.loads(l); print(d['config']['workload'], d['config']['shape'], round(d['value']), round(d['ms_per_step'],4), round(d['roofline']['frac'],3))
PY
