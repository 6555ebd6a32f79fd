#!/bin/bash
# Round 2, first single-GPU call: whole GPU suite (incl. the BASELINE-size tests), smoke, the default bench with its
# configs block, the reference arm, the FP16 stream-only experiment and ncu captures of configs[2] (384x128x128 D3Q27 KBC)
# and of the FP16 kernel.
#   gpurun --timeout 1500 -- 'bash tools/gpu_round_r2a.sh r2a'
tag=${1:-r2a}
out=gpurun_out
mkdir -p $out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > $out/${tag}_gpu.txt; nproc >> $out/${tag}_gpu.txt; free -g | head -2 >> $out/${tag}_gpu.txt
timeout 1500 python -m pytest tests -m gpu -q -x --durations=15 > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest=$?"; tail -n 25 $out/${tag}_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $out/${tag}_smoke.log 2>&1; echo "smoke=$?"; tail -n 2 $out/${tag}_smoke.log
timeout 900 python bench.py --steps 20 --warmup 5 > $out/${tag}_bench_default.json 2> $out/${tag}_bench_default.err; echo "bench=$?"; cut -c1-400 $out/${tag}_bench_default.json; tail -n 5 $out/${tag}_bench_default.err
timeout 300 python bench.py --impl reference --steps 20 --warmup 5 > $out/${tag}_bench_reference.json 2> $out/${tag}_bench_reference.err; echo "ref=$?"; cut -c1-300 $out/${tag}_bench_reference.json
# FP16: same loads / shifts / stores without the collision
timeout 300 python tools/sweep.py run cavity768_d3q19_bgk_fp16 --steps 30 > $out/${tag}_sweep_f16_stream_only.log 2>&1; tail -n 4 $out/${tag}_sweep_f16_stream_only.log
# ncu: launch list of the default command, then full captures of the two kernels under work
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $out/${tag}_launches_default_bench.csv python bench.py --steps 4 --warmup 3 --no-e2e --no-cpu-baseline --no-configs > /dev/null 2>&1; echo "ncu_list=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:step_bulk -s 6 -c 1 -o $out/${tag}_ncu_c3_sphere384 -f python bench.py --workload sphere384_d3q27_kbc_fp32 --steps 4 --warmup 3 --no-e2e --no-cpu-baseline --no-configs > /dev/null 2>&1; echo "ncu_c3=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:step_vector_h2 -s 6 -c 1 -o $out/${tag}_ncu_f16 -f python bench.py --workload cavity768_d3q19_bgk_fp16 --steps 4 --warmup 3 --no-e2e --no-cpu-baseline --no-configs > /dev/null 2>&1; echo "ncu_f16=$?"
ls -la $out | tail -n 15
