#!/bin/bash
# Round 2, single-GPU call after the machine loss: the no-alloc test alone in a fresh process, then the whole GPU suite, smoke,
# the default bench (configs block) + reference arm, the FP16 stream-only experiment and the ncu captures for profiles/.
#   gpurun --timeout 1800 -- 'bash tools/gpu_round_r2b.sh r2b'
tag=${1:-r2b}
out=gpurun_out
mkdir -p $out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > $out/${tag}_gpu.txt; nproc >> $out/${tag}_gpu.txt
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > $out/${tag}_pytest_noalloc_alone.log 2>&1; echo "noalloc_alone=$?"; tail -n 3 $out/${tag}_pytest_noalloc_alone.log
timeout 1500 python -m pytest tests -m gpu -q -x --durations=12 > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest=$?"; tail -n 22 $out/${tag}_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $out/${tag}_smoke.log 2>&1; echo "smoke=$?"; tail -n 2 $out/${tag}_smoke.log
timeout 900 python bench.py --steps 20 --warmup 5 > $out/${tag}_bench_default.json 2> $out/${tag}_bench_default.err; echo "bench=$?"; cut -c1-300 $out/${tag}_bench_default.json; tail -n 3 $out/${tag}_bench_default.err
timeout 300 python bench.py --impl reference --steps 20 --warmup 5 > $out/${tag}_bench_reference.json 2> $out/${tag}_bench_reference.err; echo "ref=$?"; cut -c1-200 $out/${tag}_bench_reference.json
if ls xlb_b200/_lib/sweep/*.so > /dev/null 2>&1; then
  timeout 300 python tools/sweep.py run cavity768_d3q19_bgk_fp16 --steps 30 > $out/${tag}_sweep_f16_stream_only.log 2>&1; tail -n 4 $out/${tag}_sweep_f16_stream_only.log
fi
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $out/${tag}_launches_default_bench.csv python bench.py --steps 4 --warmup 3 --no-e2e --no-cpu-baseline --no-configs > /dev/null 2>&1; echo "ncu_list=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:step_bulk -s 6 -c 1 -o $out/${tag}_ncu_c3_sphere384 -f python bench.py --workload sphere384_d3q27_kbc_fp32 --steps 4 --warmup 3 --no-e2e --no-cpu-baseline --no-configs > /dev/null 2>&1; echo "ncu_c3=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:step_vector_h2 -s 6 -c 1 -o $out/${tag}_ncu_f16 -f python bench.py --workload cavity768_d3q19_bgk_fp16 --steps 4 --warmup 3 --no-e2e --no-cpu-baseline --no-configs > /dev/null 2>&1; echo "ncu_f16=$?"
ls -la $out | tail -n 16
