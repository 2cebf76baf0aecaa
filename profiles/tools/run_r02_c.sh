#!/bin/bash
# round 2, GPU call C: where does the deferred-shading time go (defer_min sweep + per-launch list), small frames after the launch
# fusion, FAST-mode tests
mkdir -p gpurun_out
for d in 0 2 4 6; do
  GPT_SHADE_DEFER_MIN=$d timeout -k 5 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-configs > gpurun_out/bench_defer_$d.json 2> gpurun_out/bench_defer_$d.err
done
M=gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,launch__grid_size
for d in 0 6; do
  GPT_SHADE_DEFER_MIN=$d timeout 600 ncu --profile-from-start off --clock-control none --csv --log-file gpurun_out/ncu_defer_$d.csv --metrics $M python profiles/tools/ncu_frame.py bench 16 > gpurun_out/ncu_defer_$d.log 2>&1
done
GPT_SHADE_DEFER_MIN=0 python profiles/tools/small_frames.py > gpurun_out/small_frames.log 2>&1
( time timeout -k 10 1500 python -m pytest tests -m gpu -q ) > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
ls -la gpurun_out
