#!/bin/bash
# round 2, GPU call A: smoke gate for the pooled extend kernel, full GPU suite, bench A/B of the extend schedules, small frames,
# peaks, per-frame counter passes (see profiles/r02_*)
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt
timeout -k 5 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; rc=$?; echo "smoke rc=$rc" >> gpurun_out/smoke.log
if [ $rc -ne 0 ]; then export GPT_WF_EXTEND=ww; echo "POOLED KERNEL FAILED SMOKE: falling back to GPT_WF_EXTEND=ww for the rest" >> gpurun_out/smoke.log; fi
( time timeout -k 10 1500 python -m pytest tests -m gpu -q ) > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
for v in "ww 0 10" "pq 0 10" "pq 0 8" "pq 12 10" "pq 28 10"; do set -- $v
  GPT_WF_EXTEND=$1 GPT_WF_HOLD_K=$2 GPT_WF_PQ_BLOCKS=$3 timeout -k 5 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$1_$2_$3.json 2> gpurun_out/bench_$1_$2_$3.err
done
python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/bench.err
python -c "
import json, gpt_b200 as g
with g.Context(0) as c: print(json.dumps(c.measure_peaks()))" > gpurun_out/peaks.json 2> gpurun_out/peaks.err
python profiles/tools/small_frames.py > gpurun_out/small_frames.log 2>&1
for e in ww pq; do GPT_WF_EXTEND=$e timeout -k 5 900 python profiles/tools/configs_bench.py cfg2 cfg4 > gpurun_out/configs_$e.log 2>&1; done
M=gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,l1tex__t_bytes.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,l1tex__t_sector_hit_rate.pct,lts__t_sector_hit_rate.pct
for s in bench cfg2_dragon100k_sbvh cfg2_dragon1p6M; do
  timeout 600 ncu --profile-from-start off --clock-control none --csv --log-file gpurun_out/ncu_frame_$s.csv --metrics $M python profiles/tools/ncu_frame.py $s 16 > gpurun_out/ncu_frame_$s.log 2>&1
done
timeout 900 ncu --profile-from-start off --clock-control none --csv --log-file gpurun_out/ncu_frame_cfg4_20M.csv --metrics $M python profiles/tools/ncu_frame.py cfg4_20M 4 > gpurun_out/ncu_frame_cfg4_20M.log 2>&1
python profiles/tools/ncu_frame_summary.py gpurun_out/ncu_frame_*.csv > gpurun_out/frame_counters.json 2> gpurun_out/frame_counters.err
ls -la gpurun_out
