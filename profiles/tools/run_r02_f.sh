#!/bin/bash
# round 2, GPU call F (1 GPU): the records of the final build -- GPU suite, per-frame ncu counters (-> profiles/r02_frame_counters.json),
# oracle algorithmic bytes (-> profiles/r02_algorithmic_bytes.json), bench (both arms), launch list, one --set full capture
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt
( time timeout -k 10 1500 python -m pytest tests -m gpu -q ) > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
M=gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,l1tex__t_bytes.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,l1tex__t_sector_hit_rate.pct,lts__t_sector_hit_rate.pct
for s in bench cfg2_dragon100k_sbvh cfg2_dragon1p6M; do
  timeout 600 ncu --profile-from-start off --clock-control none --csv --log-file gpurun_out/ncu_frame_$s.csv --metrics $M python profiles/tools/ncu_frame.py $s 16 > gpurun_out/ncu_frame_$s.log 2>&1
done
timeout 900 ncu --profile-from-start off --clock-control none --csv --log-file gpurun_out/ncu_frame_cfg4_20M.csv --metrics $M python profiles/tools/ncu_frame.py cfg4_20M 4 > gpurun_out/ncu_frame_cfg4_20M.log 2>&1
python profiles/tools/ncu_frame_summary.py gpurun_out/ncu_frame_*.csv > gpurun_out/frame_counters.json 2> gpurun_out/frame_counters.err
python profiles/tools/algo_bytes.py > gpurun_out/algo_bytes.json 2> gpurun_out/algo_bytes.err
# the bench reads the two committed files: use this run's
python -c "import json; json.load(open('gpurun_out/frame_counters.json'))" && cp gpurun_out/frame_counters.json profiles/r02_frame_counters.json
python -c "import json; json.load(open('gpurun_out/algo_bytes.json'))" && cp gpurun_out/algo_bytes.json profiles/r02_algorithmic_bytes.json
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/bench.err
python bench.py --impl reference > gpurun_out/bench_reference_arm.json 2> gpurun_out/bench_reference_arm.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-configs > gpurun_out/ncu_launches.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:k_wf_extend_ww|k_wf_shade" -s 120 -c 2 -o gpurun_out/r02_full python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-configs > gpurun_out/ncu_full.log 2>&1
ncu -i gpurun_out/r02_full.ncu-rep --page raw --csv > gpurun_out/r02_full_raw.csv 2> gpurun_out/ncu_full_raw.err
python profiles/tools/small_frames.py > gpurun_out/small_frames.log 2>&1
rm -f gpurun_out/r02_full.ncu-rep.tmp; ls -la gpurun_out
