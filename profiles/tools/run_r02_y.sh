#!/bin/bash
# round 2, GPU call Y (1 GPU): source-level instruction counts of one secondary shade and one secondary extend launch (rebuilt tree)
mkdir -p gpurun_out
timeout 900 ncu --section SourceCounters --section InstructionStats --clock-control none --import-source on -k "regex:k_wf_extend_ww|k_wf_shade" -s 200 -c 2 -o gpurun_out/src python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-configs > gpurun_out/ncu_src.log 2>&1
ncu -i gpurun_out/src.ncu-rep --page source --csv --print-source sass > gpurun_out/src_sass.csv 2> gpurun_out/src_sass.err
ncu -i gpurun_out/src.ncu-rep --page source --csv --print-source cuda > gpurun_out/src_cuda.csv 2> gpurun_out/src_cuda.err
ls -la gpurun_out; head -c 600 gpurun_out/src_cuda.csv
