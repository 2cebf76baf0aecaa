#!/bin/bash
# round 2, GPU call I (1 GPU): gpt_render_static's test alone in a fresh process, then the records of the final build (run_r02_f.sh:
# whole GPU suite -- the new test collected last --, frame counters, algorithmic bytes, bench both arms, launch list, ncu capture)
mkdir -p gpurun_out
( time timeout -k 10 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q ) > gpurun_out/pytest_no_alloc_alone.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_no_alloc_alone.log
bash profiles/tools/run_r02_f.sh
tail -3 gpurun_out/pytest_no_alloc_alone.log gpurun_out/pytest_gpu.log; cat gpurun_out/bench.json
