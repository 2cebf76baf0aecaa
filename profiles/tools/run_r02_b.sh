#!/bin/bash
# round 2, GPU call B: extend-schedule A/B (ww / p3 / p1 / r1) on the bench scene and BASELINE configs 2, 4; config parity re-run
mkdir -p gpurun_out
for v in "ww 0" "p3 0" "p3 26" "p1 0" "p1 10" "r1 0" "r1 8" "r1 24"; do set -- $v
  GPT_WF_EXTEND=$1 GPT_WF_HOLD_K=$2 timeout -k 5 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-configs > gpurun_out/bench_$1_$2.json 2> gpurun_out/bench_$1_$2.err
done
for e in p3 p1 r1; do GPT_WF_EXTEND=$e timeout -k 5 900 python profiles/tools/configs_bench.py cfg2 cfg4 > gpurun_out/configs_$e.log 2>&1; done
for e in p1 r1; do GPT_WF_EXTEND=$e GPT_WF_HOLD_K=8 timeout -k 5 900 python profiles/tools/configs_bench.py cfg2 cfg4 > gpurun_out/configs_${e}_8.log 2>&1; done
( time timeout -k 10 900 python -m pytest tests/test_gpu_configs.py -q ) > gpurun_out/pytest_configs.log 2>&1
GPT_WF_EXTEND=p3 timeout -k 10 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_wide.py -q -x > gpurun_out/pytest_p3.log 2>&1
GPT_WF_EXTEND=r1 timeout -k 10 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_wide.py -q -x > gpurun_out/pytest_r1.log 2>&1
python profiles/tools/algo_bytes.py > gpurun_out/algo_bytes.json 2> gpurun_out/algo_bytes.err
python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/bench.err
ls -la gpurun_out
