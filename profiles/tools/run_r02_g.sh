#!/bin/bash
# round 2, GPU call G: A/B of load-instruction variants (256-bit node loads, streaming hints for path state) -- variants/*.so built here
mkdir -p gpurun_out
cp g.p.u-pathtracer_b200/libgpt_b200.so /tmp/default.so
for f in /tmp/default.so variants/*.so; do
  cp $f g.p.u-pathtracer_b200/libgpt_b200.so
  n=$(basename $f .so)
  for rep in 1 2; do
    timeout -k 5 300 python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-configs > gpurun_out/bench_${n}_$rep.json 2> gpurun_out/bench_${n}_$rep.err
  done
  timeout -k 5 900 python profiles/tools/configs_bench.py cfg2 cfg4 > gpurun_out/configs_$n.log 2>&1
done
cp /tmp/default.so g.p.u-pathtracer_b200/libgpt_b200.so
timeout -k 10 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_wide.py tests/test_gpu_configs.py -q -x > gpurun_out/pytest_ld256.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_ld256.log
python -c "
import json, gpt_b200 as g
with g.Context(0) as c: print(json.dumps(c.measure_peaks()))" > gpurun_out/peaks.json 2> gpurun_out/peaks.err
ls -la gpurun_out
