#!/bin/bash
# round 2, GPU call H: 32-byte ray records with one 256-bit access (default) against two 128-bit accesses (variants/ray128.so)
mkdir -p gpurun_out
timeout -k 10 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_wide.py tests/test_gpu_build.py tests/test_gpu_configs.py -q -x > gpurun_out/pytest_ray32.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_ray32.log
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
ls -la gpurun_out
