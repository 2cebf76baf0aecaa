#!/bin/bash
# round 2, GPU call J (1 GPU): A/B of software prefetch (first triangle of a postponed leaf, nearest pushed child) and streaming
# hints for the path state -- variants/*.so built beforehand -- then the bench record of the default build (both arms)
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
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/bench.err
python bench.py --impl reference > gpurun_out/bench_reference_arm.json 2> gpurun_out/bench_reference_arm.err
for f in gpurun_out/bench_*_?.json; do python -c "
import json,sys
d=json.loads(open('$f').read().strip().splitlines()[-1]); print('$f', round(d['value']), d['roofline']['kernel_class_ms_per_step'])" ; done
grep -h "Mrays\|ms" gpurun_out/configs_*.log | cut -c1-240 | tail -40
