#!/bin/bash
# round 2, GPU call Q (1 GPU): A/B of library variants (variants/*.so) with shade_ab.py (compat + FAST) on the bench mesh and the dragon
mkdir -p gpurun_out
cp g.p.u-pathtracer_b200/libgpt_b200.so /tmp/default.so
for f in /tmp/default.so variants/*.so; do
  cp $f g.p.u-pathtracer_b200/libgpt_b200.so; n=$(basename $f .so)
  for rep in 1 2; do GPT_REBUILD=1 python profiles/tools/shade_ab.py 2> gpurun_out/ab_${n}_$rep.err | tail -1 > gpurun_out/ab_${n}_$rep.json; echo $n bunny $(cat gpurun_out/ab_${n}_$rep.json); done
  GPT_REBUILD=1 python profiles/tools/shade_ab.py dragon_cornell 2> gpurun_out/abd_${n}.err | tail -1 > gpurun_out/abd_${n}.json; echo $n dragon $(cat gpurun_out/abd_${n}.json)
done
cp /tmp/default.so g.p.u-pathtracer_b200/libgpt_b200.so
