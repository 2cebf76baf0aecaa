#!/bin/bash
# round 2, GPU call O (1 GPU): occupancy of the extend / shade kernels and the refill / grid knobs re-measured ON THE REBUILT TREE
# (the extend kernel issues 0.72 of the peak there instead of 0.80: more resident warps may pay now); variants/*.so built beforehand
mkdir -p gpurun_out
cp g.p.u-pathtracer_b200/libgpt_b200.so /tmp/default.so
one() { tag=$1; shift; for rep in 1 2; do env GPT_REBUILD=1 "$@" timeout 300 python profiles/tools/rebuild_sweep.py 2> gpurun_out/sw_${tag}_$rep.err | tail -1 > gpurun_out/sw_${tag}_$rep.json; done; }
one default
for f in variants/*.so; do cp $f g.p.u-pathtracer_b200/libgpt_b200.so; one $(basename $f .so); done
cp /tmp/default.so g.p.u-pathtracer_b200/libgpt_b200.so
for r in 4 12 16 24; do one refill$r GPT_WF_REFILL_MIN=$r; done
for w in 8 32; do one waves$w GPT_SHADE_WAVES=$w; done
for k in 4 6; do one k$k GPT_WF_NODE_K=$k; done
for f in gpurun_out/sw_*.json; do python -c "
import json
d=json.loads(open('$f').read())
print('$f', ' | '.join('%s %.2f ms ext %.2f' % (k, v['ms'], v['extend_ms']) for k, v in d.items() if k != 'env'))" ; done
