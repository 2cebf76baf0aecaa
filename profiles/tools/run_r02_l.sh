#!/bin/bash
# round 2, GPU call L (1 GPU): builder knobs for the rebuilt tree (SAH node cost, leaf size, PLOC radius) on the bench mesh and the
# 100 k-triangle dragon (reference SBVH arrays -> rebuilt), GPT_REBUILD=1 so that no trial noise is in the numbers
mkdir -p gpurun_out
run() { tag=$1; shift; env GPT_REBUILD=1 "$@" timeout 300 python profiles/tools/rebuild_sweep.py 2> gpurun_out/sweep_$tag.err | tail -1 > gpurun_out/sweep_$tag.json; }
run base
for c in 0.6 0.8 1.3 1.7 2.5; do run cost$c GPT_SAH_NODE_COST=$c; done
for l in 2 4 16; do run leaf$l GPT_REBUILD_MAX_LEAF=$l; done
for r in 8 16 32; do run rad$r GPT_PLOC_RADIUS=$r; done
run k3 GPT_WF_NODE_K=3
run keep GPT_REBUILD=0
for f in gpurun_out/sweep_*.json; do python -c "
import json
d=json.loads(open('$f').read())
print('$f', ' | '.join('%s %.2f ms ext %.2f nodes %d depth %d' % (k, v['ms'], v['extend_ms'], v['rebuilt']['inner_nodes'], v['rebuilt']['depth']) for k, v in d.items() if k != 'env'))" ; done
