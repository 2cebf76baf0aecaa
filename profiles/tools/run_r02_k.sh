#!/bin/bash
# round 2, GPU call K (1 GPU): the rebuilt tree (gpt_rebuild_info) -- its parity tests, the config tests under the new default policy,
# the bench headline with it, and the node-loop exit policy (GPT_WF_NODE_K) re-tuned for trees with small leaves
mkdir -p gpurun_out
( time timeout -k 10 900 python -m pytest tests/test_gpu_rebuild.py tests/test_gpu_configs.py tests/test_zz_no_alloc.py -m gpu -q -x ) > gpurun_out/pytest_rebuild.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_rebuild.log
tail -5 gpurun_out/pytest_rebuild.log
python bench.py --steps 8 --warmup 3 --no-cpu-baseline > gpurun_out/bench_rebuild.json 2> gpurun_out/bench_rebuild.err; echo "bench rc=$?"
for k in 2 3 4 6 8; do
  GPT_WF_NODE_K=$k timeout 300 python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-configs > gpurun_out/bench_k$k.json 2> gpurun_out/bench_k$k.err
done
GPT_REBUILD=0 timeout 300 python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-configs > gpurun_out/bench_norebuild.json 2> gpurun_out/bench_norebuild.err
for f in gpurun_out/bench_*.json; do python -c "
import json
d=json.loads(open('$f').read().strip().splitlines()[-1]); print('$f', round(d['value']), d['run'].get('tree',{}).get('in_use'), d['roofline']['kernel_class_ms_per_step'], (d.get('tree_as_uploaded') or {}).get('value'))" ; done
( time timeout -k 10 1500 python -m pytest tests -m gpu -q -x ) > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
