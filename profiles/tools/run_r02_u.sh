#!/bin/bash
# round 2, GPU call U (1 GPU): the split-leaf tree -- parity tests (forced and under the default policy), config tests, timing
mkdir -p gpurun_out
( time timeout -k 10 1200 python -m pytest tests/test_gpu_rebuild.py tests/test_gpu_configs.py -m gpu -q -x ) > gpurun_out/pytest_split.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_split.log
tail -25 gpurun_out/pytest_split.log
GPT_SPLIT=1 python profiles/tools/shade_ab.py dragon_cornell 2>/dev/null | tail -1
GPT_SPLIT=1 python profiles/tools/shade_ab.py bunny_cornell 2>/dev/null | tail -1
GPT_REBUILD=1 python profiles/tools/shade_ab.py dragon_cornell 2>/dev/null | tail -1
python bench.py --steps 8 --warmup 3 --no-cpu-baseline > gpurun_out/bench_split.json 2> gpurun_out/bench_split.err; echo "bench rc=$?"
python -c "
import json
d=json.loads(open('gpurun_out/bench_split.json').read().strip().splitlines()[-1]); print(round(d['value']), d['run']['tree']['in_use'], d['run']['tree']['trees'])
for k,c in d['configs'].items(): print(k, round(c['value']), c['tree_in_use'], c['trees']['ns_per_ray'])"
