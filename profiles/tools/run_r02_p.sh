#!/bin/bash
# round 2, GPU call P (1 GPU): shade kernel with the normalisations and the material tail hoisted out of the divergent branches --
# whole GPU suite (bit-exactness against the oracle) and the bench line
mkdir -p gpurun_out
( time timeout -k 10 1500 python -m pytest tests -m gpu -q -x ) > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
for rep in 1 2; do python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-configs > gpurun_out/bench_$rep.json 2> gpurun_out/bench_$rep.err; done
for f in gpurun_out/bench_?.json; do python -c "
import json
d=json.loads(open('$f').read().strip().splitlines()[-1]); print('$f', round(d['value']), d['roofline']['kernel_class_ms_per_step'], d['fast_mode']['value'], d['tree_as_uploaded']['value'])"; done
