#!/bin/bash
# round 2, GPU call W (1 GPU): the megakernel's own tree trial -- rebuild tests, whole GPU suite, small frames
mkdir -p gpurun_out
( timeout -k 10 1500 python -m pytest tests -m gpu -q -x ) > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log | cut -c1-600
python profiles/tools/small_frames.py > gpurun_out/small_frames.log 2>&1
python -c "
import json
for l in open('gpurun_out/small_frames.jsonl'):
    d=json.loads(l); print(d['scene'], d['size'], round(d['reference_kernel_ms'],3), 'wf', round(d['gpt_launch_wavefront_ms'],3), 'mega', round(d['gpt_launch_megakernel_ms'],3), 'default', round(d['gpt_launch_default_ms'],3), d['tree_default'], round(d['reference_kernel_ms']/d['gpt_launch_default_ms'],2))"
