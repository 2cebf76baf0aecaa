#!/bin/bash
# round 2, GPU call N (2 GPUs): the multi-GPU path of the final build with the rebuilt tree -- every rank runs the trial before the
# timed region, reduce_check compares the peer-reduced frame with rank 0 alone -- weak, strong, and the 2-GPU tests
mkdir -p gpurun_out
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) "$@"; }
for mode in weak strong; do
  timeout -k 10 300 bash -c "$(declare -f run); run bench.py --gpus 2 --steps 10 --warmup 3 --scaling $mode --no-cpu-baseline --no-configs" > gpurun_out/scale_${mode}_n2.json 2> gpurun_out/scale_${mode}_n2.err
done
timeout -k 10 300 bash -c "$(declare -f run); run bench.py --gpus 2 --steps 3 --warmup 3 --scaling strong --frame-spp 256 --no-cpu-baseline --no-configs" > gpurun_out/scale_strong256_n2.json 2> gpurun_out/scale_strong256_n2.err
( timeout -k 10 600 python -m pytest tests/test_gpu_multigpu.py -m gpu -q ) > gpurun_out/pytest_multigpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_multigpu.log
tail -3 gpurun_out/pytest_multigpu.log
for f in gpurun_out/scale_*_n2.json; do python -c "
import json
d=json.loads(open('$f').read().strip().splitlines()[-1]); print('$f', round(d['value']), round(d['ms_per_step'],3), d['run']['tree'], d['reduce_check'])" || tail -5 ${f%.json}.err; done
