#!/bin/bash
# round 2, GPU call E2 (8 GPUs, charged 8x: kept short): the final build at N = 8 and 4 -- weak and strong scaling with reduce_check --
# and BASELINE configs 3 and 4 at their full sample counts over 8 GPUs (N = 1, 2 are in profiles/r02_call_d/ and the 1-GPU records)
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name,clocks.sm --format=csv > gpurun_out/gpus.txt
run() { n=$1; shift; python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) "$@"; }
timeout -k 10 300 bash -c "$(declare -f run); run 8 bench.py --gpus 8 --steps 10 --warmup 3 --scaling weak --no-cpu-baseline --no-configs" > gpurun_out/scale_weak_n8.json 2> gpurun_out/scale_weak_n8.err
timeout -k 10 300 bash -c "$(declare -f run); run 8 bench.py --gpus 8 --steps 10 --warmup 3 --scaling strong --no-cpu-baseline --no-configs" > gpurun_out/scale_strong_n8.json 2> gpurun_out/scale_strong_n8.err
timeout -k 10 300 bash -c "$(declare -f run); run 8 bench.py --gpus 8 --steps 3 --warmup 3 --scaling strong --frame-spp 256 --no-cpu-baseline --no-configs" > gpurun_out/scale_strong256_n8.json 2> gpurun_out/scale_strong256_n8.err
timeout -k 10 300 bash -c "$(declare -f run); run 4 bench.py --gpus 4 --steps 10 --warmup 3 --scaling weak --no-cpu-baseline --no-configs" > gpurun_out/scale_weak_n4.json 2> gpurun_out/scale_weak_n4.err
timeout -k 10 300 bash -c "$(declare -f run); run 4 bench.py --gpus 4 --steps 10 --warmup 3 --scaling strong --no-cpu-baseline --no-configs" > gpurun_out/scale_strong_n4.json 2> gpurun_out/scale_strong_n4.err
timeout -k 10 900 bash -c "$(declare -f run); run 8 profiles/tools/full_size_multi.py cfg3 cfg4" > gpurun_out/full_size_multi.log 2>&1
ls -la gpurun_out; cat gpurun_out/scale_*_n8.json | cut -c1-400; cat gpurun_out/full_size_multi.jsonl | cut -c1-500
