#!/bin/bash
# round 2, GPU call D (2 GPUs): the 2-GPU test, bench weak + strong scaling with reduce_check
mkdir -p gpurun_out
( timeout -k 10 600 python -m pytest tests/test_gpu_multigpu.py -q ) > gpurun_out/pytest_multigpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_multigpu.log
for mode in weak strong; do
  timeout -k 10 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 --scaling $mode > gpurun_out/bench_n2_$mode.json 2> gpurun_out/bench_n2_$mode.err; echo "rc=$?" >> gpurun_out/bench_n2_$mode.err
done
timeout -k 10 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 5 --warmup 3 --scaling strong --frame-spp 256 > gpurun_out/bench_n2_strong256.json 2> gpurun_out/bench_n2_strong256.err
timeout -k 10 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-configs > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err
ls -la gpurun_out
