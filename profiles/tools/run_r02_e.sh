#!/bin/bash
# round 2, GPU call E (8 GPUs): weak and strong scaling at N = 1, 2, 4, 8 (16-spp and 256-spp frames), device-timed, max over ranks
mkdir -p gpurun_out
for n in 1 2 4 8; do
  for mode in weak strong; do
    if [ $n -eq 1 ]; then cmd="python bench.py"; else cmd="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + n)) bench.py"; fi
    timeout -k 10 400 $cmd --gpus $n --steps 10 --warmup 3 --scaling $mode --no-cpu-baseline --no-configs > gpurun_out/scale_${mode}_n$n.json 2> gpurun_out/scale_${mode}_n$n.err
  done
  if [ $n -eq 1 ]; then cmd="python bench.py"; else cmd="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29700 + n)) bench.py"; fi
  timeout -k 10 400 $cmd --gpus $n --steps 3 --warmup 3 --scaling strong --frame-spp 256 --no-cpu-baseline --no-configs > gpurun_out/scale_strong256_n$n.json 2> gpurun_out/scale_strong256_n$n.err
done
ls -la gpurun_out
