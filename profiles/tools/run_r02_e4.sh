#!/bin/bash
# round 2, GPU call E4 (8 GPUs, kept short): scaling of the final build at N = 8 and 4 (weak, strong, strong at 256 spp)
mkdir -p gpurun_out
run() { n=$1; shift; python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) "$@"; }
timeout -k 10 300 bash -c "$(declare -f run); run 8 bench.py --gpus 8 --steps 10 --warmup 3 --scaling weak --no-cpu-baseline --no-configs" > gpurun_out/scale_weak_n8.json 2> gpurun_out/scale_weak_n8.err
timeout -k 10 300 bash -c "$(declare -f run); run 8 bench.py --gpus 8 --steps 10 --warmup 3 --scaling strong --no-cpu-baseline --no-configs" > gpurun_out/scale_strong_n8.json 2> gpurun_out/scale_strong_n8.err
timeout -k 10 300 bash -c "$(declare -f run); run 8 bench.py --gpus 8 --steps 3 --warmup 3 --scaling strong --frame-spp 256 --no-cpu-baseline --no-configs" > gpurun_out/scale_strong256_n8.json 2> gpurun_out/scale_strong256_n8.err
timeout -k 10 300 bash -c "$(declare -f run); run 4 bench.py --gpus 4 --steps 10 --warmup 3 --scaling weak --no-cpu-baseline --no-configs" > gpurun_out/scale_weak_n4.json 2> gpurun_out/scale_weak_n4.err
timeout -k 10 300 bash -c "$(declare -f run); run 4 bench.py --gpus 4 --steps 10 --warmup 3 --scaling strong --no-cpu-baseline --no-configs" > gpurun_out/scale_strong_n4.json 2> gpurun_out/scale_strong_n4.err
for f in gpurun_out/scale_*.json; do python -c "
import json
d=json.loads(open('$f').read().strip().splitlines()[-1]); rc=d.get('reduce_check') or {}
print('$f', round(d['value']), round(d['ms_per_step'],3), d['run']['tree']['in_use'], d['run']['tree']['ranks_agreed'], rc.get('max_abs_diff'), rc.get('tolerance'), rc.get('count_plane_exact'))" || tail -3 ${f%.json}.err; done
