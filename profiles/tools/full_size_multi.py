#!/usr/bin/env python
"""BASELINE.json configs 3 and 4 as they are specified -- FULL sample counts, spp-partitioned over the GPUs of one box: 4K x 256 spp
on the 16-instance car scene and 4K x 1024 spp on the 20 M-triangle synthetic scene.  Every rank builds the (deterministic) tree
itself, renders 16 spp of each 16 x N-spp progressive frame into its IPC frame buffer, and the root sums the peers' frames over
NVLink, accumulates, resolves and reads the RGBA8 frame back (multigpu.PeerFrames).  Timed on the device, max over ranks.
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 profiles/tools/full_size_multi.py
-> gpurun_out/full_size_multi.jsonl (rank 0)"""
import json
import os
import sys
import time
from importlib import import_module

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import gpt_b200 as gpt  # noqa: E402
from support import procedural, scenes  # noqa: E402
mg = import_module("gpt_b200.multigpu")

world = int(os.environ.get("WORLD_SIZE", "1"))
rank = int(os.environ.get("RANK", "0"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
D = dist if world > 1 else None


def sync_all():
    torch.cuda.synchronize(dev)
    if D is not None:
        D.barrier()
        torch.cuda.synchronize(dev)


def run(name, v, f, W, H, spp_total, mat, per_gpu=16):
    info = gpt.default_kernel_info(W, H)
    info.depth, info.triCurrentMat = 8, mat
    stream = torch.cuda.current_stream(dev)
    with gpt.Context(local) as ctx:
        t0 = time.time()
        b = ctx.build_bvh(v, f)
        ctx.upload_built(b, gpt.default_spheres())
        torch.cuda.synchronize(dev)
        t_build = time.time() - t0
        scratch = torch.zeros(H, W, 4, dtype=torch.float32, device=dev)
        for r in range(10):                                 # warm-up + per-scene trial (up to nine calls), not part of the image
            ctx.render(info, scratch, 99, 10_000_000 + r * per_gpu, per_gpu, stream.cuda_stream)
        del scratch
        total = torch.zeros(H, W, 4, dtype=torch.float32, device=dev) if rank == 0 else None
        rgba = torch.zeros(H, W, dtype=torch.int32, device=dev) if rank == 0 else None
        host = torch.zeros(H, W, dtype=torch.int32).pin_memory() if rank == 0 else None
        peer = mg.PeerFrames(ctx, W, H, dist=D, device=dev)

        def render_ptr(ptr, begin, count):
            ctx.render(info, ptr, 0, begin, count, stream.cuda_stream)

        per_frame = per_gpu * world
        sync_all()
        ctx.reset_stats()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        frames = 0
        for s in range(0, spp_total, per_frame):
            peer.frame(render_ptr, s, min(per_frame, spp_total - s), total, rgba, None, stream)
            if rank == 0:
                peer.wait_reduced(stream)          # the reduce runs on the root's side stream
                host.copy_(rgba, non_blocking=True)
            frames += 1
        e1.record(stream)
        sync_all()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        rays = torch.tensor([float(ctx.stats().rays)], dtype=torch.float64, device=dev)
        if D is not None:
            D.all_reduce(ms, op=D.ReduceOp.MAX)
            D.all_reduce(rays)
        if rank == 0:
            dt = float(ms.item()) * 1e-3
            img = total[..., :3] / total[..., 3:4].clamp(min=1)
            d = {"config": name, "n_gpus": world, "triangles": int(len(f)), "width": W, "height": H, "spp": spp_total, "depth": 8, "seconds": dt,
                 "rays": float(rays.item()), "Mrays_per_s": float(rays.item()) / dt / 1e6, "progressive_frames": frames, "spp_per_gpu_per_frame": per_gpu,
                 "ms_per_progressive_frame": dt * 1e3 / frames, "build_and_upload_s": t_build, "wide_nodes": ctx.wide_info()[0], "reorder": ctx.reorder_info()[0],
                 "mean_radiance": float(img.mean()), "count_min_max": [float(total[..., 3].min()), float(total[..., 3].max())],
                 "reduce": "fused peer-memory reduce + accumulate + resolve on the root (CUDA IPC over NVLink)" if world > 1 else "none"}
            print(json.dumps(d), flush=True)
            os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
            with open(os.path.join(ROOT, "gpurun_out", "full_size_multi.jsonl"), "a") as fh:
                fh.write(json.dumps(d) + "\n")
        peer.close()
        ctx.free_bvh(b)


which = set(sys.argv[1:]) or {"cfg3", "cfg4"}
if "cfg3" in which:
    sc = scenes.load_scene("gto_sixteen")
    v, f = procedural.instance_grid(sc["verts"], sc["faces"], 4, 4, 0.4, 9.5, -35.0, 15.0)
    run("cfg3 gto_sixteen x16 instances, 4K, 256 spp", v, f, 3840, 2160, 256, gpt.METAL)
if "cfg4" in which:
    v, f = procedural.synthetic_scene(20_000_000, seed=1234)
    run("cfg4 synthetic 20M, 4K, 1024 spp", v, f, 3840, 2160, 1024, gpt.DIFF)
if D is not None:
    D.barrier()
    D.destroy_process_group()
