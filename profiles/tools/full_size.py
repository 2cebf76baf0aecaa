#!/usr/bin/env python
"""BASELINE.json configs 3 and 4 at their FULL sample counts on one B200 (configs_bench.py times a reduced spp): 4K x 256 spp on the
16-instance car scene and 4K x 1024 spp on the 20 M-triangle synthetic scene, rendered as progressive frames of 16 spp with a
resolve + RGBA8 read-back per frame.  Reports wall time for the whole image, rays, Mrays/s.  gpurun_out/full_size.jsonl"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import gpt_b200 as gpt  # noqa: E402
from support import procedural, scenes  # noqa: E402


def run(name, v, f, W, H, spp_total, mat, per_frame=16):
    info = gpt.default_kernel_info(W, H)
    info.depth, info.triCurrentMat = 8, mat
    with gpt.Context(0) as ctx:
        t0 = time.time()
        b = ctx.build_bvh(v, f)
        ctx.upload_built(b, gpt.default_spheres())
        torch.cuda.synchronize()
        t_build = time.time() - t0
        acc = torch.zeros(H, W, 4, dtype=torch.float32, device="cuda")
        rgba = torch.zeros(H, W, dtype=torch.int32, device="cuda")
        host = torch.zeros(H, W, dtype=torch.int32).pin_memory()
        for r in range(7):                                  # warm-up + per-scene trial, not part of the image
            ctx.render(info, acc, 99, 10_000_000 + r * per_frame, per_frame)
        torch.cuda.synchronize()
        acc.zero_(); ctx.reset_stats()
        t0 = time.time()
        for s in range(0, spp_total, per_frame):
            ctx.render(info, acc, 0, s, min(per_frame, spp_total - s))
            ctx.resolve(acc, W, H, rgba, None)
            host.copy_(rgba, non_blocking=True)
        torch.cuda.synchronize()
        dt = time.time() - t0
        st = ctx.stats()
        img = acc[..., :3] / acc[..., 3:4].clamp(min=1)
        d = {"config": name, "triangles": int(len(f)), "width": W, "height": H, "spp": spp_total, "depth": 8, "seconds": dt, "rays": int(st.rays),
             "Mrays_per_s": st.rays / dt / 1e6, "ms_per_16spp_frame": dt * 1e3 / (spp_total / per_frame), "build_and_upload_s": t_build,
             "wide_nodes": ctx.wide_info()[0], "reorder": ctx.reorder_info()[0], "mean_radiance": float(img.mean()), "count_min_max": [float(acc[..., 3].min()), float(acc[..., 3].max())]}
        print(json.dumps(d), flush=True)
        with open(os.path.join(ROOT, "gpurun_out", "full_size.jsonl"), "a") as fh:
            fh.write(json.dumps(d) + "\n")
        ctx.free_bvh(b)


sc = scenes.load_scene("gto_sixteen")
v, f = procedural.instance_grid(sc["verts"], sc["faces"], 4, 4, 0.4, 9.5, -35.0, 15.0)
run("cfg3 gto_sixteen x16 instances, 4K, 256 spp", v, f, 3840, 2160, 256, gpt.METAL)
v, f = procedural.synthetic_scene(20_000_000, seed=1234)
run("cfg4 synthetic 20M, 4K, 1024 spp", v, f, 3840, 2160, 1024, gpt.DIFF)
