#!/usr/bin/env python
"""Does the extend kernel prefer smaller leaves than the SAH (cost 1/1) chooses?  Device LBVH with max_leaf = 1..8 on the
bench scene and the 100 k dragon, 1080p 16 spp depth 8; reference-built SBVH beside it."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import gpt_b200 as gpt  # noqa: E402
from support import scenes  # noqa: E402

sph = gpt.default_spheres()


def timed(ctx, info, spp=16, reps=3):
    acc = torch.zeros(info.height, info.width, 4, dtype=torch.float32, device="cuda")
    ctx.render(info, acc, 0, 0, spp)
    torch.cuda.synchronize()
    ctx.reset_stats()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for r in range(reps):
        ctx.render(info, acc, 0, (r + 1) * spp, spp)
    e1.record()
    torch.cuda.synchronize()
    kt = ctx.kernel_times()
    return e0.elapsed_time(e1) / reps, kt["extend"][0] / reps, kt["shade"][0] / reps


for name in ("bunny_cornell", "dragon_cornell"):
    sc = scenes.load_scene(name)
    info = gpt.default_kernel_info(1920, 1080)
    info.depth = 8
    with gpt.Context(0, gpt.FLAG_TIME_KERNELS) as ctx:
        ctx.upload_scene(sc["nodes"], sc["tris"], sc["indices"], sph)
        ms, ext, sh = timed(ctx, info)
        print(json.dumps({"scene": name, "tree": "reference SBVH", "ms": ms, "extend_ms": ext, "shade_ms": sh}), flush=True)
        for ml in (1, 2, 3, 4, 6, 8):
            b = ctx.build_bvh(sc["verts"], sc["faces"], max_leaf=ml)
            ctx.upload_built(b, sph)
            ms, ext, sh = timed(ctx, info)
            print(json.dumps({"scene": name, "tree": "LBVH max_leaf=%d" % ml, "inner": b.n_inner_nodes, "leaves": b.n_leaves, "depth": b.depth,
                              "ms": ms, "extend_ms": ext, "shade_ms": sh}), flush=True)
            ctx.free_bvh(b)
