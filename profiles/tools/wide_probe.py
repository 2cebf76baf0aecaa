#!/usr/bin/env python
"""One synthetic scene, one traversal structure, a few frames: the command profiled by ncu for the wide-vs-binary question.
usage: wide_probe.py <triangles> <binary|wide> [width height spp]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import gpt_b200 as gpt  # noqa: E402
from support import procedural  # noqa: E402

n = int(sys.argv[1]); kind = sys.argv[2]
W, H, spp = (int(a) for a in sys.argv[3:6]) if len(sys.argv) >= 6 else (1920, 1080, 2)
v, f = procedural.synthetic_scene(n, seed=1234)
info = gpt.default_kernel_info(W, H)
info.depth = 8
with gpt.Context(0, gpt.FLAG_TIME_KERNELS | (gpt.FLAG_WIDE_BVH if kind == "wide" else gpt.FLAG_BINARY_BVH)) as ctx:
    b = ctx.build_bvh(v, f)
    ctx.upload_built(b, gpt.default_spheres())
    print("wide_info", ctx.wide_info())
    acc = torch.zeros(H, W, 4, dtype=torch.float32, device="cuda")
    for r in range(7):                       # warm-up; also lets the per-scene trial (up to 6 calls) finish
        ctx.render(info, acc, 0, 100 + r * spp, spp)
    torch.cuda.synchronize()
    print("reorder_info", ctx.reorder_info(), "trial", ctx.trial_info(), "wide", ctx.wide_info())
    ctx.reset_stats()
    ctx.render(info, acc, 0, spp, spp)
    torch.cuda.synchronize()
    st = ctx.stats(); kt = ctx.kernel_times()
    print(kind, n, "rays", st.rays, "extend ms", kt["extend"][0], "shade ms", kt["shade"][0], "Mrays/s", st.rays / (sum(v[0] for v in kt.values())) / 1e3)
    ctx.free_bvh(b)
