#!/usr/bin/env python
"""Drop-in regime: ONE 1-spp launchKernel call per displayed frame (BasicScene.cpp:395-404), small frames.  Times
gpt_launch (wavefront and megakernel schedules) against the reference's own kernel on the same kernelInfo."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import gpt_b200 as gpt  # noqa: E402
from support import refgpu, scenes  # noqa: E402

refk = refgpu.RefGpuKernel(0) if refgpu.available() else None
sph = gpt.default_spheres()
for name, W, H, depth in [("cornell", 512, 512, 4), ("dragon_cornell", 1280, 720, 4), ("bunny_cornell", 1920, 1080, 8), ("bunny_cornell", 256, 144, 4),
                          ("bunny_cornell", 640, 360, 8), ("bunny_cornell", 800, 600, 8), ("bunny_cornell", 1024, 576, 8), ("dragon_cornell", 640, 480, 4)]:
    sc = scenes.load_scene(name)
    info = gpt.default_kernel_info(W, H)
    info.depth = depth
    rs = refgpu.RefScene(torch, info, sc, sph)
    res = {"scene": name, "size": [W, H], "depth": depth}

    def loop(fn, n=200):
        for f in range(20):
            info.hash, info.constantPdf = gpt.hash64(f), f + 1
            fn(info)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for f in range(n):
            info.hash, info.constantPdf = gpt.hash64(20 + f), 21 + f
            fn(info)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n
    if refk is not None:
        res["reference_kernel_ms"] = loop(refk.launch)
    for label, flags in (("wavefront", gpt.FLAG_WAVEFRONT), ("megakernel", gpt.FLAG_MEGAKERNEL), ("default", 0)):
        with gpt.Context(0, flags) as ctx:
            res["gpt_launch_%s_ms" % label] = loop(ctx.launch)
            res["tree_%s" % label] = ctx.tree_info()["in_use"] if W * H >= 393216 or label == "wavefront" else ctx.megakernel_tree_info()["in_use"]
    print(json.dumps(res), flush=True)
    out = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(out):
        open(os.path.join(out, "small_frames.jsonl"), "a").write(json.dumps(res) + "\n")
