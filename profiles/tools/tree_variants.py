#!/usr/bin/env python
"""Experiment: the extend kernel on trees the REFERENCE's SplitBVHBuilder makes with other Platform parameters (leaf-size cap, SAH
triangle cost; built on the CPU into variants_data/*.npz), traversed exactly as uploaded, against the same scene's default SBVH and
this library's rebuilt tree.  1080p x 16 spp x depth 8, ms per frame."""
import glob
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import gpt_b200 as gpt  # noqa: E402

info = gpt.default_kernel_info(1920, 1080)
info.depth = 8


def time_tree(arr, flags):
    with gpt.Context(0, gpt.FLAG_TIME_KERNELS | flags) as ctx:
        ctx.upload_scene(arr["nodes"], arr["tris"], arr["indices"], gpt.default_spheres())
        acc = torch.zeros(1080, 1920, 4, dtype=torch.float32, device="cuda")
        for w in range(4):
            ctx.render(info, acc, 0, 1000 + 16 * w, 16)
        torch.cuda.synchronize()
        ctx.reset_stats()
        for r in range(4):
            ctx.render(info, acc, 0, 16 * r, 16)
        torch.cuda.synchronize()
        kt = ctx.kernel_times()
        return {"extend_ms": kt["extend"][0] / 4, "shade_ms": kt["shade"][0] / 4, "rebuilt_in_use": ctx.rebuild_info()["state"], "mean": float(acc[..., :3].mean())}


for f in sorted(glob.glob(os.path.join(ROOT, "variants_data", "*.npz"))):
    z = np.load(f)
    arr = {"nodes": z["nodes"].reshape(-1, 4), "tris": z["tris"].reshape(-1, 4), "indices": z["indices"]}
    out = {"tree": os.path.basename(f), "inner_nodes": len(arr["nodes"]) // 4, "as_uploaded": time_tree(arr, gpt.FLAG_KEEP_TREE)}
    if f.endswith("_base.npz"):
        out["rebuilt"] = time_tree(arr, 0)          # run with GPT_REBUILD=1
    print(json.dumps(out), flush=True)
