#!/usr/bin/env python
"""shade / extend ms per 1080p x 16 spp x depth 8 frame of the bench mesh, compat and GPT_FLAG_FAST arithmetic (A/B of library builds)"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import gpt_b200 as gpt  # noqa: E402
from support import scenes  # noqa: E402

sc = scenes.load_scene(sys.argv[1] if len(sys.argv) > 1 else "bunny_cornell")
info = gpt.default_kernel_info(1920, 1080)
info.depth = 8
out = {}
for label, flags in (("compat", 0), ("fast", gpt.FLAG_FAST)):
    with gpt.Context(0, gpt.FLAG_TIME_KERNELS | flags) as ctx:
        ctx.upload_scene(sc["nodes"], sc["tris"], sc["indices"], gpt.default_spheres())
        acc = torch.zeros(1080, 1920, 4, dtype=torch.float32, device="cuda")
        for w in range(4):
            ctx.render(info, acc, 0, 1000 + 16 * w, 16)
        torch.cuda.synchronize()
        ctx.reset_stats()
        for r in range(6):
            ctx.render(info, acc, 0, 16 * r, 16)
        torch.cuda.synchronize()
        kt = ctx.kernel_times()
        out[label] = {"extend_ms": kt["extend"][0] / 6, "shade_ms": kt["shade"][0] / 6}
print(json.dumps(out))
