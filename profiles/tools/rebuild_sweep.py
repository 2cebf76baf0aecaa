#!/usr/bin/env python
"""ms per 1080p x 16 spp x depth 8 frame on the REBUILT tree of two fixture scenes (reference SBVH arrays in, gpt_rebuild_info), for
sweeps of the builder's environment knobs (one process per setting: they are read once).  Run with GPT_REBUILD=1."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import gpt_b200 as gpt  # noqa: E402
from support import scenes  # noqa: E402

out = {"env": {k: v for k, v in os.environ.items() if k.startswith("GPT_")}}
for name in sys.argv[1:] or ["bunny_cornell", "dragon_cornell"]:
    sc = scenes.load_scene(name)
    info = gpt.default_kernel_info(1920, 1080)
    info.depth = 8
    with gpt.Context(0, gpt.FLAG_TIME_KERNELS) as ctx:
        ctx.upload_scene(sc["nodes"], sc["tris"], sc["indices"], gpt.default_spheres())
        acc = torch.zeros(1080, 1920, 4, dtype=torch.float32, device="cuda")
        for w in range(4):
            ctx.render(info, acc, 0, 1000 + 16 * w, 16)
        torch.cuda.synchronize()
        ctx.reset_stats()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for r in range(4):
            ctx.render(info, acc, 0, 16 * r, 16)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 4
        kt = ctx.kernel_times()
        out[name] = {"ms": ms, "Mrays_s": ctx.stats().rays / 4 / ms / 1e3, "extend_ms": kt["extend"][0] / 4, "rebuilt": ctx.rebuild_info()}
print(json.dumps(out))
