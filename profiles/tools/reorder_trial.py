#!/usr/bin/env python
"""What the per-scene reordering trial decides (gpt_reorder_info) and what the frame rate is afterwards, on scenes of both kinds:
compact object in a box (dragon x16) and space-filling synthetic scenes.  gpurun_out/reorder_trial.jsonl"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import gpt_b200 as gpt  # noqa: E402
from support import procedural, scenes  # noqa: E402
from configs_bench import time_ours  # noqa: E402


def run(name, v, f, W, H, spp):
    info = gpt.default_kernel_info(W, H)
    info.depth = 8
    with gpt.Context(0, gpt.FLAG_TIME_KERNELS) as ctx:
        b = ctx.build_bvh(v, f)
        ctx.upload_built(b, gpt.default_spheres())
        acc = torch.zeros(H, W, 4, dtype=torch.float32, device="cuda")
        states = []
        for r in range(7):
            ctx.render(info, acc, 0, 1000 + r * spp, spp)
            states.append(int(ctx.trial_info()[0]))
        st, off, on = ctx.reorder_info()
        ms, rays, kt = time_ours(ctx, info, spp)
        d = {"scene": name, "triangles": int(len(f)), "trial_pending_after_each_call": states, "reorder": st, "trial_ns_per_ray": ctx.trial_info()[1], "wide_nodes": ctx.wide_info()[0], "Mrays_per_s": rays / ms / 1e3, "extend_ms": kt.get("extend"), "other_ms": kt.get("other")}
        print(json.dumps(d), flush=True)
        with open(os.path.join(ROOT, "gpurun_out", "reorder_trial.jsonl"), "a") as fh:
            fh.write(json.dumps(d) + "\n")
        ctx.free_bvh(b)


if len(sys.argv) > 1 and sys.argv[1] == "syn20m":
    v, f = procedural.synthetic_scene(20_000_000, seed=1234)
    run("synthetic 20M, 4K", v, f, 3840, 2160, 2)
    sys.exit(0)
sc = scenes.load_scene("dragon_cornell")
run("dragon 100k + cornell", sc["verts"], sc["faces"], 1920, 1080, 8)
v, f = procedural.subdivide(sc["verts"], sc["faces"], 1)
run("dragon x4 (400k)", v, f, 1920, 1080, 8)
v, f = procedural.subdivide(sc["verts"], sc["faces"], 2)
run("dragon x16 (1.6M)", v, f, 1920, 1080, 8)
for n in (500_000, 2_000_000, 8_000_000):
    v, f = procedural.synthetic_scene(n, seed=1234)
    run("synthetic %.1fM" % (n / 1e6), v, f, 1920, 1080, 4)
