#!/usr/bin/env python
"""Per-ray traversal statistics (inner-node visits, triangle tests, leaves; oracle counters on the arrays the device builder
emits) for the scenes of wide_ab.py -- what the extend kernel's work per ray consists of.  gpurun_out/ray_stats.jsonl"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import gpt_b200 as gpt  # noqa: E402
from support import procedural, scenes  # noqa: E402

OUT = os.path.join(ROOT, "gpurun_out", "ray_stats.jsonl")


def stats(name, verts, faces, arrays, W=96, H=54, depth=8):
    info = gpt.default_kernel_info(W, H)
    info.depth = depth
    info.cam.dist = 18.0 * H / 1080.0          # the 1080p framing
    sc = dict(verts=verts, faces=faces, **arrays)
    _, c = scenes.oracle_render(sc, gpt.default_spheres(), info, 1, 0, 1, counters=True, threads=16)
    r = float(c.rays)
    d = {"scene": name, "triangles": int(len(faces)), "rays": int(c.rays), "inner_nodes_per_ray": c.inner_nodes / r, "tri_tests_per_ray": c.tri_tests / r,
         "leaves_per_ray": c.leaf_terms / r, "tri_hit_fraction": c.tri_hits / r, "max_stack": int(c.max_stack),
         "algorithmic_bytes_per_ray": c.algorithmic_bytes() / r}
    print(json.dumps(d), flush=True)
    with open(OUT, "a") as fh:
        fh.write(json.dumps(d) + "\n")


def built(ctx, verts, faces):
    b = ctx.build_bvh(verts, faces)
    arr = ctx.built_to_host(b)
    ctx.free_bvh(b)
    return arr


with gpt.Context(0) as ctx:
    sc = scenes.load_scene("bunny_cornell")
    stats("bunny+cornell, reference SBVH", sc["verts"], sc["faces"], {k: sc[k] for k in ("nodes", "tris", "indices")})
    stats("bunny+cornell, device LBVH", sc["verts"], sc["faces"], built(ctx, sc["verts"], sc["faces"]))
    sc = scenes.load_scene("dragon_cornell")
    stats("dragon 100k, reference SBVH", sc["verts"], sc["faces"], {k: sc[k] for k in ("nodes", "tris", "indices")})
    stats("dragon 100k, device LBVH", sc["verts"], sc["faces"], built(ctx, sc["verts"], sc["faces"]))
    for n in (2_000_000, 20_000_000):
        v, f = procedural.synthetic_scene(n, seed=1234)
        stats("synthetic %dM, device LBVH" % (n // 1_000_000), v, f, built(ctx, v, f))
