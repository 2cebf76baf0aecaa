#!/usr/bin/env python
"""Differential stress of the 4-wide quantised tree against the binary tree (which equals the oracle, tests/test_gpu_parity.py):
random procedural scenes x 2 M random rays each, nearest-hit t and face index compared bit for bit.  A quantised box that failed to
contain its binary box would show as a missed or farther hit.  gpurun_out/wide_stress.jsonl"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import gpt_b200 as gpt  # noqa: E402
from support import procedural, scenes  # noqa: E402


def trace(ctx, o, d, cull):
    n = len(o)
    tri = torch.empty(n, dtype=torch.int32, device="cuda"); t = torch.empty(n, dtype=torch.float32, device="cuda")
    ctx.trace_rays(o, d, n, cull, tri, None, t)
    torch.cuda.synchronize()
    return tri, t


def rays(n, lo, hi, seed, mode):
    g = torch.Generator(device="cuda"); g.manual_seed(seed)
    lo = torch.tensor(lo, device="cuda"); hi = torch.tensor(hi, device="cuda")
    c, ext = (lo + hi) / 2, (hi - lo)
    o = c + (torch.rand(n, 3, device="cuda", generator=g) - 0.5) * ext * (3.0 if mode == 0 else 1.0)
    tgt = c + (torch.rand(n, 3, device="cuda", generator=g) - 0.5) * ext * 1.2
    d = tgt - o
    if mode == 2:                                  # axis-parallel rays: the 2^-80 guard against the quantised planes
        ax = torch.randint(0, 3, (n,), device="cuda", generator=g)
        sgn = torch.where(torch.rand(n, device="cuda", generator=g) < 0.5, -1.0, 1.0)
        d = torch.zeros(n, 3, device="cuda"); d[torch.arange(n), ax] = sgn
    d = d / d.norm(dim=1, keepdim=True)
    return o.float().contiguous(), d.float().contiguous()


N = 2_000_000
cases = [("synthetic", 200_000, s) for s in (1, 2, 3)] + [("synthetic", 3_000_000, 4)]
for name in ("bunny_cornell", "sphere_mesh", "dragon_cornell"):
    cases.append((name, 0, 0))
if len(sys.argv) > 1:
    cases = cases[:int(sys.argv[1])]
for kind, ntri, seed in cases:
    if kind == "synthetic":
        v, f = procedural.synthetic_scene(ntri, seed=seed)
        arrays = None
    else:
        sc = scenes.load_scene(kind); v, f = sc["verts"], sc["faces"]; arrays = sc
    lo, hi = v.min(0).tolist(), v.max(0).tolist()
    res = {}
    for label, flag in (("binary", gpt.FLAG_BINARY_BVH), ("wide", gpt.FLAG_WIDE_BVH)):
        with gpt.Context(0, flag) as ctx:
            if arrays is None:
                b = ctx.build_bvh(v, f); ctx.upload_built(b, None)
            else:
                ctx.upload_scene(arrays["nodes"], arrays["tris"], arrays["indices"], None)
            assert (ctx.wide_info()[0] > 0) == (label == "wide")
            out = []
            for mode in (0, 1, 2):
                for cull in (True, False):
                    o, d = rays(N, lo, hi, 100 * seed + mode, mode)
                    out.append(trace(ctx, o, d, cull))
            res[label] = out
            if arrays is None:
                ctx.free_bvh(b)
    n_t = n_tri = hits = wide_closer = wide_farther = 0
    by_mode = [0, 0, 0]
    for i, ((tb, hb), (tw, hw)) in enumerate(zip(res["binary"], res["wide"])):
        neq = hb.view(torch.int32) != hw.view(torch.int32)
        n_t += int(neq.sum()); n_tri += int((tb != tw).sum()); hits += int((hb < 3e38).sum())
        wide_closer += int((hw < hb).sum()); wide_farther += int((hw > hb).sum())
        by_mode[i // 2] += int(neq.sum())
    d = {"scene": kind, "triangles": int(len(f)), "seed": seed, "rays": 6 * N, "hits": hits, "t_mismatches": n_t, "face_index_mismatches": n_tri,
         "wide_closer": wide_closer, "wide_farther_or_missed": wide_farther, "by_ray_kind": {"from_outside": by_mode[0], "from_inside": by_mode[1], "axis_parallel": by_mode[2]}}
    print(json.dumps(d), flush=True)
    with open(os.path.join(ROOT, "gpurun_out", "wide_stress.jsonl"), "a") as fh:
        fh.write(json.dumps(d) + "\n")
