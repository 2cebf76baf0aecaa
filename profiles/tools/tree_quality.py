#!/usr/bin/env python
"""Tree quality of the device LBVH against the reference's host SBVH on the SAME mesh: SAH cost of both trees (node cost 1,
triangle cost 1, computed from the flattened arrays) and the extend-kernel time of one frame on each (binary tree forced, so only
the tree differs).  One JSON line per (scene, tree).   Usage: python profiles/tools/tree_quality.py [n_tris ...]"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def sah_cost(nodes, tris):
    """nodes: (4 * n_inner, 4) float32 in CudaBVH's compact layout, tris: (n_vec4, 4) with one 0x80000000 terminator per leaf.
    Returns (SAH cost relative to the root box, inner nodes reached, leaves reached, triangle references, depth)."""
    nodes = np.ascontiguousarray(nodes, np.float32).reshape(-1, 4)
    bits = nodes.view(np.int32)
    term = np.flatnonzero(np.ascontiguousarray(tris, np.float32).reshape(-1, 4).view(np.uint32)[:, 0] == 0x80000000)

    def boxes(rows):           # rows: vec4 index of node -> (lo, hi) per child, each (n, 2, 3)
        n0, n1, nz = nodes[rows], nodes[rows + 1], nodes[rows + 2]
        lo = np.stack([np.stack([n0[:, 0], n0[:, 2], nz[:, 0]], 1), np.stack([n1[:, 0], n1[:, 2], nz[:, 2]], 1)], 1)
        hi = np.stack([np.stack([n0[:, 1], n0[:, 3], nz[:, 1]], 1), np.stack([n1[:, 1], n1[:, 3], nz[:, 3]], 1)], 1)
        return lo, hi

    def area(lo, hi):
        d = np.maximum(hi - lo, 0.0).astype(np.float64)
        return 2.0 * (d[..., 0] * d[..., 1] + d[..., 1] * d[..., 2] + d[..., 2] * d[..., 0])

    cur = np.zeros(1, np.int64)
    lo, hi = boxes(cur)
    root_area = float(area(lo.min(1), hi.max(1))[0])
    cost, n_inner, n_leaf, n_ref, depth = 1.0, 0, 0, 0, 0         # the root itself
    while len(cur):
        depth += 1
        n_inner += len(cur)
        lo, hi = boxes(cur)
        a = area(lo, hi) / root_area                                # (n, 2)
        ch = bits[cur + 3][:, :2].astype(np.int64)
        inner = ch >= 0
        cost += float(a[inner].sum())
        leaf_start = ~ch[~inner]
        cnt = (term[np.searchsorted(term, leaf_start)] - leaf_start) // 3
        cost += float((a[~inner] * cnt).sum())
        n_leaf += int((~inner).sum()); n_ref += int(cnt.sum())
        cur = ch[inner] // 16
    return cost, n_inner, n_leaf, n_ref, depth


def main():
    import torch
    import gpt_b200 as gpt
    from support import procedural, reflib, scenes
    sizes = [int(a) for a in sys.argv[1:]] or [100_000, 300_000]
    cases = [("synthetic %d" % n, procedural.synthetic_scene(n, seed=1234)) for n in sizes]
    sc = scenes.load_scene("dragon_cornell")
    cases.append(("dragon_cornell", (sc["verts"], sc["faces"])))
    W, H, spp = 1920, 1080, 4
    info = gpt.default_kernel_info(W, H); info.depth = 8
    sph = gpt.default_spheres()
    for name, (v, f) in cases:
        ref, host_s = None, 0.0
        if len(f) <= 1_100_000 and not os.environ.get("TQ_NO_REF"):
            t0 = time.time()
            ref = reflib.ref_build_bvh(v, f)
            host_s = time.time() - t0
        with gpt.Context(0, gpt.FLAG_TIME_KERNELS | gpt.FLAG_BINARY_BVH | gpt.FLAG_WAVEFRONT) as ctx:
            ctx.set_flags(gpt.FLAG_TIME_KERNELS | gpt.FLAG_BINARY_BVH | gpt.FLAG_WAVEFRONT | gpt.FLAG_BUILD_LBVH)
            b0 = ctx.build_bvh(v, f); ctx.free_bvh(b0); b0 = ctx.build_bvh(v, f)      # second build: allocator warm
            radix = ctx.built_to_host(b0)
            ctx.set_flags(gpt.FLAG_TIME_KERNELS | gpt.FLAG_BINARY_BVH | gpt.FLAG_WAVEFRONT)
            b = ctx.build_bvh(v, f); ctx.free_bvh(b); b = ctx.build_bvh(v, f)
            ploc = ctx.built_to_host(b)
            trees = [("device radix tree", radix, {"build_ms": b0.build_ms}),
                     ("device PLOC r=%s" % os.environ.get("GPT_PLOC_RADIUS", "24"), ploc, {"build_ms": b.build_ms})]
            if ref is not None:
                trees.insert(0, ("reference SBVH (host)", ref, {"build_s": host_s}))
            for tname, arr, extra in trees:
                ctx.upload_scene(arr["nodes"], arr["tris"], arr["indices"], sph)
                acc = torch.zeros(H, W, 4, dtype=torch.float32, device="cuda")
                for w in range(3):
                    ctx.render(info, acc, 0, 1000 + w * spp, spp)
                torch.cuda.synchronize(); ctx.reset_stats()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for r in range(3):
                    ctx.render(info, acc, 0, r * spp, spp)
                e1.record(); torch.cuda.synchronize()
                kt = ctx.kernel_times()
                c, ni, nl, nr, dp = sah_cost(arr["nodes"], arr["tris"])
                print(json.dumps({"scene": name, "triangles": int(len(f)), "tree": tname, "sah_cost": c, "inner": ni, "leaves": nl, "tri_refs": nr,
                                  "depth": dp, "ms_per_frame": e0.elapsed_time(e1) / 3, "extend_ms": kt["extend"][0] / 3,
                                  "Mrays_per_s": ctx.stats().rays / e0.elapsed_time(e1) / 1e3, **extra}), flush=True)
            ctx.free_bvh(b); ctx.free_bvh(b0)


if __name__ == "__main__":
    main()
