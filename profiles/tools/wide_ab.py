#!/usr/bin/env python
"""A/B of the traversal structure on ONE B200: binary tree as uploaded (GPT_FLAG_BINARY_BVH) versus the 4-wide quantised tree
collapsed from it on the device (GPT_FLAG_WIDE_BVH, gpt_wide.cu), same frames, same images.  One JSON line per (scene, tree)
into gpurun_out/wide_ab.jsonl.   Usage: python profiles/tools/wide_ab.py [bunny dragon dragon16 gto16 syn20m ...]"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import gpt_b200 as gpt  # noqa: E402
from support import procedural, scenes  # noqa: E402
from configs_bench import time_ours  # noqa: E402

OUT = os.path.join(ROOT, "gpurun_out", "wide_ab.jsonl")


def emit(d):
    print(json.dumps(d), flush=True)
    if os.path.isdir(os.path.dirname(OUT)):
        with open(OUT, "a") as fh:
            fh.write(json.dumps(d) + "\n")


def run(name, W, H, spp, depth, mat, verts, faces, ref_arrays, sph):
    info = gpt.default_kernel_info(W, H)
    info.depth, info.triCurrentMat = depth, mat
    images = {}
    for tree in (["reference SBVH"] if ref_arrays is not None else []) + ["device LBVH"]:
        for label, flag in (("binary", gpt.FLAG_BINARY_BVH), ("wide4q", gpt.FLAG_WIDE_BVH)):
            with gpt.Context(0, gpt.FLAG_TIME_KERNELS | flag) as ctx:
                b = None
                if tree == "device LBVH":
                    b = ctx.build_bvh(verts, faces)
                    ctx.upload_built(b, sph)
                else:
                    ctx.upload_scene(ref_arrays["nodes"], ref_arrays["tris"], ref_arrays["indices"], sph)
                t0 = time.time()
                n_wide, levels = ctx.wide_info()
                torch.cuda.synchronize()
                wide_build_s = time.time() - t0
                ms, rays, kt = time_ours(ctx, info, spp)
                acc = torch.zeros(H, W, 4, dtype=torch.float32, device="cuda")
                ctx.render(info, acc, 0, 0, 1)
                torch.cuda.synchronize()
                images[(tree, label)] = acc.cpu().numpy()
                d = {"scene": name, "tree": tree, "traversal": label, "triangles": int(len(faces)), "width": W, "height": H, "spp": spp,
                     "depth": depth, "ms_per_frame": ms, "Mrays_per_s": rays / ms / 1e3, "extend_ms": kt.get("extend"), "shade_ms": kt.get("shade"),
                     "wide_nodes": n_wide, "wide_levels": levels, "wide_build_s_incl_sync": wide_build_s if n_wide else None}
                if b is not None:
                    d["binary_inner_nodes"], d["binary_depth"] = b.n_inner_nodes, b.depth
                    ctx.free_bvh(b)
                if label == "wide4q":
                    d["pixels_differing_from_binary"] = float(np.any(images[(tree, "binary")] != images[(tree, "wide4q")], axis=-1).mean())
                emit(d)


def main():
    which = set(sys.argv[1:]) or {"bunny", "dragon", "dragon16", "gto16", "syn20m"}
    sph = gpt.default_spheres()
    items = [(list(s.posRad), list(s.emi), list(s.col), s.mat) for s in sph]
    items.append(([-9.0, -9.0, -33.0, 5.0], [0.0, 0.0, 0.0], [0.9, 0.8, 0.7], gpt.METAL))
    sph9 = gpt.spheres_from_list(items)
    if "bunny" in which:
        sc = scenes.load_scene("bunny_cornell")
        run("bunny+cornell 5k (bench.py workload)", 1920, 1080, 16, 8, gpt.DIFF, sc["verts"], sc["faces"], sc, sph9)
    if "dragon" in which or "dragon16" in which:
        sc = scenes.load_scene("dragon_cornell")
        if "dragon" in which:
            run("dragon 100k + cornell", 1920, 1080, 16, 8, gpt.DIFF, sc["verts"], sc["faces"], sc, sph)
        if "dragon16" in which:
            v, f = procedural.subdivide(sc["verts"], sc["faces"], 2)
            run("dragon subdivided x16 (1.6M)", 1920, 1080, 16, 8, gpt.DIFF, v, f, None, sph)
    if "gto16" in which:
        sc = scenes.load_scene("gto_sixteen")
        v, f = procedural.instance_grid(sc["verts"], sc["faces"], 4, 4, 0.4, 9.5, -35.0, 15.0)
        run("gto_sixteen x16 instances (157k), 4K", 3840, 2160, 8, 8, gpt.METAL, v, f, None, sph)
    if "syn20m" in which:
        v, f = procedural.synthetic_scene(20_000_000, seed=1234)
        run("synthetic 20M, 4K", 3840, 2160, 4, 8, gpt.DIFF, v, f, None, sph)


if __name__ == "__main__":
    main()
