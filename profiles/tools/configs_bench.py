#!/usr/bin/env python
"""BASELINE.json configs 0..4 on ONE B200 (bench.py carries configs[1] as the headline; the others are reported here).
For every config: our wavefront path (Mrays/s, ms per spp-frame) on (a) the reference-built SBVH arrays where the
host build is feasible (committed / _big fixtures) and (b) the arrays of the device-side LBVH builder, plus the
reference's own CUDA kernel (oracle/_ref/reftrace.cubin) on the same arrays where it exists.  spp is reduced from the
config's figure where stated (throughput per sample does not depend on it).  One JSON line per measurement into
gpurun_out/configs.jsonl.   Usage: python profiles/tools/configs_bench.py [cfg ...]"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import gpt_b200 as gpt  # noqa: E402
from support import procedural, refgpu, scenes  # noqa: E402

OUT = os.path.join(ROOT, "gpurun_out", "configs.jsonl")


def emit(d):
    print(json.dumps(d), flush=True)
    if os.path.isdir(os.path.dirname(OUT)):
        with open(OUT, "a") as fh:
            fh.write(json.dumps(d) + "\n")


def time_ours(ctx, info, spp, reps=3):
    W, H = info.width, info.height
    acc = torch.zeros(H, W, 4, dtype=torch.float32, device="cuda")
    for w in range(10):                        # warm-up: allocates the wave buffers and lets the per-scene trial (<= 9 calls) end
        ctx.render(info, acc, 0, 10_000_000 + w * spp, spp)
    torch.cuda.synchronize()
    ctx.reset_stats()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for r in range(reps):
        ctx.render(info, acc, 0, (r + 1) * spp, spp)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    st = ctx.stats()
    kt = ctx.kernel_times()
    return ms, st.rays / reps, {k: v[0] / reps for k, v in kt.items() if v[1]}


def time_ref_kernel(refk, info, arrays, sph, spp, reps=2):
    if refk is None:
        return None
    i2 = type(info).from_buffer_copy(info)
    rs = refgpu.RefScene(torch, i2, arrays, sph)

    def frames(first):
        for f in range(spp):
            i2.hash = gpt.hash64(first + f)
            i2.constantPdf = f + 1
            refk.launch(i2)
    frames(0)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for r in range(reps):
        frames((r + 1) * spp)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def run(name, W, H, spp, depth, mat, verts, faces, ref_arrays, sph, refk, note, ref_spp=None):
    info = gpt.default_kernel_info(W, H)
    info.depth = depth
    info.triCurrentMat = mat
    with gpt.Context(0, gpt.FLAG_TIME_KERNELS) as ctx:
        trees = []
        if ref_arrays is not None:
            trees.append(("reference SBVH (host, SplitBVHBuilder)", ref_arrays, None))
        t0 = time.time()
        b = ctx.build_bvh(verts, faces)
        wall = time.time() - t0
        trees.append(("device LBVH (gpt_build_bvh)", None, b))
        for tname, arr, built in trees:
            if built is not None:
                ctx.upload_built(built, sph)
                sizes = (built.bvh.triNodeSize * 16, built.bvh.triSize * 16, built.bvh.triIndicesSize * 4)
            else:
                ctx.upload_scene(arr["nodes"], arr["tris"], arr["indices"], sph)
                sizes = (arr["nodes"].nbytes, arr["tris"].nbytes, arr["indices"].nbytes)
            ms, rays, kt = time_ours(ctx, info, spp)
            d = {"config": name, "tree": tname, "triangles": int(len(faces)), "width": W, "height": H, "spp": spp, "depth": depth,
                 "ms_per_frame": ms, "Mrays_per_s": rays / ms / 1e3, "rays_per_frame": rays, "kernel_class_ms": kt,
                 "scene_bytes": {"nodes": int(sizes[0]), "tris": int(sizes[1]), "indices": int(sizes[2])}, "note": note}
            if built is not None:
                d["build_ms_device"] = built.build_ms
                d["build_s_wall_incl_h2d"] = wall
                d["inner_nodes"], d["leaves"], d["tree_depth"] = built.n_inner_nodes, built.n_leaves, built.depth
            rspp = ref_spp or spp
            if refk is not None and rspp > 0:
                if built is not None:
                    arr = ctx.built_to_host(built)
                rms = time_ref_kernel(refk, info, arr, sph, rspp)
                d["reference_gpu_kernel_ms_per_frame"] = rms * spp / rspp
                d["speedup_vs_reference_gpu_kernel"] = rms * spp / rspp / ms
            emit(d)
        ctx.free_bvh(b)


def main():
    which = set(sys.argv[1:]) or {"cfg0", "cfg1", "cfg2", "cfg3", "cfg4"}
    refk = refgpu.RefGpuKernel(0) if refgpu.available() else None
    sph = gpt.default_spheres()
    items = [(list(s.posRad), list(s.emi), list(s.col), s.mat) for s in sph]
    items.append(([-9.0, -9.0, -33.0, 5.0], [0.0, 0.0, 0.0], [0.9, 0.8, 0.7], gpt.METAL))
    sph9 = gpt.spheres_from_list(items)
    if "cfg0" in which:
        sc = scenes.load_scene("cornell")
        run("cfg0 cornell.obj 512x512 1spp depth4 diffuse", 512, 512, 1, 4, gpt.DIFF, sc["verts"], sc["faces"], sc, sph, refk, "as BASELINE configs[0]")
    if "cfg1" in which:
        sc = scenes.load_scene("bunny_cornell")
        run("cfg1 bunny+cornell 1080p 16spp depth8", 1920, 1080, 16, 8, gpt.DIFF, sc["verts"], sc["faces"], sc, sph9, refk, "bench.py workload")
    if "cfg2" in which:
        sc = scenes.load_scene("dragon_cornell")
        run("cfg2 dragon(100k)+cornell 1080p 64spp depth8", 1920, 1080, 16, 8, gpt.DIFF, sc["verts"], sc["faces"], sc, sph, refk,
            "cornell_dragon.obj is a missing blob: Assets/dragon.obj (100 000 tris) + cornell.obj; 16 of 64 spp timed", ref_spp=4)
        v, f = procedural.subdivide(sc["verts"], sc["faces"], 2)
        run("cfg2 dragon subdivided x16 (1.6M tris) 1080p depth8", 1920, 1080, 16, 8, gpt.DIFF, v, f, None, sph, refk,
            "two 1->4 midpoint subdivisions stand in for the missing 870k-triangle dragon; host SBVH build infeasible here (~50 s), LBVH only", ref_spp=4)
    if "cfg3" in which:
        sc = scenes.load_scene("gto_sixteen")
        v, f = procedural.instance_grid(sc["verts"], sc["faces"], 4, 4, 0.4, 9.5, -35.0, 15.0)
        run("cfg3 gto_sixteen x16 instances (157k tris) 4K depth8", 3840, 2160, 8, 8, gpt.METAL, v, f, None, sph, refk,
            "4x4 grid (x by depth) of the decimated car scaled 0.4 to fit the 40-wide room (vertex duplication before the build); 8 of 256 spp timed on one GPU", ref_spp=2)
    if "cfg4" in which:
        t0 = time.time()
        v, f = procedural.synthetic_scene(20_000_000, seed=1234)
        gen = time.time() - t0
        run("cfg4 synthetic 20M tris 4K depth8", 3840, 2160, 4, 8, gpt.DIFF, v, f, None, sph, refk,
            "icospheres + boxes, seed 1234 (generated in %.1f s); 4 of 1024 spp timed on one GPU; host SBVH est. 15-25 min => LBVH only" % gen, ref_spp=1)
    if refk is not None:
        refk.close()


if __name__ == "__main__":
    main()
