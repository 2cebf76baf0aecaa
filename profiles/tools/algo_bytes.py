#!/usr/bin/env python
"""Algorithmic bytes per ray (64*N_inner + 48*N_tri + 16*N_leaf + 4*[tri hit] + 44*N_sphere) of the bench.py
workload and of the scenes of BASELINE configs 2-4, counted by the CPU oracle executing the REFERENCE algorithm
(cudaUtils.h:256-460) on the arrays each scene is traversed on (the reference SBVH fixture, or the device builder's arrays copied
back) -- a 480x270, 1 spp, depth 8 sample at the full frame's framing.  -> profiles/r02_algorithmic_bytes.json (bench.py reads it).
Needs a GPU for the device-built trees:  python profiles/tools/algo_bytes.py > gpurun_out/algo_bytes.json"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import gpt_b200 as gpt  # noqa: E402
from support import config_check, procedural, scenes  # noqa: E402

W, H = 480, 270


def count(name, host, sph, mat, full_h):
    info = gpt.default_kernel_info(W, H)
    info.depth = 8
    info.triCurrentMat = mat
    info.cam.dist = float(full_h // 60) * H / full_h          # BasicScene.cpp:241 at the full size, scaled to the sample's height
    _, c = scenes.oracle_render(host, sph, info, 0, 0, 1, counters=True, threads=os.cpu_count() or 8)
    r = float(c.rays)
    trav = 64 * c.inner_nodes + 48 * c.tri_tests + 16 * c.leaf_terms + 4 * c.tri_hits
    return {"rays_sampled": int(c.rays), "inner_nodes_per_ray": c.inner_nodes / r, "tri_tests_per_ray": c.tri_tests / r,
            "leaf_terminators_per_ray": c.leaf_terms / r, "tri_hits_per_ray": c.tri_hits / r, "sphere_tests_per_ray": c.sphere_tests / r,
            "max_stack": int(c.max_stack), "traversal_bytes_per_ray": trav / r, "bytes_per_ray": c.algorithmic_bytes() / r,
            "how": "oracle/liboracle.so orc_render, %dx%d, 1 spp, seed 0, depth 8, framing of the full-size frame" % (W, H)}


def main():
    import bench
    out = {}
    sc = bench.load_scene_arrays()
    out["bench"] = count("bench", sc, bench.bench_spheres(gpt), gpt.DIFF, 1080)
    names = sys.argv[1:] or ["cfg2_dragon100k_sbvh", "cfg2_dragon1p6M", "cfg3_gto16", "cfg4_20M"]
    ctx = None
    for name in names:
        verts, faces, ref, fw, fh, mat = config_check.make_scene(name, gpt, procedural, scenes)
        if ref is None:
            if ctx is None:
                ctx = gpt.Context(0)
            b = ctx.build_bvh(verts, faces)
            host = dict(ctx.built_to_host(b), verts=verts, faces=faces)
            ctx.free_bvh(b)
        else:
            host = ref
        out[name] = dict(count(name, host, gpt.default_spheres(), mat, fh), triangles=int(len(faces)),
                         tree="reference SBVH" if ref is not None else "device-built (gpt_build_bvh)")
    json.dump(out, sys.stdout, indent=1)
    print()


if __name__ == "__main__":
    main()
