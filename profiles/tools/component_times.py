#!/usr/bin/env python
"""Times gpt_render on variants of the bench workload to split the cost of a ray into its parts
(development tool; run on the GPU box):  spheres only / mesh+spheres, depth 1 / depth 8."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch
import bench
import gpt_b200 as gpt

class A: width, height, depth = 1920, 1080, 8

sc = bench.load_scene_arrays()
sph = bench.bench_spheres(gpt)

def run(label, mesh, depth, spp=16, reps=3):
    ctx = gpt.Context(0)
    if mesh:
        ctx.upload_scene(sc["nodes"], sc["tris"], sc["indices"], sph)
    else:
        ctx.upload_scene(None, None, None, sph)
    info = gpt.default_kernel_info(A.width, A.height); info.depth = depth
    acc = torch.zeros(A.height, A.width, 4, device="cuda")
    ctx.render(info, acc, 0, 0, spp); torch.cuda.synchronize()
    ctx.reset_stats()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(reps):
        ctx.render(info, acc, 0, i * spp, spp)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    rays = ctx.stats().rays / reps
    print("%-28s %8.2f ms  %7.1f Mrays  %8.0f Mrays/s  %.3f ns/ray" % (label, ms, rays / 1e6, rays / ms / 1e3, ms * 1e6 / rays))
    ctx.close()

run("spheres only, depth 8", False, 8)
run("spheres only, depth 1", False, 1)
run("mesh+spheres, depth 1", True, 1)
run("mesh+spheres, depth 2", True, 2)
run("mesh+spheres, depth 8", True, 8)
