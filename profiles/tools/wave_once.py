#!/usr/bin/env python
"""Renders the bench frame with ONE sample per pixel (development tool for ncu launch lists):
wave_once.py [mega|wavefront] [spp]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch
import bench
import gpt_b200 as gpt

class A: width, height, depth = 1920, 1080, 8

mode = sys.argv[1] if len(sys.argv) > 1 else "wavefront"
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 1
sc = bench.load_scene_arrays()
ctx = gpt.Context(0, 0 if mode == "wavefront" else gpt.FLAG_MEGAKERNEL)
ctx.upload_scene(sc["nodes"], sc["tris"], sc["indices"], bench.bench_spheres(gpt))
info = bench.bench_info(gpt, A)
acc = torch.zeros(A.height, A.width, 4, device="cuda")
for i in range(2):
    ctx.render(info, acc, 0, i * spp, spp)
torch.cuda.synchronize()
print("ok", ctx.stats().rays)
