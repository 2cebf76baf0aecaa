#!/usr/bin/env python
"""Counts the ALGORITHMIC bytes per ray of the bench workload with the CPU oracle (the reference algorithm on
the reference BVH2 arrays; SURVEY 8-d formula) and writes profiles/r01_algorithmic_bytes.json, the per-unit
figure bench.py's roofline uses.  Full 1080p frame, 1 spp (16.6 M rays), seed 0."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402
import gpt_b200 as gpt  # noqa: E402
from support import scenes  # noqa: E402


class A:
    width, height, depth = 1920, 1080, 8


sc = bench.load_scene_arrays()
info = bench.bench_info(gpt, A)
acc, cnt = scenes.oracle_render(sc, bench.bench_spheres(gpt), info, 0, 0, 1, counters=True, threads=os.cpu_count())
out = {bench.WORKLOAD: {
    "rays": cnt.rays, "rays_by_bounce": list(cnt.rays_by_bounce)[:8],
    "inner_nodes_per_ray": cnt.inner_nodes / cnt.rays, "tri_tests_per_ray": cnt.tri_tests / cnt.rays,
    "leaf_terminators_per_ray": cnt.leaf_terms / cnt.rays, "tri_hits_per_ray": cnt.tri_hits / cnt.rays,
    "sphere_tests_per_ray": cnt.sphere_tests / cnt.rays, "max_stack": cnt.max_stack,
    "bytes_per_ray": cnt.algorithmic_bytes() / cnt.rays,
    "formula": "64*N_inner + 48*N_tri + 16*N_leaf + 4*[tri hit] + 44*N_sphere (SURVEY 8-d); +32 B/pixel/launch accum",
    "how": "oracle/liboracle.so orc_render, 1920x1080, 1 spp, seed 0, depth 8, 9 spheres"}}
with open(os.path.join(ROOT, "profiles", "r01_algorithmic_bytes.json"), "w") as fh:
    json.dump(out, fh, indent=1)
print(json.dumps(out, indent=1))
