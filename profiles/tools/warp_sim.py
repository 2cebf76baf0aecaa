#!/usr/bin/env python
"""Driver of profiles/tools/warp_sim.cpp: dumps a scene (reference-built SBVH arrays + the default spheres) to a flat file.
   python profiles/tools/warp_sim.py bunny_cornell /tmp/sim/bunny.bin
   python profiles/tools/warp_sim.py synthetic_1000000 /tmp/sim/syn1m.bin      (builds the tree with the reference's SplitBVHBuilder)"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import gpt_b200 as gpt  # noqa: E402
from support import procedural, reflib, scenes  # noqa: E402

name, out = sys.argv[1], sys.argv[2]
if name.startswith("synthetic_"):
    v, f = procedural.synthetic_scene(int(name.split("_")[1]), seed=1234)
    sc = reflib.ref_build_bvh(v, f)
elif name == "dragon16":
    s0 = scenes.load_scene("dragon_cornell")
    v, f = procedural.subdivide(s0["verts"], s0["faces"], 2)
    sc = reflib.ref_build_bvh(v, f)
else:
    sc = scenes.load_scene(name)
sph = gpt.default_spheres()
sp = np.frombuffer(bytes(sph), dtype=np.float32).copy()
with open(out, "wb") as fh:
    np.array([sc["nodes"].shape[0], sc["tris"].shape[0], len(sph)], np.int32).tofile(fh)
    sc["nodes"].astype(np.float32).tofile(fh)
    sc["tris"].astype(np.float32).tofile(fh)
    sp.tofile(fh)
print(name, sc["nodes"].shape, sc["tris"].shape, len(sph))
