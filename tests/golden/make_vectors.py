#!/usr/bin/env python
"""Mints golden output vectors from the CPU oracle (oracle/liboracle.so) for the committed scenes:
primary-hit maps (jitter off) and a multi-spp radiance sum at thumbnail size.  Run after a deliberate
change to the numeric contract; tests/test_oracle_render.py::test_golden_vectors and the GPU parity
tests compare against these files."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import gpt_b200 as gpt  # noqa: E402  (host-only helpers: default spheres / camera; no GPU needed)
from support import reflib, scenes  # noqa: E402

W, H, DEPTH, SEED, SPP = 96, 54, 6, 2017, 4

for name in ["cornell", "bunny_cornell"]:
    sc = scenes.load_scene(name)
    sph = gpt.default_spheres()
    info = gpt.default_kernel_info(W, H)
    info.depth = DEPTH
    info.cam.dist = 12.0 * H / 720.0
    geom, tri, prim, t, _ = scenes.oracle_primary(sc, sph, info, SEED, 0, flags=reflib.ORC_FLAG_NO_JITTER)
    acc = scenes.oracle_render(sc, sph, info, SEED, 0, SPP)
    out = os.path.join(HERE, "golden_%s.npz" % name)
    np.savez_compressed(out, meta=np.array([W, H, DEPTH, SEED, SPP]), geom=geom.astype(np.int8), tri_idx=tri,
                        prim_idx=prim.astype(np.int8), t=t, sum_rgb=acc[..., :3])
    print(out, os.path.getsize(out) // 1024, "KB", "tri pixels", int((geom == 0).sum()))
