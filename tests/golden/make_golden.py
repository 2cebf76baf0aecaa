#!/usr/bin/env python
"""Regenerates the scene fixtures from the reference's own assets and builder.

Needs /root/reference (assets) and oracle/_ref/librefbvh.so (the reference's SplitBVHBuilder +
CudaBVH compiled in place by oracle/Makefile).  Small scenes are committed under tests/golden/;
the two larger ones (BASELINE configs 2 and 3) are committed too, without the unused Woop array, so that the
-m gpu parity tests and bench.py's per-config records find them on any box.

  cornell        Assets/cornell.obj                                   (BASELINE config 0)
  bunny_cornell  Assets/cornell.obj + Assets/bunny_low.obj translated (0,-15.2,-40)  (config 1)
  sphere_mesh    Assets/sphere.obj (64 faces, 48 quads fan-triangulated) tiny case
  dragon_cornell Assets/cornell.obj + Assets/dragon.obj (100 000 tris)               (config 2 stand-in; cornell_dragon.obj is a missing blob)
  gto_sixteen    Assets/gto_sixteen.obj                                              (config 3 base mesh, instanced 4x4 by the tests)
"""
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from support import objio, reflib  # noqa: E402

ASSETS = "/root/reference/Assets"


def build(name, meshes, big=False):
    v, f = objio.merge(meshes)
    t0 = time.time()
    b = reflib.ref_build_bvh(v, f)
    path = os.path.join(HERE, "scene_%s.npz" % name)
    arrays = dict(verts=v, faces=f, nodes=b["nodes"], tris=b["tris"], indices=b["indices"])
    if not big:
        arrays["woop"] = b["woop"]          # CudaBVH.cpp:244-262; nothing reads it, kept for the small fixtures only
    np.savez_compressed(path, **arrays)
    print("%-16s %7d tris  %6d inner nodes  %7d tri slots  build %.2fs  -> %s (%.0f KB)" % (
        name, len(f), b["nodes"].shape[0] // 4, b["tris"].shape[0], time.time() - t0, os.path.relpath(path), os.path.getsize(path) / 1024))


def main():
    if not os.path.isdir(ASSETS):
        print("no /root/reference: keeping committed fixtures")
        return 0
    if not reflib.have_refbvh():
        print("oracle/_ref/librefbvh.so missing: run `make -C oracle` first")
        return 1
    cornell = objio.load_obj(os.path.join(ASSETS, "cornell.obj"))
    bunny_v, bunny_f = objio.load_obj(os.path.join(ASSETS, "bunny_low.obj"))
    bunny = (bunny_v + np.array([0.0, -15.2, -40.0], np.float32), bunny_f)
    build("cornell", [cornell])
    build("bunny_cornell", [cornell, bunny])
    # cube.obj (12 tris) and plane.obj (1 tri) make the reference's createCompact assert: the SAH keeps the
    # root a leaf (CudaBVH.cpp:141) -- the reference cannot render them either, so they are not fixtures.
    build("sphere_mesh", [objio.load_obj(os.path.join(ASSETS, "sphere.obj"))])
    if "--small" not in sys.argv:
        build("gto_sixteen", [objio.load_obj(os.path.join(ASSETS, "gto_sixteen.obj"))], big=True)
        build("dragon_cornell", [cornell, objio.load_obj(os.path.join(ASSETS, "dragon.obj"))], big=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
