"""Minimal OBJ reader used by the fixture generator (tests/golden/make_golden.py).

Mirrors what the reference's loader keeps (GpuPathTracer/utilfun.cpp:454-524
``loadIndexedTris``): every ``v`` line in file order, and per face the 0-based
``vertex_index`` of each corner; polygons are fan-triangulated like tinyobjloader
1.0.x does (v0, v[k], v[k+1]).  Materials / normals / uvs are dropped.
"""
import numpy as np


def load_obj(path):
    verts, faces = [], []
    with open(path, "r") as fh:
        for line in fh:
            if line.startswith("v "):
                p = line.split()
                verts.append((float(p[1]), float(p[2]), float(p[3])))
            elif line.startswith("f "):
                idx = []
                for tok in line.split()[1:]:
                    i = int(tok.split("/")[0])
                    idx.append(i - 1 if i > 0 else len(verts) + i)
                for k in range(1, len(idx) - 1):
                    faces.append((idx[0], idx[k], idx[k + 1]))
    return np.asarray(verts, dtype=np.float32), np.asarray(faces, dtype=np.int32)


def merge(meshes):
    """Concatenate (verts, tris) meshes into one shape (the reference loader asserts one shape)."""
    vs, ts, off = [], [], 0
    for v, t in meshes:
        vs.append(v)
        ts.append(t + off)
        off += len(v)
    return np.concatenate(vs).astype(np.float32), np.concatenate(ts).astype(np.int32)


def write_obj(path, verts, tris):
    with open(path, "w") as fh:
        fh.write("o mesh\n")
        for v in verts:
            fh.write("v %.9g %.9g %.9g\n" % (v[0], v[1], v[2]))
        for t in tris:
            fh.write("f %d %d %d\n" % (t[0] + 1, t[1] + 1, t[2] + 1))
