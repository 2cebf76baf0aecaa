"""Scene fixtures + conversion between the product's kernelInfo and the oracle's parameter structs."""
import ctypes as C
import os

import numpy as np

from . import reflib

GOLDEN = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "golden")


def scene_path(name):
    p = os.path.join(GOLDEN, "scene_%s.npz" % name)
    return p if os.path.exists(p) else None


def load_scene(name):
    """dict(verts, faces, nodes(n,4), tris(n,4), indices[, woop(n,4)]) built by the REFERENCE's SplitBVHBuilder+CudaBVH."""
    p = scene_path(name)
    if p is None:
        raise FileNotFoundError("scene fixture %s missing (run tests/golden/make_golden.py where /root/reference is mounted)" % name)
    z = np.load(p)
    return {k: np.ascontiguousarray(z[k]) for k in z.files}


def orc_spheres(gpt_spheres):
    """product Sphere ctypes array -> oracle sphere array (same 44-byte layout)."""
    n = len(gpt_spheres) if gpt_spheres is not None else 0
    arr = (reflib.OrcSphere * max(n, 1))()
    if n:
        C.memmove(arr, gpt_spheres, n * 44)
    return arr, n


def orc_scene(scene, spheres, n_spheres, boxes=None, n_boxes=0, tri_materials=None, face_material=None):
    """tri_materials: product TriMaterial ctypes array (28 B each = the oracle's 7-float rows); face_material: int32 array."""
    sc = reflib.OrcScene()
    if face_material is not None and len(face_material):
        sc.tri_materials = C.addressof(tri_materials)
        sc.n_tri_materials = len(tri_materials)
        sc.face_material = face_material.ctypes.data
        sc.n_faces = len(face_material)
    if scene is not None:
        sc.nodes = scene["nodes"].ctypes.data
        sc.n_node_vec4 = scene["nodes"].shape[0]
        sc.tris = scene["tris"].ctypes.data
        sc.n_tri_vec4 = scene["tris"].shape[0]
        sc.indices = scene["indices"].ctypes.data
        sc.n_indices = scene["indices"].shape[0]
    sc.spheres = C.addressof(spheres) if n_spheres else None
    sc.n_spheres = n_spheres
    sc.boxes = C.addressof(boxes) if n_boxes else None
    sc.n_boxes = n_boxes
    return sc


def orc_params(info, seed, flags=0):
    """kernelInfo (product ctypes struct) -> OrcParams."""
    p = reflib.OrcParams()
    p.width, p.height = info.width, info.height
    for k in ("front", "right", "up", "pos"):
        getattr(p.cam, k)[:] = list(getattr(info.cam, k))
    p.cam.dist, p.cam.aspect, p.cam.fov = info.cam.dist, info.cam.aspect, info.cam.fov
    p.cull_back_faces = int(info.cullBackFaces)
    p.depth = info.depth
    p.air_ref_index, p.glass_ref_index = info.air_ref_index, info.glass_ref_index
    p.tri_mat = info.triCurrentMat
    p.bk_color[:] = list(info.bkColor)
    p.phong_expo = info.phongExpo
    p.emi[:] = list(info.emi)
    p.col[:] = list(info.col)
    p.seed = seed
    p.flags = flags
    return p


def oracle_render(scene, spheres, info, seed, sample_begin, sample_count, flags=0, threads=8, counters=False, boxes=None,
                  tri_materials=None, face_material=None):
    """float32 (H,W,4) sums + optional counters, via orc_render."""
    L = reflib.oracle()
    sph, ns = orc_spheres(spheres)
    nb = len(boxes) if boxes is not None else 0
    barr = None
    if nb:
        barr = (reflib.OrcBox * nb)()
        C.memmove(barr, boxes, nb * 52)
    if face_material is not None:
        face_material = np.ascontiguousarray(face_material, dtype=np.int32)
    sc = orc_scene(scene, sph, ns, barr, nb, tri_materials, face_material)
    p = orc_params(info, seed, flags)
    acc = np.zeros((info.height, info.width, 4), dtype=np.float32)
    cnt = reflib.OrcCounters()
    rc = L.orc_render(C.byref(sc), C.byref(p), C.c_uint64(sample_begin), C.c_uint32(sample_count),
                      acc.ctypes.data_as(C.c_void_p), C.byref(cnt) if counters else None, C.c_int(threads))
    assert rc == 0
    return (acc, cnt) if counters else acc


def oracle_primary(scene, spheres, info, seed, sample, flags=0, boxes=None):
    L = reflib.oracle()
    sph, ns = orc_spheres(spheres)
    nb = len(boxes) if boxes is not None else 0
    barr = None
    if nb:
        barr = (reflib.OrcBox * nb)()
        C.memmove(barr, boxes, nb * 52)
    sc = orc_scene(scene, sph, ns, barr, nb)
    p = orc_params(info, seed, flags)
    n = info.width * info.height
    geom = np.zeros(n, np.int32); tri = np.zeros(n, np.int32); prim = np.zeros(n, np.int32); t = np.zeros(n, np.float32)
    cnt = reflib.OrcCounters()
    rc = L.orc_primary_hits(C.byref(sc), C.byref(p), C.c_uint64(sample), geom.ctypes.data_as(C.c_void_p), tri.ctypes.data_as(C.c_void_p),
                            prim.ctypes.data_as(C.c_void_p), t.ctypes.data_as(C.c_void_p), C.byref(cnt))
    assert rc == 0
    return geom, tri, prim, t, cnt


def oracle_trace_rays(scene, o, d, cull=True, brute=False):
    L = reflib.oracle()
    sph, ns = orc_spheres(None)
    sc = orc_scene(scene, sph, 0)
    o = np.ascontiguousarray(o, np.float32); d = np.ascontiguousarray(d, np.float32)
    n = o.shape[0]
    tri = np.zeros(n, np.int32); addr = np.zeros(n, np.int32); t = np.zeros(n, np.float32)
    if brute:
        rc = L.orc_trace_rays_bruteforce(C.byref(sc), o.ctypes.data_as(C.c_void_p), d.ctypes.data_as(C.c_void_p), C.c_size_t(n), C.c_int(int(cull)),
                                         addr.ctypes.data_as(C.c_void_p), t.ctypes.data_as(C.c_void_p))
        assert rc == 0
        return None, addr, t, None
    cnt = reflib.OrcCounters()
    rc = L.orc_trace_rays(C.byref(sc), o.ctypes.data_as(C.c_void_p), d.ctypes.data_as(C.c_void_p), C.c_size_t(n), C.c_int(int(cull)),
                          tri.ctypes.data_as(C.c_void_p), addr.ctypes.data_as(C.c_void_p), t.ctypes.data_as(C.c_void_p), C.byref(cnt))
    assert rc == 0
    return tri, addr, t, cnt


def random_rays(scene, n, seed=0):
    """Incoherent rays aimed through the mesh's bounding box (plus a few that miss)."""
    rng = np.random.default_rng(seed)
    v = scene["verts"]
    lo, hi = v.min(0), v.max(0)
    c, ext = (lo + hi) / 2, (hi - lo)
    o = (c + (rng.random((n, 3)) - 0.5) * ext * 3.0).astype(np.float32)
    tgt = (c + (rng.random((n, 3)) - 0.5) * ext * 1.2).astype(np.float32)
    d = tgt - o
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    return o, d.astype(np.float32)
