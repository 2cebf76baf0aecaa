"""Seeded procedural meshes for BASELINE config 5: tessellated icospheres + random axis-aligned
boxes as 12-triangle meshes, uniformly placed in x[-18,18] y[-14,14] z[-300,-25] (in front of the default camera),
radii U[0.5,3].  Triangle areas stay far above the reference's ABSOLUTE determinant epsilon 1e-5
(cudaUtils.h:139): the smallest icosphere (r=0.5) at the deepest subdivision used has triangles of area
> 1e-3.  Pure numpy; returns an indexed mesh (verts float32 (n,3), faces int32 (m,3)) like uf::loadIndexedTris
(utilfun.cpp:454-524)."""
import numpy as np

_T = (1.0 + 5.0 ** 0.5) / 2.0
_ICO_V = np.array([[-1, _T, 0], [1, _T, 0], [-1, -_T, 0], [1, -_T, 0], [0, -1, _T], [0, 1, _T], [0, -1, -_T], [0, 1, -_T],
                   [_T, 0, -1], [_T, 0, 1], [-_T, 0, -1], [-_T, 0, 1]], dtype=np.float64)
_ICO_F = np.array([[0, 11, 5], [0, 5, 1], [0, 1, 7], [0, 7, 10], [0, 10, 11], [1, 5, 9], [5, 11, 4], [11, 10, 2], [10, 7, 6],
                   [7, 1, 8], [3, 9, 4], [3, 4, 2], [3, 2, 6], [3, 6, 8], [3, 8, 9], [4, 9, 5], [2, 4, 11], [6, 2, 10],
                   [8, 6, 7], [9, 8, 1]], dtype=np.int64)
_CACHE = {}


def icosphere(level):
    """unit icosphere subdivided `level` times: 20 * 4**level outward-facing triangles"""
    if level in _CACHE:
        return _CACHE[level]
    v = _ICO_V / np.linalg.norm(_ICO_V, axis=1, keepdims=True)
    f = _ICO_F
    for _ in range(level):
        e = np.concatenate([f[:, [0, 1]], f[:, [1, 2]], f[:, [2, 0]]])
        es = np.sort(e, axis=1)
        uniq, inv = np.unique(es, axis=0, return_inverse=True)
        mid = v[uniq[:, 0]] + v[uniq[:, 1]]
        mid /= np.linalg.norm(mid, axis=1, keepdims=True)
        m = len(v) + inv.reshape(3, -1).T                         # midpoint vertex ids per face: (ab, bc, ca)
        v = np.concatenate([v, mid])
        a, b, c = f[:, 0], f[:, 1], f[:, 2]
        ab, bc, ca = m[:, 0], m[:, 1], m[:, 2]
        f = np.concatenate([np.stack([a, ab, ca], 1), np.stack([b, bc, ab], 1), np.stack([c, ca, bc], 1), np.stack([ab, bc, ca], 1)])
    _CACHE[level] = (v.astype(np.float32), f.astype(np.int32))
    return _CACHE[level]


_BOX_F = np.array([[0, 2, 1], [0, 3, 2], [4, 5, 6], [4, 6, 7], [0, 1, 5], [0, 5, 4], [1, 2, 6], [1, 6, 5], [2, 3, 7], [2, 7, 6],
                   [3, 0, 4], [3, 4, 7]], dtype=np.int32)


def synthetic_scene(n_tris_target, seed=1234, level=4, box_fraction=0.02):
    """~n_tris_target triangles: icospheres of `level` subdivisions + box_fraction of the triangles as boxes."""
    rng = np.random.default_rng(seed)
    sv, sf = icosphere(level)
    n_box = int(n_tris_target * box_fraction) // 12
    n_sph = max(1, (n_tris_target - 12 * n_box) // len(sf))
    lo, hi = np.array([-18.0, -14.0, -300.0]), np.array([18.0, 14.0, -25.0])
    c = (lo + rng.random((n_sph, 3)) * (hi - lo)).astype(np.float32)
    r = (0.5 + 2.5 * rng.random(n_sph)).astype(np.float32)
    verts = (sv[None, :, :] * r[:, None, None] + c[:, None, :]).reshape(-1, 3)
    faces = (sf[None, :, :] + (np.arange(n_sph, dtype=np.int64) * len(sv))[:, None, None]).reshape(-1, 3)
    if n_box:
        bc = lo + rng.random((n_box, 3)) * (hi - lo)
        bh = 0.25 + 1.25 * rng.random((n_box, 3))
        corners = np.array([[-1, -1, -1], [1, -1, -1], [1, 1, -1], [-1, 1, -1], [-1, -1, 1], [1, -1, 1], [1, 1, 1], [-1, 1, 1]], dtype=np.float64)
        bv = (bc[:, None, :] + corners[None, :, :] * bh[:, None, :]).reshape(-1, 3).astype(np.float32)
        bf = (_BOX_F[None, :, :] + (len(verts) + np.arange(n_box, dtype=np.int64) * 8)[:, None, None]).reshape(-1, 3)
        verts = np.concatenate([verts, bv])
        faces = np.concatenate([faces, bf])
    return np.ascontiguousarray(verts, dtype=np.float32), np.ascontiguousarray(faces, dtype=np.int32)


def subdivide(verts, faces, times=1):
    """1 -> 4 midpoint subdivision (shared midpoints welded); keeps the surface, multiplies the triangle count by 4."""
    v, f = verts.astype(np.float32), faces.astype(np.int64)
    for _ in range(times):
        e = np.concatenate([f[:, [0, 1]], f[:, [1, 2]], f[:, [2, 0]]])
        uniq, inv = np.unique(np.sort(e, axis=1), axis=0, return_inverse=True)
        mid = 0.5 * (v[uniq[:, 0]] + v[uniq[:, 1]])
        m = len(v) + inv.reshape(3, -1).T
        v = np.concatenate([v, mid.astype(np.float32)])
        a, b, c = f[:, 0], f[:, 1], f[:, 2]
        ab, bc, ca = m[:, 0], m[:, 1], m[:, 2]
        f = np.concatenate([np.stack([a, ab, ca], 1), np.stack([b, bc, ab], 1), np.stack([c, ca, bc], 1), np.stack([ab, bc, ca], 1)])
    return np.ascontiguousarray(v, np.float32), np.ascontiguousarray(f, np.int32)


def instance_grid(verts, faces, nx, nz, scale, x_step, z_first, z_step, floor_y=-15.0):
    """nx x nz copies of a mesh (vertex duplication before the BVH build, SURVEY cfg4 "multi-instance"): each copy is
    scaled about its bounding-box centre, put on the room's floor (y = floor_y; the default room is x[-20,20]
    y[-15,15], BasicScene.cpp:180-202) and placed on a grid in x and depth."""
    lo, hi = verts.min(0), verts.max(0)
    c = 0.5 * (lo + hi)
    base = (verts - c) * scale
    half_h = 0.5 * (hi[1] - lo[1]) * scale
    vs, fs = [], []
    k = 0
    for j in range(nz):
        for i in range(nx):
            off = np.array([(i - (nx - 1) / 2.0) * x_step, floor_y + half_h + 0.01, z_first - j * z_step], np.float32)
            vs.append((base + off).astype(np.float32))
            fs.append(faces + k * len(verts))
            k += 1
    return np.ascontiguousarray(np.concatenate(vs), np.float32), np.ascontiguousarray(np.concatenate(fs), np.int32)
