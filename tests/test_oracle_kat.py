"""Known-answer tests that pin the oracle (oracle/oracle.cpp).  The reference ships no tests or golden
vectors for this path, so the oracle is pinned by published vectors (Philox),
closed-form geometry, an independent Python evaluation of the reference formulas, and an O(N)
brute-force sweep that cross-checks the BVH traversal on the reference-built trees."""
import ctypes as C

import numpy as np
import pytest

from support import reflib, scenes

L = None


def setup_module(_m):
    global L
    L = reflib.oracle()


def f3(*v):
    return (C.c_float * 3)(*v)


# ---- RNG -------------------------------------------------------------------------------------
def test_philox_known_answers():
    # Random123 kat_vectors: philox4x32-10
    kats = [((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
            ((0xffffffff,) * 4, (0xffffffff,) * 2, (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
            ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0), (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1))]
    for ctr, key, want in kats:
        out = (C.c_uint32 * 4)()
        L.orc_philox4x32_10((C.c_uint32 * 4)(*ctr), (C.c_uint32 * 2)(*key), out)
        assert tuple(out) == want


def test_uniform_range_is_half_open_like_curand():
    # curand_uniform returns (0,1]
    assert L.orc_u01(0) == pytest.approx(2.0 ** -24) and L.orc_u01(0) > 0.0
    assert L.orc_u01(0xFFFFFFFF) == 1.0
    assert L.orc_u01(0x80000000) == pytest.approx(0.5, abs=1e-7)


def test_hash64_matches_python_bigint():
    # utilfun.cpp:380-389 evaluated with Python integers
    def h(key):
        m = (1 << 64) - 1
        key = ((~key) + (key << 21)) & m
        key ^= key >> 24
        key = ((key + (key << 3)) + (key << 8)) & m
        key ^= key >> 14
        key = ((key + (key << 2)) + (key << 4)) & m
        key ^= key >> 28
        key = (key + (key << 31)) & m
        return key
    for k in [0, 1, 2, 12345, 2 ** 40 + 7, 2 ** 64 - 1]:
        assert L.orc_hash64(C.c_uint64(k)) == h(k)


# ---- deterministic transcendental functions ----------------------------------------------------
def test_sincos2pi_accuracy():
    us = np.concatenate([np.linspace(0, 1, 4097), np.random.default_rng(0).random(4096)]).astype(np.float32)
    s, c = C.c_float(), C.c_float()
    err = 0.0
    for u in us:
        L.orc_sincos2pi(C.c_float(u), C.byref(s), C.byref(c))
        err = max(err, abs(s.value - np.sin(2 * np.pi * float(u))), abs(c.value - np.cos(2 * np.pi * float(u))))
    assert err < 4e-7


def test_powf_accuracy_and_edges():
    rng = np.random.default_rng(1)
    for x, y in zip(rng.random(2000).astype(np.float32), (1.0 / (1.0 + rng.random(2000) * 200)).astype(np.float32)):
        got = L.orc_powf(C.c_float(x), C.c_float(y))
        assert got == pytest.approx(float(x) ** float(y), rel=3e-6)
    assert L.orc_powf(C.c_float(0.0), C.c_float(1 / 31)) == 0.0   # r2 == 1 -> cosTheta == 0
    assert L.orc_powf(C.c_float(1.0), C.c_float(1 / 31)) == 1.0


# ---- triangle: cudaUtils.h:135-172 ----------------------------------------------------------------
def tri(v0, v1, v2, o, d, cull=1):
    out = (C.c_float * 3)()
    hit = L.orc_intersect_triangle(f3(*v0), f3(*v1), f3(*v2), f3(*o), f3(*d), cull, out)
    return hit, tuple(out)


def test_triangle_known_answers():
    v0, v1, v2 = (0, 0, -5), (1, 0, -5), (0, 1, -5)          # CCW seen from +z
    hit, (u, v, t) = tri(v0, v1, v2, (0.25, 0.25, 0), (0, 0, -1))
    assert hit and (u, v, t) == (0.25, 0.25, 5.0)
    assert tri(v0, v1, v2, (0.25, 0.25, -10), (0, 0, 1), cull=1)[0] == 0      # back face culled
    hit, (_, _, t) = tri(v0, v1, v2, (0.25, 0.25, -10), (0, 0, 1), cull=0)     # back face kept when cull off
    assert hit and t == 5.0
    assert tri(v0, v1, v2, (0.25, 0.25, -6), (0, 0, -1))[0] == 0              # behind the origin (t<0)
    assert tri(v0, v1, v2, (0.8, 0.8, 0), (0, 0, -1))[0] == 0                 # u+v > 1
    assert tri(v0, v1, v2, (-0.01, 0.2, 0), (0, 0, -1))[0] == 0               # u < 0
    hit, (u, v, t) = tri(v0, v1, v2, (0.0, 0.0, 0), (0, 0, -1))               # vertex: inclusive bounds (:160-161)
    assert hit and u == 0.0 and v == 0.0
    hit, (u, v, t) = tri(v0, v1, v2, (0.5, 0.5, 0), (0, 0, -1))               # hypotenuse u+v == 1 accepted
    assert hit
    assert tri(v0, v1, v2, (0.2, 0.2, -5), (1, 0, 0))[0] == 0                 # parallel: det == 0 inside the eps band


def test_triangle_absolute_det_epsilon():
    # det = area-scaled; |det| < 1e-5 is rejected regardless of geometry (SURVEY Q8 / hard part)
    s = 0.003                                                                # det = s*s = 9e-6 < 1e-5
    assert tri((0, 0, -1), (s, 0, -1), (0, s, -1), (s / 4, s / 4, 0), (0, 0, -1))[0] == 0
    s = 0.004                                                                # det = 1.6e-5 > 1e-5
    assert tri((0, 0, -1), (s, 0, -1), (0, s, -1), (s / 4, s / 4, 0), (0, 0, -1))[0] == 1


# ---- sphere: CommomStructs.hpp:23-31 ---------------------------------------------------------------
def sph(o, d, c=(0, 0, -10), r=2.0):
    s = reflib.OrcSphere()
    s.posRad[:] = [c[0], c[1], c[2], r]
    L.orc_sphere_intersect.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    return L.orc_sphere_intersect(C.byref(s), f3(*o), f3(*d))


def test_sphere_known_answers():
    assert sph((0, 0, 0), (0, 0, -1)) == 8.0                      # front hit
    assert sph((0, 0, -10), (0, 0, -1)) == 2.0                    # from inside: far root
    assert sph((0, 0, 0), (0, 0, 1)) == 0.0                       # behind
    assert sph((3, 0, 0), (0, 0, -1)) == 0.0                      # miss (disc < 0)
    assert sph((2, 0, 0), (0, 0, -1)) == 10.0                     # tangent: disc == 0
    assert sph((0, 0, -8.005), (0, 0, 1)) == 0.0                  # both roots <= eps 0.01 -> 0


# ---- box spec: cudaUtils.h:43-58 ---------------------------------------------------------------------
def test_box_known_answers():
    b = reflib.OrcBox()
    b.mn[:] = [-1, -1, -6]
    b.mx[:] = [1, 1, -4]
    L.orc_box_intersect.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    assert L.orc_box_intersect(C.byref(b), f3(0.1, 0.2, 0), f3(0, 0, -1)) == 4.0
    assert L.orc_box_intersect(C.byref(b), f3(3, 0.2, 0), f3(0, 0, -1)) == 0.0
    assert L.orc_box_intersect(C.byref(b), f3(0.1, 0.2, -5), f3(0, 0, -1)) == 0.0   # inside: entry t<0 -> spec returns 0


# ---- camera: cudaUtils.h:111-134 ---------------------------------------------------------------------
def test_camera_ray_formula(gpt):
    info = gpt.default_kernel_info(1280, 720)
    p = scenes.orc_params(info, 0)
    o, d = (C.c_float * 3)(), (C.c_float * 3)()
    L.orc_camera_ray(C.byref(p), 640, 360, C.c_float(0), C.c_float(0), o, d)
    # independent evaluation (float64) of the reference expression
    dist, asp, fov = 12.0, 1280 / 720, np.tan(np.radians(45.0))
    xs = (640 - 640 + 0.5) * dist * asp * fov / 1279
    ys = (360 - 360 + 0.5) * dist * fov / 719
    dirv = np.array([0, 0, -1.0]) * dist + np.array([1.0, 0, 0]) * xs + np.array([0, 1.0, 0]) * ys
    assert info.cam.dist == 12.0                                   # 720/60 integer division
    np.testing.assert_allclose(np.array(o[:]), dirv, rtol=2e-6, atol=2e-6)     # origin ON the image plane (pos = 0)
    np.testing.assert_allclose(np.array(d[:]), dirv / np.linalg.norm(dirv), rtol=2e-6, atol=2e-6)


# ---- BVH traversal vs. brute force on the reference-built trees ------------------------------------------
@pytest.mark.parametrize("name,n", [("cornell", 4000), ("sphere_mesh", 4000), ("bunny_cornell", 1500)])
@pytest.mark.parametrize("cull", [True, False])
def test_traversal_equals_bruteforce(name, n, cull):
    sc = scenes.load_scene(name)
    o, d = scenes.random_rays(sc, n, seed=7)
    tri_i, addr, t, cnt = scenes.oracle_trace_rays(sc, o, d, cull=cull)
    _, baddr, bt, _ = scenes.oracle_trace_rays(sc, o, d, cull=cull, brute=True)
    assert (t < 3e38).sum() > n // 10, "ray set should hit the mesh"
    np.testing.assert_array_equal(t, bt)                           # nearest distance: exact
    # same triangle unless two slots tie on t (SBVH duplicates references; shared edges)
    differ = addr != baddr
    for i in np.nonzero(differ)[0]:
        o1, d1 = o[i:i + 1], d[i:i + 1]
        # the BVH-visited slot must itself be a hit at exactly that t
        tv = sc["tris"][addr[i]:addr[i] + 3, :3]
        out = (C.c_float * 3)()
        assert L.orc_intersect_triangle(f3(*tv[0]), f3(*tv[1]), f3(*tv[2]), f3(*o1[0]), f3(*d1[0]), int(cull), out)
        assert out[2] == t[i]
    hit = addr >= 0
    np.testing.assert_array_equal(tri_i[hit], sc["indices"][addr[hit]])       # remap, cudaUtils.h:452-456
    assert (tri_i[~hit] == -1).all()
    assert cnt.max_stack < 64


def test_bvh_layout_invariants():
    """What CudaBVH::createCompact promises (CudaBVH.cpp:121-270) and the device re-pack relies on."""
    for name in ["cornell", "sphere_mesh", "bunny_cornell"]:
        sc = scenes.load_scene(name)
        nodes = sc["nodes"].view(np.int32).reshape(-1, 4, 4)
        tris_u = sc["tris"].view(np.uint32)
        n_leaf = 0
        for i, nd in enumerate(nodes):
            for c in nd[3, :2]:
                if c >= 0:
                    assert c % 64 == 0 and c // 64 > i and c // 64 < len(nodes)
                else:
                    n_leaf += 1
                    slot = ~c
                    k = 0
                    while tris_u[slot, 0] != 0x80000000:
                        slot += 3
                        k += 1
                    assert k >= 1
        assert n_leaf == len(nodes) + 1
        assert (tris_u[:, 0] == 0x80000000).sum() >= n_leaf
        assert len(sc["indices"]) == len(sc["tris"])
