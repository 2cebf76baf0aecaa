"""CPU tests of the oracle's bounce loop / accumulation semantics (tracer.cu:27-400) and of the
committed golden vectors minted from it (tests/golden/golden_*.npz, made by tests/golden/make_vectors.py)."""
import ctypes as C
import os

import numpy as np
import pytest

from support import reflib, scenes

GOLDEN = scenes.GOLDEN


def small_info(gpt, w=64, h=36, depth=4):
    info = gpt.default_kernel_info(w, h)
    info.depth = depth
    # keep the 1280x720 framing at thumbnail size: dist = h/60 would be 0 (integer division)
    info.cam.dist = 12.0 * h / 720.0
    return info


def test_render_is_deterministic_and_counts_samples(gpt):
    sc = scenes.load_scene("cornell")
    info = small_info(gpt)
    a = scenes.oracle_render(sc, gpt.default_spheres(), info, 5, 0, 3, threads=4)
    b = scenes.oracle_render(sc, gpt.default_spheres(), info, 5, 0, 3, threads=1)
    np.testing.assert_array_equal(a, b)
    assert (a[..., 3] == 3.0).all() and np.isfinite(a).all() and (a[..., :3] >= 0).all()
    c = scenes.oracle_render(sc, gpt.default_spheres(), info, 6, 0, 3)
    assert not np.array_equal(a, c)                       # seed matters


def test_sample_slices_partition_the_sum(gpt):
    """RNG keyed by GLOBAL sample index: any split of [0,N) over ranks gives the same samples."""
    sc = scenes.load_scene("cornell")
    info = small_info(gpt)
    sph = gpt.default_spheres()
    whole = scenes.oracle_render(sc, sph, info, 1, 0, 4)
    parts = scenes.oracle_render(sc, sph, info, 1, 0, 1) + scenes.oracle_render(sc, sph, info, 1, 1, 2) + scenes.oracle_render(sc, sph, info, 1, 3, 1)
    np.testing.assert_allclose(parts, whole, rtol=3e-7, atol=1e-7)
    np.testing.assert_array_equal(parts[..., 3], whole[..., 3])


def test_miss_returns_background_unweighted(gpt):
    """Q2: a miss returns bkColor, dropping mask and accumulated emission (tracer.cu:140-142)."""
    sc = scenes.load_scene("cornell")
    info = small_info(gpt)
    info.bkColor[:] = [0.25, 0.5, 0.75]
    info.emi[:] = [0.3, 0.3, 0.3]                        # emissive triangles: would be visible if accu survived
    acc = scenes.oracle_render(sc, None, info, 0, 0, 1)
    rgb = acc[..., :3]
    is_bk = np.all(rgb == np.array([0.25, 0.5, 0.75], np.float32), axis=-1)
    # without spheres every path eventually either escapes (exact bk) or survives `depth` bounces (pure emission sum)
    other = rgb[~is_bk]
    assert is_bk.mean() > 0.5
    assert np.all(other >= 0.3 - 1e-6)


def test_depth_zero_and_depth_one(gpt):
    sc = scenes.load_scene("cornell")
    sph = gpt.default_spheres()
    info = small_info(gpt, depth=0)
    assert (scenes.oracle_render(sc, sph, info, 0, 0, 2)[..., :3] == 0).all()
    info.depth = 1                                       # only first-hit emission mask*emi (tracer.cu:120,133,305)
    acc = scenes.oracle_render(sc, sph, info, 0, 0, 1, flags=reflib.ORC_FLAG_NO_JITTER)
    geom, tri, prim, t, _ = scenes.oracle_primary(sc, sph, info, 0, 0, flags=reflib.ORC_FLAG_NO_JITTER)
    rgb = acc[..., :3].reshape(-1, 3)
    assert (rgb[geom == 0] == 0).all()                   # triangles: global emi = 0
    for sid in np.unique(prim[geom == 1]):
        np.testing.assert_array_equal(rgb[(geom == 1) & (prim == sid)], np.tile(np.array(list(sph[sid].emi), np.float32), (((geom == 1) & (prim == sid)).sum(), 1)))


def test_legacy_running_mean_clamps_every_frame(gpt):
    """Q3: acc = clamp((acc*(N-1)+s)/N, 0, 1) per frame (tracer.cu:386-391); pack truncates (Q11)."""
    L = reflib.oracle()
    sc = scenes.load_scene("cornell")
    sph, ns = scenes.orc_spheres(gpt.default_spheres())
    osc = scenes.orc_scene(sc, sph, ns)
    info = small_info(gpt)
    acc = np.zeros((info.height, info.width, 3), np.float32)
    draw = np.zeros((info.height, info.width), np.uint32)
    mean = np.zeros_like(acc)
    for frame in range(3):
        seed = gpt.hash64(frame)
        p = scenes.orc_params(info, seed)
        assert L.orc_launch_legacy(C.byref(osc), C.byref(p), C.c_uint64(frame + 1), acc.ctypes.data_as(C.c_void_p), draw.ctypes.data_as(C.c_void_p), 4) == 0
        s = scenes.oracle_render(sc, gpt.default_spheres(), info, seed, 0, 1)[..., :3]
        n = np.float32(frame + 1)
        mean = np.clip((mean * np.float32(frame) + s) * (np.float32(1.0) / n), 0, 1).astype(np.float32)
        np.testing.assert_array_equal(acc, mean)
    assert acc.max() <= 1.0
    r = (draw & 0xFF).astype(np.int32)
    np.testing.assert_array_equal(r, (np.float32(255.0) * acc[..., 0]).astype(np.int32))     # truncation, not rounding
    assert (draw >> 24 == 0).all()                                                            # 0x00BBGGRR


def test_resolve_matches_mean_and_pack(gpt):
    L = reflib.oracle()
    rng = np.random.default_rng(3)
    acc = (rng.random((8, 8, 4)) * 40).astype(np.float32)
    acc[..., 3] = 16.0
    acc[0, 0] = [100, -3, 8, 16]                        # clamped both ways at resolve only
    rgba = np.zeros((8, 8), np.uint32)
    rgb = np.zeros((8, 8, 3), np.float32)
    assert L.orc_resolve(acc.ctypes.data_as(C.c_void_p), 8, 8, rgba.ctypes.data_as(C.c_void_p), rgb.ctypes.data_as(C.c_void_p)) == 0
    np.testing.assert_array_equal(rgb, acc[..., :3] * (np.float32(1.0) / np.float32(16.0)))
    cl = np.clip(rgb, 0, 1)
    want = (np.float32(255) * cl).astype(np.uint32)
    np.testing.assert_array_equal(rgba, want[..., 0] | (want[..., 1] << 8) | (want[..., 2] << 16))


@pytest.mark.parametrize("name", ["cornell", "bunny_cornell"])
def test_golden_vectors(gpt, name):
    """Committed outputs of the oracle (tests/golden/make_vectors.py): guards the oracle against drift and gives
    the GPU tests a fixed target that does not depend on re-running the oracle."""
    path = os.path.join(GOLDEN, "golden_%s.npz" % name)
    z = np.load(path)
    sc = scenes.load_scene(name)
    sph = gpt.default_spheres()
    w, h, depth, seed, spp = [int(x) for x in z["meta"]]
    info = small_info(gpt, w, h, depth)
    geom, tri, prim, t, _ = scenes.oracle_primary(sc, sph, info, seed, 0, flags=reflib.ORC_FLAG_NO_JITTER)
    np.testing.assert_array_equal(geom.astype(np.int8), z["geom"])
    np.testing.assert_array_equal(tri, z["tri_idx"])
    np.testing.assert_array_equal(prim.astype(np.int8), z["prim_idx"])
    np.testing.assert_array_equal(t, z["t"])
    acc = scenes.oracle_render(sc, sph, info, seed, 0, spp)
    np.testing.assert_array_equal(acc[..., :3], z["sum_rgb"])
