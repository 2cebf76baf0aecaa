"""GPU parity tests (run with -m gpu on the B200 box): the CUDA path, called through the C-ABI, against the CPU
oracle on the same seeded inputs, against the committed golden vectors, and -- at BASELINE's full size --
through size-independent properties.  Integer/index outputs and hit distances must match BIT FOR BIT; radiance
is compared bit-exact per pixel with a stated allowance for the (rare) pixels where it is not."""
import ctypes as C
import os

import numpy as np
import pytest

from support import reflib, scenes

pytestmark = pytest.mark.gpu

# Radiance tolerance (north_star: RMSE <= 1e-3 linear radiance on converged images).  The CUDA kernels and the
# oracle evaluate the same fp32 formulas with the same explicit FMAs (host libm sqrt/div and the device's are both
# IEEE), so per-sample sums are bit-identical: ZERO differing pixels is both expected and what every run has shown,
# and a one-pixel regression must fail.
MAX_DIFF_PIXEL_FRACTION = 0.0
MAX_RMSE = 0.0


def small_info(gpt, w=96, h=54, depth=6):
    info = gpt.default_kernel_info(w, h)
    info.depth = depth
    info.cam.dist = 12.0 * h / 720.0
    return info


@pytest.fixture(scope="module")
def ctxs(gpt, torch_cuda):
    made = {}

    def get(name, spheres="default", boxes=None, flags=0):
        key = (name, spheres if isinstance(spheres, str) else id(spheres), id(boxes), flags)
        if key not in made:
            ctx = gpt.Context(0, flags)
            sc = scenes.load_scene(name) if name else None
            sph = gpt.default_spheres() if isinstance(spheres, str) else spheres
            if sc is not None:
                ctx.upload_scene(sc["nodes"], sc["tris"], sc["indices"], sph, boxes)
            else:
                ctx.upload_scene(None, None, None, sph, boxes)
            made[key] = (ctx, sc, sph)
        return made[key]
    yield get
    for ctx, _, _ in made.values():
        ctx.close()


def dev(torch, arr):
    return torch.from_numpy(np.ascontiguousarray(arr)).cuda()


def gpu_render(torch, ctx, info, seed, begin, count):
    acc = torch.zeros(info.height, info.width, 4, dtype=torch.float32, device="cuda")
    ctx.render(info, acc, seed, begin, count)
    torch.cuda.synchronize()
    return acc.cpu().numpy()


def assert_radiance_close(got, want):
    assert got.shape == want.shape
    diff = np.any(got != want, axis=-1)
    frac = diff.mean()
    rmse = float(np.sqrt(np.mean((got.astype(np.float64) - want.astype(np.float64)) ** 2)))
    assert frac <= MAX_DIFF_PIXEL_FRACTION, "pixels differing from the oracle: %.4f%%" % (100 * frac)
    assert rmse <= MAX_RMSE, rmse


# ------------------------------------------------------------------ traversal: arbitrary incoherent rays
@pytest.mark.parametrize("name,n", [("cornell", 20000), ("sphere_mesh", 20000), ("bunny_cornell", 40000)])
@pytest.mark.parametrize("cull", [True, False])
def test_trace_rays_bit_exact(gpt, torch_cuda, ctxs, name, n, cull):
    torch = torch_cuda
    ctx, sc, _ = ctxs(name)
    o, d = scenes.random_rays(sc, n, seed=11)
    # edge cases: axis-parallel directions (the 2^-80 guard, cudaUtils.h:283-286) and rays starting inside the box
    d[:64] = 0.0
    d[:64, 2] = -1.0
    d[64:128] = 0.0
    d[64:128, 0] = 1.0
    tri_o, addr_o, t_o, _ = scenes.oracle_trace_rays(sc, o, d, cull=cull)
    tri = torch.empty(n, dtype=torch.int32, device="cuda")
    addr = torch.empty(n, dtype=torch.int32, device="cuda")
    t = torch.empty(n, dtype=torch.float32, device="cuda")
    ctx.trace_rays(dev(torch, o), dev(torch, d), n, cull, tri, addr, t)
    torch.cuda.synchronize()
    assert (t_o < 3e38).sum() > n // 20
    np.testing.assert_array_equal(t.cpu().numpy(), t_o)            # hit distance: bit exact
    np.testing.assert_array_equal(addr.cpu().numpy(), addr_o)      # same triangle slot: no t-tie exceptions needed (same visit order)
    np.testing.assert_array_equal(tri.cpu().numpy(), tri_o)        # remapped OBJ face index


def test_trace_rays_empty_and_ragged(gpt, torch_cuda, ctxs):
    torch = torch_cuda
    ctx, sc, _ = ctxs("cornell")
    ctx.trace_rays(None, None, 0, True)                            # n == 0 is a no-op
    for n in [1, 31, 33, 129]:                                     # ragged warps
        o, d = scenes.random_rays(sc, n, seed=n)
        _, addr_o, t_o, _ = scenes.oracle_trace_rays(sc, o, d)
        addr = torch.empty(n, dtype=torch.int32, device="cuda")
        t = torch.empty(n, dtype=torch.float32, device="cuda")
        ctx.trace_rays(dev(torch, o), dev(torch, d), n, True, None, addr, t)
        torch.cuda.synchronize()
        np.testing.assert_array_equal(t.cpu().numpy(), t_o)
        np.testing.assert_array_equal(addr.cpu().numpy(), addr_o)


# ------------------------------------------------------------------ primary hits
@pytest.mark.parametrize("name,w,h", [("cornell", 512, 512), ("bunny_cornell", 480, 270), ("sphere_mesh", 200, 120)])
@pytest.mark.parametrize("jitter", [False, True])
def test_primary_hits_exact(gpt, torch_cuda, ctxs, name, w, h, jitter):
    torch = torch_cuda
    flags = 0 if jitter else gpt.FLAG_NO_JITTER
    ctx, sc, sph = ctxs(name, flags=flags)
    info = gpt.default_kernel_info(w, h)
    if name == "sphere_mesh":
        gpt.set_camera(info, (0.0, 0.0, 30.0), 0.0, 0.0)           # sphere.obj: radius 9.35 at the origin
    geom_o, tri_o, prim_o, t_o, _ = scenes.oracle_primary(sc, sph, info, 99, 3, flags=0 if jitter else reflib.ORC_FLAG_NO_JITTER)
    n = w * h
    geom = torch.empty(n, dtype=torch.int32, device="cuda"); tri = torch.empty_like(geom); prim = torch.empty_like(geom)
    t = torch.empty(n, dtype=torch.float32, device="cuda")
    ctx.trace_primary(info, 99, 3, geom, tri, prim, t)
    torch.cuda.synchronize()
    assert (geom_o == 0).sum() > 100
    np.testing.assert_array_equal(geom.cpu().numpy(), geom_o)
    np.testing.assert_array_equal(tri.cpu().numpy(), tri_o)        # nearest-hit triangle IDs agree exactly
    np.testing.assert_array_equal(prim.cpu().numpy(), prim_o)
    np.testing.assert_array_equal(t.cpu().numpy(), t_o)


@pytest.mark.parametrize("name", ["cornell", "bunny_cornell"])
def test_against_golden_vectors(gpt, torch_cuda, ctxs, name):
    torch = torch_cuda
    z = np.load(os.path.join(scenes.GOLDEN, "golden_%s.npz" % name))
    w, h, depth, seed, spp = [int(x) for x in z["meta"]]
    info = small_info(gpt, w, h, depth)
    ctx, sc, sph = ctxs(name, flags=gpt.FLAG_NO_JITTER)
    n = w * h
    geom = torch.empty(n, dtype=torch.int32, device="cuda"); tri = torch.empty_like(geom); prim = torch.empty_like(geom)
    t = torch.empty(n, dtype=torch.float32, device="cuda")
    ctx.trace_primary(info, seed, 0, geom, tri, prim, t)
    torch.cuda.synchronize()
    np.testing.assert_array_equal(geom.cpu().numpy().astype(np.int8), z["geom"])
    np.testing.assert_array_equal(tri.cpu().numpy(), z["tri_idx"])
    np.testing.assert_array_equal(t.cpu().numpy(), z["t"])
    ctx2, _, _ = ctxs(name)
    acc = gpu_render(torch, ctx2, info, seed, 0, spp)
    assert (acc[..., 3] == spp).all()
    assert_radiance_close(acc[..., :3], z["sum_rgb"])


# ------------------------------------------------------------------ bounce loop: every material, every quirk
@pytest.mark.parametrize("tri_mat", [0, 1, 2, 3])                   # DIFF, METAL, SPEC, REFR (Mat enum order)
def test_render_matches_oracle_per_material(gpt, torch_cuda, ctxs, tri_mat):
    torch = torch_cuda
    ctx, sc, sph = ctxs("bunny_cornell")
    info = small_info(gpt, 160, 90, depth=8)
    info.triCurrentMat = tri_mat
    want = scenes.oracle_render(sc, sph, info, 7, 2, 3)
    got = gpu_render(torch, ctx, info, 7, 2, 3)
    np.testing.assert_array_equal(got[..., 3], want[..., 3])
    assert_radiance_close(got[..., :3], want[..., :3])


def test_render_mixed_sphere_materials_and_metal_fixed(gpt, torch_cuda, ctxs):
    torch = torch_cuda
    items = [(list(s.posRad), list(s.emi), list(s.col), s.mat) for s in gpt.default_spheres()]
    items.append(([-9.0, -9.0, -33.0, 5.0], [0, 0, 0], [0.9, 0.8, 0.7], gpt.METAL))
    items.append(([0.0, 4.0, -36.0, 4.0], [0, 0, 0], [1, 1, 1], gpt.REFR))
    sph = gpt.spheres_from_list(items)
    for flags, oflags in [(0, 0), (gpt.FLAG_METAL_FIXED, reflib.ORC_FLAG_METAL_FIXED)]:
        ctx, sc, _ = ctxs("bunny_cornell", spheres=sph, flags=flags)
        info = small_info(gpt, 160, 90, depth=8)
        info.phongExpo = 12.0
        want = scenes.oracle_render(sc, sph, info, 3, 0, 2, flags=oflags)
        got = gpu_render(torch, ctx, info, 3, 0, 2)
        assert_radiance_close(got[..., :3], want[..., :3])


def test_emitting_triangles_and_overflowing_mask(gpt, torch_cuda, ctxs):
    """The wavefront pipeline keeps the accumulated colour in memory and touches it only on a miss, on an emitting hit, or
    when the colour mask is no longer finite (then mask * 0 is NaN in tracer.cu:120/:133, and must be here).  Emitting
    triangles exercise the read-modify-write; a colour of 1e30 overflows the mask to inf within two bounces."""
    torch = torch_cuda
    for col, emi, depth in (((0.8, 0.7, 0.6), (0.25, 0.5, 0.125), 6), ((1e30, 1e30, 0.5), (0.0, 0.0, 0.0), 5), ((1e30, 0.9, -2.0), (0.0, 1.0, 0.0), 5)):
        for flags in (gpt.FLAG_WAVEFRONT, gpt.FLAG_MEGAKERNEL):
            ctx, sc, sph = ctxs("bunny_cornell", flags=flags)
            info = small_info(gpt, 128, 72, depth=depth)
            info.col[:] = col
            info.emi[:] = emi
            want = scenes.oracle_render(sc, sph, info, 5, 0, 2)
            got = gpu_render(torch, ctx, info, 5, 0, 2)
            g, w = got[..., :3], want[..., :3]
            # NaN counts as equal to NaN (the payload differs between x86 and the GPU); everything else bit for bit
            diff = np.any((g != w) & ~(np.isnan(g) & np.isnan(w)), axis=-1)
            assert np.isnan(w).any() == (col[0] > 1e20)           # the overflow cases do reach NaN in the reference arithmetic
            assert diff.mean() <= MAX_DIFF_PIXEL_FRACTION, (col, emi, flags, diff.mean())


def test_boxes_and_sphere_only_scene(gpt, torch_cuda, ctxs):
    """N2: Box primitive from the reference's commented spec; also a scene without any mesh."""
    torch = torch_cuda
    boxes = (gpt.Box * 2)()
    boxes[0].mn[:] = [-8, -15, -40]; boxes[0].mx[:] = [-2, -6, -34]; boxes[0].col[:] = [0.8, 0.3, 0.3]; boxes[0].mat = gpt.DIFF
    boxes[1].mn[:] = [3, -15, -45]; boxes[1].mx[:] = [9, -3, -39]; boxes[1].col[:] = [1, 1, 1]; boxes[1].mat = gpt.SPEC
    sph = gpt.default_spheres()
    ctx, _, _ = ctxs(None, spheres=sph, boxes=boxes)
    info = small_info(gpt, 160, 90, depth=5)
    want = scenes.oracle_render(None, sph, info, 1, 0, 2, boxes=boxes)
    got = gpu_render(torch, ctx, info, 1, 0, 2)
    assert_radiance_close(got[..., :3], want[..., :3])
    geom_o, _, prim_o, t_o, _ = scenes.oracle_primary(None, sph, info, 1, 0, boxes=boxes)
    assert (geom_o == 2).sum() > 50
    n = info.width * info.height
    geom = torch.empty(n, dtype=torch.int32, device="cuda"); prim = torch.empty_like(geom)
    ctx.trace_primary(info, 1, 0, geom, None, prim, None)
    torch.cuda.synchronize()
    np.testing.assert_array_equal(geom.cpu().numpy(), geom_o)
    np.testing.assert_array_equal(prim.cpu().numpy(), prim_o)


def test_miss_background_and_cull_off(gpt, torch_cuda, ctxs):
    torch = torch_cuda
    empty = (gpt.Sphere * 0)()
    ctx, sc, _ = ctxs("cornell", spheres=empty)
    info = small_info(gpt, 128, 72, depth=4)
    info.bkColor[:] = [0.25, 0.5, 0.75]
    info.emi[:] = [0.3, 0.2, 0.1]
    for cull in (1, 0):
        info.cullBackFaces = cull
        want = scenes.oracle_render(sc, None, info, 0, 0, 2)
        got = gpu_render(torch, ctx, info, 0, 0, 2)
        assert_radiance_close(got[..., :3], want[..., :3])


def test_depth_edge_cases(gpt, torch_cuda, ctxs):
    torch = torch_cuda
    ctx, sc, sph = ctxs("cornell")
    for depth in (0, 1, 16):
        info = small_info(gpt, 64, 36, depth=depth)
        want = scenes.oracle_render(sc, sph, info, 0, 0, 2)
        got = gpu_render(torch, ctx, info, 0, 0, 2)
        assert_radiance_close(got[..., :3], want[..., :3])
        if depth == 0:
            assert (got[..., :3] == 0).all()


def test_ragged_image_sizes(gpt, torch_cuda, ctxs):
    """Widths/heights that are not multiples of the 8x4 job tile, incl. the minimum 2x2."""
    torch = torch_cuda
    ctx, sc, sph = ctxs("cornell")
    for (w, h) in [(2, 2), (7, 5), (33, 17), (130, 67)]:
        info = small_info(gpt, w, h, depth=3)
        want = scenes.oracle_render(sc, sph, info, 4, 0, 2)
        got = gpu_render(torch, ctx, info, 4, 0, 2)
        assert (got[..., 3] == 2).all()
        assert_radiance_close(got[..., :3], want[..., :3])


# ------------------------------------------------------------------ accumulation contracts
def test_sample_slices_any_partition(gpt, torch_cuda, ctxs):
    """Multi-GPU contract: slices of the global sample range add up to the whole range."""
    torch = torch_cuda
    ctx, sc, sph = ctxs("bunny_cornell")
    info = small_info(gpt, 160, 90, depth=6)
    whole = gpu_render(torch, ctx, info, 5, 0, 8)
    acc = torch.zeros(info.height, info.width, 4, dtype=torch.float32, device="cuda")
    for (b, c) in [(0, 3), (3, 1), (4, 4)]:
        ctx.render(info, acc, 5, b, c)
    torch.cuda.synchronize()
    parts = acc.cpu().numpy()
    np.testing.assert_array_equal(parts[..., 3], whole[..., 3])
    np.testing.assert_allclose(parts[..., :3], whole[..., :3], rtol=5e-7, atol=1e-6)
    # 1-spp slices ARE bit-identical to the oracle's samples taken one at a time
    one = gpu_render(torch, ctx, info, 5, 6, 1)
    want = scenes.oracle_render(sc, sph, info, 5, 6, 1)
    assert_radiance_close(one[..., :3], want[..., :3])


def test_launch_is_a_drop_in_for_launchKernel(gpt, torch_cuda):
    """gpt_launch with the reference's kernelInfo contract: device pointers inside the struct (BasicScene.cpp:294-313),
    hash(frame) seed, constantPdf running mean with per-frame clamp, RGBA8 pack (tracer.cu:386-398)."""
    torch = torch_cuda
    L = reflib.oracle()
    sc = scenes.load_scene("bunny_cornell")
    sph = gpt.default_spheres()
    w, h = 160, 90
    d_nodes, d_tris, d_idx = dev(torch, sc["nodes"]), dev(torch, sc["tris"]), dev(torch, sc["indices"])
    d_sph = torch.from_numpy(np.frombuffer(bytes(sph), dtype=np.uint8).copy()).cuda()
    info = small_info(gpt, w, h, depth=4)
    info.bvhData.dev_triNode = d_nodes.data_ptr(); info.bvhData.triNodeSize = sc["nodes"].shape[0]
    info.bvhData.dev_triPtr = d_tris.data_ptr(); info.bvhData.triSize = sc["tris"].shape[0]
    info.bvhData.dev_triIndicesPtr = d_idx.data_ptr(); info.bvhData.triIndicesSize = sc["indices"].shape[0]
    info.sphereTex = d_sph.data_ptr(); info.numSpheres = len(sph)
    accum = torch.zeros(h * w * 3, dtype=torch.float32, device="cuda")
    draw = torch.zeros(h * w, dtype=torch.int32, device="cuda")
    info.accumBuffer = accum.data_ptr(); info.dev_drawRes = draw.data_ptr()
    osph, ns = scenes.orc_spheres(sph)
    osc = scenes.orc_scene(sc, osph, ns)
    o_acc = np.zeros((h, w, 3), np.float32); o_draw = np.zeros((h, w), np.uint32)
    with gpt.Context(0) as ctx:                                     # NOTE: no gpt_upload_scene: pointers travel in kernelInfo
        for frame in range(4):
            info.hash = gpt.hash64(frame)
            info.constantPdf = frame + 1
            ctx.launch(info)
            ctx.sync()
            p = scenes.orc_params(info, info.hash)
            assert L.orc_launch_legacy(C.byref(osc), C.byref(p), C.c_uint64(frame + 1), o_acc.ctypes.data_as(C.c_void_p), o_draw.ctypes.data_as(C.c_void_p), 8) == 0
            got = accum.cpu().numpy().reshape(h, w, 3)
            assert_radiance_close(got, o_acc)
            same = np.all(got == o_acc, axis=-1)
            np.testing.assert_array_equal(draw.cpu().numpy().view(np.uint32).reshape(h, w)[same], o_draw[same])
        # reset semantics: constantPdf = 1 overwrites without clearing (BasicScene.cpp:399)
        info.hash = gpt.hash64(0); info.constantPdf = 1
        ctx.launch(info); ctx.sync()
        first = scenes.oracle_render(sc, sph, info, info.hash, 0, 1)[..., :3]
        assert_radiance_close(accum.cpu().numpy().reshape(h, w, 3), np.clip(first, 0, 1))


def test_basic_scene_mirror_runs_progressive_frames(gpt, torch_cuda):
    sc = scenes.load_scene("cornell")
    bs = gpt.BasicScene(128, 72, sc["nodes"], sc["tris"], sc["indices"])
    acc, draw = bs.run(3)
    assert bs.info.constantPdf == 3 and bs.frameNumber == 3
    a = acc.cpu().numpy()
    assert a.min() >= 0.0 and a.max() <= 1.0 and a.mean() > 0.05


def test_resolve_matches_oracle(gpt, torch_cuda, ctxs):
    torch = torch_cuda
    L = reflib.oracle()
    ctx, _, _ = ctxs("cornell")
    rng = np.random.default_rng(5)
    acc = (rng.random((37, 53, 4)) * 30).astype(np.float32)
    acc[..., 3] = 16.0
    acc[0, 0] = [1e4, -2, 16, 16]
    acc[1, 1] = [1, 2, 3, 0]                                      # count 0 -> 0
    rgba_o = np.zeros((37, 53), np.uint32); rgb_o = np.zeros((37, 53, 3), np.float32)
    L.orc_resolve(acc.ctypes.data_as(C.c_void_p), 53, 37, rgba_o.ctypes.data_as(C.c_void_p), rgb_o.ctypes.data_as(C.c_void_p))
    rgba = torch.zeros(37, 53, dtype=torch.int32, device="cuda"); rgb = torch.zeros(37, 53, 3, dtype=torch.float32, device="cuda")
    ctx.resolve(dev(torch, acc), 53, 37, rgba, rgb)
    torch.cuda.synchronize()
    np.testing.assert_array_equal(rgb.cpu().numpy(), rgb_o)
    np.testing.assert_array_equal(rgba.cpu().numpy().view(np.uint32), rgba_o)


# ------------------------------------------------------------------ errors
def test_error_behaviour(gpt, torch_cuda):
    torch = torch_cuda
    with gpt.Context(0) as ctx:
        info = small_info(gpt)
        acc = torch.zeros(info.height, info.width, 4, device="cuda")
        with pytest.raises(gpt.GptError, match="no scene"):
            ctx.render(info, acc, 0, 0, 1)
        sc = scenes.load_scene("cornell")
        bad = sc["nodes"].copy()
        bad.view(np.int32)[3, 0] = 0                              # child pointing back at the root: a cycle
        with pytest.raises(gpt.GptError, match="malformed"):
            ctx.upload_scene(bad, sc["tris"], sc["indices"], gpt.default_spheres())
        bad = sc["nodes"].copy()
        bad.view(np.int32)[3, 1] = ~(10 ** 6)                     # leaf far outside the triangle array
        with pytest.raises(gpt.GptError, match="malformed"):
            ctx.upload_scene(bad, sc["tris"], sc["indices"], gpt.default_spheres())
        ctx.upload_scene(sc["nodes"], sc["tris"], sc["indices"], gpt.default_spheres())
        info.width = 1
        with pytest.raises(gpt.GptError, match="width/height"):
            ctx.render(info, acc, 0, 0, 1)
        info = small_info(gpt)
        info.accumBuffer = None
        with pytest.raises(gpt.GptError, match="accumBuffer"):
            ctx.launch(info)


# ------------------------------------------------------------------ BASELINE config size: properties
def test_full_size_properties(gpt, torch_cuda, ctxs):
    """BASELINE configs[1] (bunny in Cornell box, 1080p, 16 spp, depth 8): too big for the CPU oracle in a test,
    so check what must hold at any size."""
    torch = torch_cuda
    ctx, sc, sph = ctxs("bunny_cornell")
    info = gpt.default_kernel_info(1920, 1080)
    info.depth = 8
    ctx.reset_stats()
    a = gpu_render(torch, ctx, info, 0, 0, 16)
    st = ctx.stats()
    assert st.paths == 1920 * 1080 * 16
    assert 1920 * 1080 * 16 <= st.rays <= 1920 * 1080 * 16 * 8
    assert (a[..., 3] == 16).all() and np.isfinite(a).all() and (a[..., :3] >= 0).all()
    b = gpu_render(torch, ctx, info, 0, 0, 16)
    np.testing.assert_array_equal(a, b)                          # run-to-run deterministic (no float atomics)
    # a 64x64 crop must equal the oracle evaluated on just those pixels' samples
    L = reflib.oracle()
    osph, ns = scenes.orc_spheres(sph)
    osc = scenes.orc_scene(sc, osph, ns)
    p = scenes.orc_params(info, 0)
    rgb = (C.c_float * 3)()
    rng = np.random.default_rng(0)
    bad = 0
    for _ in range(300):
        px, py = int(rng.integers(0, 1920)), int(rng.integers(0, 1080))
        s = np.zeros(3, np.float32)
        for k in range(16):
            L.orc_sample(C.byref(osc), C.byref(p), px, py, C.c_uint64(k), rgb, None)
            s = (s + np.array(rgb[:], np.float32)).astype(np.float32)
        bad += int(not np.array_equal(s, a[py, px, :3]))
    assert bad == 0, bad


# ------------------------------------------------------------------ wavefront pipeline == megakernel == oracle
@pytest.mark.parametrize("name,tri_mat,depth,w,h", [("bunny_cornell", 0, 8, 160, 90), ("bunny_cornell", 1, 5, 160, 90),
                                                    ("bunny_cornell", 3, 6, 131, 67), ("cornell", 2, 1, 64, 36),
                                                    ("sphere_mesh", 0, 4, 96, 54), (None, 0, 5, 96, 54)])
def test_wavefront_pipeline_bit_identical(gpt, torch_cuda, ctxs, name, tri_mat, depth, w, h):
    """Both schedules evaluate the same per-sample arithmetic: images must be IDENTICAL, and equal to the oracle."""
    torch = torch_cuda
    ctx_m, sc, sph = ctxs(name, flags=gpt.FLAG_MEGAKERNEL)
    ctx_w, _, _ = ctxs(name)
    info = small_info(gpt, w, h, depth=depth)
    info.triCurrentMat = tri_mat
    a = gpu_render(torch, ctx_m, info, 21, 1, 3)
    ctx_w.reset_stats()
    b = gpu_render(torch, ctx_w, info, 21, 1, 3)
    st = ctx_w.stats()
    np.testing.assert_array_equal(a, b)
    assert st.paths == 3 * w * h and st.rays >= 3 * w * h
    want = scenes.oracle_render(sc, sph, info, 21, 1, 3)
    assert_radiance_close(b[..., :3], want[..., :3])
    # depth 0 and an early-terminating scene (no spheres: misses end paths, queues shrink)
    if name == "cornell":
        empty = (gpt.Sphere * 0)()
        ctx_e, _, _ = ctxs("cornell", spheres=empty)
        info2 = small_info(gpt, 128, 72, depth=6)
        got = gpu_render(torch, ctx_e, info2, 2, 0, 2)
        want = scenes.oracle_render(sc, None, info2, 2, 0, 2)
        assert_radiance_close(got[..., :3], want[..., :3])
        info2.depth = 0
        assert (gpu_render(torch, ctx_e, info2, 2, 0, 2)[..., :3] == 0).all()


def test_both_pipelines_legacy_launch(gpt, torch_cuda):
    torch = torch_cuda
    sc = scenes.load_scene("bunny_cornell")
    w, h = 160, 90
    outs = []
    for flags in (gpt.FLAG_MEGAKERNEL, 0):
        bs = gpt.BasicScene(w, h, sc["nodes"], sc["tris"], sc["indices"], flags=flags)
        acc, draw = bs.run(3)
        outs.append((acc.cpu().numpy().copy(), draw.cpu().numpy().copy()))
    np.testing.assert_array_equal(outs[0][0], outs[1][0])
    np.testing.assert_array_equal(outs[0][1], outs[1][1])


def test_per_triangle_materials_match_oracle(gpt, torch_cuda):
    """N4: a material table indexed by the remapped OBJ face index replaces the single global triangle material
    (tracer.cu:131-135); faces outside the table and a cleared table fall back to info.col/emi/triCurrentMat."""
    torch = torch_cuda
    sc = scenes.load_scene("bunny_cornell")
    n_faces = len(sc["faces"])
    mats = (gpt.TriMaterial * 5)()
    for m, (col, emi, mat) in zip(mats, [((0.8, 0.2, 0.2), (0, 0, 0), gpt.DIFF), ((1, 1, 1), (3, 3, 2.5), gpt.DIFF), ((0.9, 0.9, 0.9), (0, 0, 0), gpt.SPEC),
                                         ((0.7, 0.8, 0.9), (0, 0, 0), gpt.METAL), ((1, 1, 1), (0, 0, 0), gpt.REFR)]):
        m.col[:] = col; m.emi[:] = emi; m.mat = mat
    rng = np.random.default_rng(4)
    face_mat = rng.integers(0, 5, n_faces).astype(np.int32)
    face_mat[::7] = -1                                              # "no usemtl": global material
    face_mat = face_mat[: n_faces - 100]                            # table shorter than the mesh: tail uses the global material
    sph = gpt.default_spheres()
    info = small_info(gpt, 160, 90, depth=7)
    info.triCurrentMat = gpt.DIFF
    for flags in (0, gpt.FLAG_MEGAKERNEL):
        with gpt.Context(0, flags) as ctx:
            ctx.upload_scene(sc["nodes"], sc["tris"], sc["indices"], sph)
            ctx.set_triangle_materials(mats, face_mat)
            got = gpu_render(torch, ctx, info, 21, 0, 3)
            want = scenes.oracle_render(sc, sph, info, 21, 0, 3, tri_materials=mats, face_material=face_mat)
            assert_radiance_close(got[..., :3], want[..., :3])
            plain = scenes.oracle_render(sc, sph, info, 21, 0, 3)
            assert np.mean(np.any(want != plain, axis=-1)) > 0.2    # the table really changes the image
            ctx.set_triangle_materials(None, None)
            assert_radiance_close(gpu_render(torch, ctx, info, 21, 0, 3)[..., :3], plain[..., :3])


def test_small_calls_take_the_megakernel_same_image(gpt, torch_cuda):
    """Default flags, GPT_SMALL_FRAME_PATHS unset (a fresh process: read once, conftest sets it to 0 for the other tests): a call of
    fewer than 393216 paths runs as ONE megakernel launch, a larger one through the wavefront pipeline (many launches); forcing
    either pipeline gives the identical frame; gpt_launch (legacy contract) switches the same way."""
    import json
    import os
    import subprocess
    import sys
    code = r'''
import json, sys, numpy as np, torch
sys.path.insert(0, %r); sys.path.insert(0, %r)
import gpt_b200 as gpt
from support import scenes
sc = scenes.load_scene("cornell")
res = {}
for label, (w, h) in (("small", (512, 512)), ("large", (768, 576))):
    info = gpt.default_kernel_info(w, h); info.depth = 4; info.cam.dist = 12.0 * h / 720.0
    frames, launches = [], []
    for flags in (0, gpt.FLAG_WAVEFRONT, gpt.FLAG_MEGAKERNEL):
        with gpt.Context(0, flags) as ctx:
            ctx.upload_scene(sc["nodes"], sc["tris"], sc["indices"], gpt.default_spheres(), None)
            acc = torch.zeros(h, w, 4, dtype=torch.float32, device="cuda")
            ctx.render(info, acc, 11, 0, 1)
            torch.cuda.synchronize()
            launches.append(ctx.stats().kernel_launches); frames.append(acc.cpu().numpy())
    res[label] = {"launches": launches, "same": bool(np.array_equal(frames[0], frames[1]) and np.array_equal(frames[0], frames[2])),
                  "lit": float(frames[0][..., :3].max())}
print(json.dumps(res))
''' % (os.path.dirname(os.path.dirname(os.path.abspath(__file__))), os.path.dirname(os.path.abspath(__file__)))
    env = {k: v for k, v in os.environ.items() if k != "GPT_SMALL_FRAME_PATHS"}
    out = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    r = json.loads(out.stdout.strip().splitlines()[-1])
    for label in ("small", "large"):
        assert r[label]["same"] and r[label]["lit"] > 0.0, r
    d, wf, mk = r["small"]["launches"]
    assert d == mk and wf > mk, r                      # 262144 paths: the default is the megakernel
    d, wf, mk = r["large"]["launches"]
    assert d == wf and wf > mk, r                      # 442368 paths: the default is the wavefront pipeline


# ------------------------------------------------------------------ round 2: pointer-identity rule, depth limit, NO_TRIAL
def test_invalidate_scene_after_in_place_rewrite(gpt, torch_cuda):
    """gpt_launch traverses a re-packed copy cached by pointer identity (gpt.h, POINTER-IDENTITY RULE): rewriting the caller's BVH
    arrays IN PLACE is invisible until gpt_invalidate_scene(); then the next launch re-packs and equals the oracle on the new arrays."""
    torch = torch_cuda
    sc = scenes.load_scene("bunny_cornell")
    sph = gpt.default_spheres()
    w, h = 128, 72
    # a second tree with the same array sizes: the same mesh moved 2.5 units in x (child boxes and vertices; the leaf terminators
    # -- rows whose x is the bit pattern of -0.0, cudaUtils.h:413 -- stay as they are)
    nodes2, tris2 = sc["nodes"].copy(), sc["tris"].copy()
    n = nodes2.reshape(-1, 4, 4)
    n[:, 0, 0:2] += 2.5
    n[:, 1, 0:2] += 2.5
    term = tris2.view(np.uint32)[:, 0] == 0x80000000
    tris2[~term, 0] += 2.5
    sc2 = dict(sc, nodes=nodes2, tris=tris2)
    d_nodes, d_tris, d_idx = dev(torch, sc["nodes"]), dev(torch, sc["tris"]), dev(torch, sc["indices"])
    d_sph = torch.from_numpy(np.frombuffer(bytes(sph), dtype=np.uint8).copy()).cuda()
    info = small_info(gpt, w, h, depth=3)
    info.bvhData.dev_triNode = d_nodes.data_ptr(); info.bvhData.triNodeSize = sc["nodes"].shape[0]
    info.bvhData.dev_triPtr = d_tris.data_ptr(); info.bvhData.triSize = sc["tris"].shape[0]
    info.bvhData.dev_triIndicesPtr = d_idx.data_ptr(); info.bvhData.triIndicesSize = sc["indices"].shape[0]
    info.sphereTex = d_sph.data_ptr(); info.numSpheres = len(sph)
    accum = torch.zeros(h * w * 3, dtype=torch.float32, device="cuda")
    info.accumBuffer = accum.data_ptr()
    info.hash, info.constantPdf = gpt.hash64(7), 1

    def frame():
        ctx.launch(info); ctx.sync()
        return accum.cpu().numpy().reshape(h, w, 3).copy()
    with gpt.Context(0) as ctx:
        a = frame()
        want_a = np.clip(scenes.oracle_render(sc, sph, info, info.hash, 0, 1)[..., :3], 0, 1)
        assert np.array_equal(a, want_a)
        d_nodes.copy_(dev(torch, nodes2)); d_tris.copy_(dev(torch, tris2))      # same addresses, same sizes, new content
        torch.cuda.synchronize()
        assert np.array_equal(frame(), a)                            # documented: the stale packed tree is still traversed
        ctx.invalidate_scene()
        b = frame()
        want_b = np.clip(scenes.oracle_render(sc2, sph, info, info.hash, 0, 1)[..., :3], 0, 1)
        assert not np.array_equal(b, a)
        assert np.array_equal(b, want_b)


def test_tree_deeper_than_the_stack_is_rejected(gpt, torch_cuda):
    """A chain of 70 inner nodes (each: one leaf, one inner child): valid CudaBVH layout, but 70 levels would overrun the 64-entry
    traversal stack (cudaUtils.h:22,259) -- GPT_ERR_LAYOUT instead of a silent out-of-bounds write; 60 levels are accepted."""
    def chain(levels):
        nodes = np.zeros((levels * 4, 4), np.float32)
        tris, idx = [], []
        ni = nodes.view(np.int32)
        for i in range(levels):
            z = -30.0 - i
            for leaf in range(2 if i == levels - 1 else 1):
                slot = len(tris)
                x0 = float(leaf)
                tris += [[x0, 0, z, 0], [x0 + 1, 0, z, 0], [x0, 1, z, 0]]
                idx += [len(idx) // 4, 0, 0]
                t = np.zeros(4, np.float32); t.view(np.uint32)[0] = 0x80000000
                tris.append(list(t)); idx.append(0)
                ni[4 * i + 3, leaf if i == levels - 1 else 0] = ~slot
            if i < levels - 1:
                ni[4 * i + 3, 1] = 64 * (i + 1)
            nodes[4 * i + 0] = [-1, 3, -1, 3]; nodes[4 * i + 1] = [-1, 3, -1, 3]; nodes[4 * i + 2] = [-200, 0, -200, 0]
        t = np.array(tris, np.float32)
        t.view(np.uint32)[3::4, 0] = 0x80000000
        return nodes, t, np.array(idx, np.int32)
    with gpt.Context(0) as ctx:
        n, t, i = chain(60)
        ctx.upload_scene(n, t, i, gpt.default_spheres())
        n, t, i = chain(70)
        with pytest.raises(gpt.GptError, match="traversal stack"):
            ctx.upload_scene(n, t, i, gpt.default_spheres())


def test_no_trial_flag_keeps_calls_asynchronous(gpt, torch_cuda):
    """GPT_FLAG_NO_TRIAL: a scene large enough for the timed per-scene trial never enters it (no render call synchronises on its
    account), and the image equals the default context's."""
    torch = torch_cuda
    from support import procedural
    v, f = procedural.synthetic_scene(120000, seed=3)
    info = gpt.default_kernel_info(1280, 1024)                       # > 1 Mi paths per call: trial territory
    info.depth = 3
    out = []
    for flags in (gpt.FLAG_NO_TRIAL, 0):
        with gpt.Context(0, flags) as ctx:
            b = ctx.build_bvh(v, f)
            ctx.upload_built(b, gpt.default_spheres())
            acc = torch.zeros(1024, 1280, 4, dtype=torch.float32, device="cuda")
            pend0 = ctx.trial_info()[0]
            ctx.render(info, acc, 1, 0, 1)
            torch.cuda.synchronize()
            out.append((pend0, ctx.trial_info()[0], acc.cpu().numpy()))
            ctx.free_bvh(b)
    assert out[0][0] is False and out[0][1] is False                 # never pending with the flag
    assert out[1][0] is True                                         # the default context does put this scene on trial
    assert np.mean(np.any(out[0][2] != out[1][2], axis=-1)) <= 1e-5  # binary vs (possibly) 4-wide candidate: t-ties only
