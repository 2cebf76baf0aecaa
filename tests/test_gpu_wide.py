"""4-wide quantised BVH (g.p.u-pathtracer_b200/csrc/gpt_wide.cu) against the binary tree and the oracle (-m gpu).

The wide tree is collapsed on the device from the packed binary tree; its quantised boxes contain the binary boxes, so the
nearest hit (t and the remapped face index) must be IDENTICAL to the binary traversal / the oracle.  The triangle slot may be
the other copy of a spatially split triangle (bit-equal t, different visit order) -- nothing else may differ."""
import numpy as np
import pytest

from support import procedural, scenes

pytestmark = pytest.mark.gpu


def dev(torch, arr):
    return torch.from_numpy(np.ascontiguousarray(arr)).cuda()


def trace(torch, ctx, o, d, cull):
    n = len(o)
    tri = torch.empty(n, dtype=torch.int32, device="cuda"); addr = torch.empty_like(tri)
    t = torch.empty(n, dtype=torch.float32, device="cuda")
    ctx.trace_rays(dev(torch, o), dev(torch, d), n, cull, tri, addr, t)
    torch.cuda.synchronize()
    return tri.cpu().numpy(), addr.cpu().numpy(), t.cpu().numpy()


@pytest.mark.parametrize("name,n", [("cornell", 20000), ("sphere_mesh", 20000), ("bunny_cornell", 40000)])
@pytest.mark.parametrize("cull", [True, False])
def test_wide_trace_equals_oracle(gpt, torch_cuda, name, n, cull):
    torch = torch_cuda
    sc = scenes.load_scene(name)
    o, d = scenes.random_rays(sc, n, seed=23)
    d[:64] = 0.0; d[:64, 2] = -1.0            # axis-parallel directions: the 2^-80 guard against the quantised planes
    d[64:128] = 0.0; d[64:128, 0] = 1.0
    d[128:192] = 0.0; d[128:192, 1] = -1.0
    tri_o, addr_o, t_o, _ = scenes.oracle_trace_rays(sc, o, d, cull=cull)
    with gpt.Context(0, gpt.FLAG_WIDE_BVH) as ctx:
        ctx.upload_scene(sc["nodes"], sc["tris"], sc["indices"], gpt.default_spheres(), None)
        n_wide, levels = ctx.wide_info()
        assert 0 < n_wide <= len(sc["nodes"]) // 4 and 0 < levels <= 30
        tri, addr, t = trace(torch, ctx, o, d, cull)
    assert (t_o < 3e38).sum() > n // 20
    np.testing.assert_array_equal(t, t_o)                          # hit distance: bit exact
    np.testing.assert_array_equal(tri, tri_o)                      # remapped OBJ face index
    other_copy = addr != addr_o                                    # allowed only between copies of one split triangle
    assert other_copy.mean() <= 0.01
    assert np.array_equal(tri[other_copy], tri_o[other_copy])


def test_wide_flags_and_binary_default(gpt, torch_cuda):
    sc = scenes.load_scene("bunny_cornell")
    with gpt.Context(0) as ctx:                                    # small scene: binary by default
        ctx.upload_scene(sc["nodes"], sc["tris"], sc["indices"], gpt.default_spheres(), None)
        assert ctx.wide_info() == (0, 0)
        ctx.set_flags(gpt.FLAG_WIDE_BVH)
        n_wide, levels = ctx.wide_info()
        assert n_wide > 0 and levels > 0
        ctx.set_flags(gpt.FLAG_BINARY_BVH)
        assert ctx.wide_info() == (0, 0)
        ctx.set_flags(gpt.FLAG_WIDE_BVH | gpt.FLAG_MEGAKERNEL)     # the megakernel walks the binary tree
        assert ctx.wide_info() == (0, 0)
    with gpt.Context(0, gpt.FLAG_WIDE_BVH) as ctx:                 # sphere-only scene: nothing to collapse
        ctx.upload_scene(None, None, None, gpt.default_spheres(), None)
        assert ctx.wide_info() == (0, 0)


@pytest.mark.parametrize("name,mat,cull,depth", [("bunny_cornell", 0, 1, 6), ("bunny_cornell", 3, 0, 5), ("sphere_mesh", 2, 1, 4),
                                                 ("cornell", 1, 1, 5)])
def test_wide_render_equals_binary_render(gpt, torch_cuda, name, mat, cull, depth):
    """same frame through both trees, wavefront pipeline, native and legacy contract: bit-identical images"""
    torch = torch_cuda
    sc = scenes.load_scene(name)
    w, h = 240, 136
    info = gpt.default_kernel_info(w, h)
    info.cam.dist = 12.0 * h / 720.0
    if name == "sphere_mesh":
        gpt.set_camera(info, (0.0, 0.0, 30.0), 0.0, 0.0)
    info.depth, info.triCurrentMat, info.cullBackFaces = depth, mat, cull
    out = []
    for flags in (gpt.FLAG_BINARY_BVH, gpt.FLAG_WIDE_BVH):
        with gpt.Context(0, flags) as ctx:
            ctx.upload_scene(sc["nodes"], sc["tris"], sc["indices"], gpt.default_spheres(), None)
            acc = torch.zeros(h, w, 4, dtype=torch.float32, device="cuda")
            ctx.render(info, acc, 99, 0, 4)
            torch.cuda.synchronize()
            assert (ctx.wide_info()[0] > 0) == (flags == gpt.FLAG_WIDE_BVH)
            out.append(acc.cpu().numpy())
    assert out[0][..., :3].max() > 0
    differing = np.any(out[0] != out[1], axis=-1).mean()
    assert differing <= 1e-4, differing                            # a t-tie between two copies of a split triangle cannot change the image


def test_wide_on_built_lbvh_large_scene(gpt, torch_cuda):
    """1.2 M-triangle procedural scene, device-built LBVH: the wide tree's hits equal the oracle's on the binary arrays, and a
    frame equals the binary tree's frame"""
    torch = torch_cuda
    verts, faces = procedural.synthetic_scene(1_200_000, seed=77)
    n = 4000
    rng = np.random.default_rng(5)
    o = np.zeros((n, 3), np.float32)
    tgt = np.stack([rng.uniform(-18, 18, n), rng.uniform(-14, 14, n), rng.uniform(-300, -25, n)], 1)
    d = (tgt / np.linalg.norm(tgt, axis=1, keepdims=True)).astype(np.float32)
    w, h = 256, 144
    info = gpt.default_kernel_info(w, h)
    info.cam.dist = 12.0 * h / 720.0
    info.depth = 4
    frames = []
    for flags in (0, gpt.FLAG_BINARY_BVH):                         # 1.2 M triangles: the wide tree is the default
        with gpt.Context(0, flags) as ctx:
            b = ctx.build_bvh(verts, faces)
            ctx.upload_built(b, gpt.default_spheres())
            n_wide, levels = ctx.wide_info()
            if not flags:
                assert n_wide > 0 and 3 * levels + 2 <= 96, (n_wide, levels)
                assert n_wide < b.n_inner_nodes
                arr = ctx.built_to_host(b)
                host_scene = dict(verts=verts, faces=faces, **arr)
                tri_o, addr_o, t_o, _ = scenes.oracle_trace_rays(host_scene, o, d, cull=True)
                tri, addr, t = trace(torch, ctx, o, d, True)
                assert (t_o < 3e38).mean() > 0.5
                np.testing.assert_array_equal(t, t_o)
                np.testing.assert_array_equal(tri, tri_o)
                print("wide tree: %d nodes in %d levels over %d binary nodes (depth %d)" % (n_wide, levels, b.n_inner_nodes, b.depth))
            else:
                assert n_wide == 0
            acc = torch.zeros(h, w, 4, dtype=torch.float32, device="cuda")
            ctx.render(info, acc, 5, 0, 2)
            torch.cuda.synchronize()
            frames.append(acc.cpu().numpy())
            ctx.free_bvh(b)
    assert np.any(frames[0] != frames[1], axis=-1).mean() <= 1e-4
    # (GPT_SORT_MIN_TRIS=1000000 in the environment of this test run, see conftest) above that size the secondary rays are re-ordered before each extend (k_wf_sort_keys + radix sort): only the fetch
    # order changes, so the frame still equals the oracle's on the built arrays, pixel for pixel
    want = scenes.oracle_render(host_scene, gpt.default_spheres(), info, 5, 0, 2)
    assert np.any(frames[1] != want, axis=-1).mean() == 0.0


def test_traversal_trial_runs_and_keeps_images(gpt, torch_cuda, tmp_path):
    """Without GPT_WF_SORT / GPT_SORT_MIN_TRIS / GPT_WIDE_MIN_TRIS in the environment (a fresh process: they are read once) a scene
    of >= 256 Ki triangles goes through the nine-call trial -- warm-up + two timed calls (the faster counts) for the binary tree,
    the wide tree, and the faster of the two with rays re-ordered -- and then reports decisions; the SAME sample rendered under all candidates gives
    bit-identical frames."""
    import json
    import os
    import subprocess
    import sys
    code = r'''
import json, sys, numpy as np, torch
sys.path.insert(0, %r); sys.path.insert(0, %r)
import gpt_b200 as gpt
from support import procedural
v, f = procedural.synthetic_scene(300000, seed=9)
info = gpt.default_kernel_info(1024, 1024); info.depth = 4; info.cam.dist = 12.0 * 1024 / 720.0
with gpt.Context(0) as ctx:
    b = ctx.build_bvh(v, f); ctx.upload_built(b, gpt.default_spheres())
    pending, frames = [ctx.trial_info()[0]], []
    for call in range(10):
        acc = torch.zeros(1024, 1024, 4, dtype=torch.float32, device="cuda")
        ctx.render(info, acc, 3, 0, 1)                  # the same sample every call
        torch.cuda.synchronize()
        frames.append(acc.cpu().numpy()); pending.append(ctx.trial_info()[0])
    print(json.dumps({"pending": pending, "ns": ctx.trial_info()[1], "reorder": ctx.reorder_info()[0], "wide_nodes": ctx.wide_info()[0],
                      "same": bool(all(np.array_equal(frames[0], x) for x in frames[1:])), "lit": float(frames[0][..., :3].max())}))
    ctx.free_bvh(b)
''' % (os.path.dirname(os.path.dirname(os.path.abspath(__file__))), os.path.dirname(os.path.abspath(__file__)))
    env = {k: v for k, v in os.environ.items() if k not in ("GPT_WF_SORT", "GPT_SORT_MIN_TRIS", "GPT_WIDE_MIN_TRIS")}
    out = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    r = json.loads(out.stdout.strip().splitlines()[-1])
    assert r["pending"] == [True] * 9 + [False, False], r
    assert all(x > 0.0 for x in r["ns"]), r
    assert r["reorder"] in (0, 1)
    assert r["same"] and r["lit"] > 0.0, r
