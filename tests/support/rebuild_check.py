#!/usr/bin/env python
"""The rebuilt tree (gpt_rebuild_info: the same triangle references under a tree of the library's own builder) against the CPU oracle
executing the reference algorithm on the arrays AS UPLOADED.  Run as a subprocess by tests/test_gpu_rebuild.py with GPT_REBUILD=1
(the rebuilt tree is always traversed) or with no GPT_* variable at all (--trial: the library's own policy, a timed trial).

  python tests/support/rebuild_check.py <fixture scene> [--trial]

Checks (fixture scenes are the reference's SBVH arrays, spatial-split duplicates included):
  rays      20 000 incoherent rays: t bit for bit; face index equal wherever no second triangle lies at the same t; the reported slot
            holds the reported face (it may be the other copy of a split triangle)
  primary   the 640x360 primary-hit map: geometry type, t, face index
  frames    a 200x150 x 6 spp x depth 7 frame equals the oracle's bit for bit in every pixel but t-ties (<= 2 pixels), with the global
            triangle material and with a per-face material table
  trial     (--trial) 1080p x 4 spp calls until the trial has ended: the rebuilt tree was timed, every call's frame equal up to ties
Prints one JSON object; exit code 0 iff every check passed."""
import argparse
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(HERE))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("scene")
    ap.add_argument("--trial", action="store_true")
    args = ap.parse_args()
    import torch
    import gpt_b200 as gpt
    from support import reflib, scenes
    from support.config_check import incoherent_rays
    if not torch.cuda.is_available():
        raise SystemExit("rebuild_check.py needs a GPU: the product has no CPU path")
    sc = scenes.load_scene(args.scene)
    sph = gpt.default_spheres()
    out = {"scene": args.scene, "env": {k: v for k, v in os.environ.items() if k.startswith("GPT_")}, "triangles": int(len(sc["faces"]))}
    fails = []
    ctx = gpt.Context(0, 0)
    ctx.upload_scene(sc["nodes"], sc["tris"], sc["indices"], sph)
    keep = gpt.Context(0, gpt.FLAG_KEEP_TREE)
    keep.upload_scene(sc["nodes"], sc["tris"], sc["indices"], sph)
    out["rebuilt"] = ctx.rebuild_info()
    out["trees"] = ctx.tree_info()
    out["rebuilt_under_keep_tree"] = keep.rebuild_info()
    kt = keep.tree_info()
    if keep.rebuild_info()["inner_nodes"] != 0 or keep.rebuild_info()["state"] != 0 or kt["inner_nodes"]["split"] != 0 or kt["in_use"] not in ("binary", "4-wide"):
        fails.append("GPT_FLAG_KEEP_TREE still built a second tree")
    if out["rebuilt"]["inner_nodes"] <= 0:
        fails.append("no rebuilt tree")
    forced = "split" if os.environ.get("GPT_SPLIT") == "1" else "rebuilt"
    if forced == "split" and out["trees"]["inner_nodes"]["split"] <= out["trees"]["inner_nodes"]["binary"]:
        fails.append("no split-leaf tree")

    if not args.trial:
        if out["trees"]["in_use"] != forced:
            fails.append("the %s tree is forced by the environment but %s is in use" % (forced, out["trees"]["in_use"]))
        # ---- rays
        n = 20000
        o, d = incoherent_rays(sc["verts"], n, 11)
        tri_o, addr_o, t_o, _ = scenes.oracle_trace_rays(sc, o, d, cull=True)
        tri = torch.empty(n, dtype=torch.int32, device="cuda"); addr = torch.empty_like(tri)
        t = torch.empty(n, dtype=torch.float32, device="cuda")
        ctx.trace_rays(torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda(), n, True, tri, addr, t)
        torch.cuda.synchronize()
        tri, addr, t = tri.cpu().numpy(), addr.cpu().numpy(), t.cpu().numpy()
        idx = np.asarray(sc["indices"])
        hit = t < 3e38
        slot_face_ok = bool((idx[addr[hit]] == tri[hit]).all()) if hit.any() else True
        out["rays"] = {"n": n, "hit_fraction": float(hit.mean()), "t_mismatch": int((t != t_o).sum()), "face_mismatch": int((tri != tri_o).sum()),
                       "slot_differs_same_face": int(((addr != addr_o) & (tri == tri_o)).sum()), "slot_holds_reported_face": slot_face_ok}
        if out["rays"]["hit_fraction"] < 0.05 or out["rays"]["t_mismatch"] or out["rays"]["face_mismatch"] > max(1, n // 5000) or not slot_face_ok:
            fails.append("rays differ from the oracle on the uploaded arrays")
        # ---- primary map
        pw, ph = 640, 360
        info = gpt.default_kernel_info(pw, ph)
        info.cam.dist = 12.0 * ph / 720.0
        ctx.set_flags(gpt.FLAG_NO_JITTER)
        geom_o, ptri_o, prim_o, pt_o, _ = scenes.oracle_primary(sc, sph, info, 0, 0, flags=reflib.ORC_FLAG_NO_JITTER)
        g = torch.empty(pw * ph, dtype=torch.int32, device="cuda"); pt_i = torch.empty_like(g); pp = torch.empty_like(g)
        ptt = torch.empty(pw * ph, dtype=torch.float32, device="cuda")
        ctx.trace_primary(info, 0, 0, g, pt_i, pp, ptt)
        torch.cuda.synchronize()
        ctx.set_flags(0)
        out["primary"] = {"mesh_pixels": int((geom_o == 0).sum()), "geom_mismatch": int((g.cpu().numpy() != geom_o).sum()),
                          "t_mismatch": int((ptt.cpu().numpy() != pt_o).sum()), "face_mismatch": int((pt_i.cpu().numpy() != ptri_o).sum())}
        if out["primary"]["geom_mismatch"] or out["primary"]["t_mismatch"] or out["primary"]["face_mismatch"] > 2:
            fails.append("primary-hit map differs")
        # ---- frames against the oracle, global material and per-face table
        info = gpt.default_kernel_info(200, 150)
        info.depth = 7
        info.cam.dist = 12.0 * 150 / 720.0
        n_faces = len(sc["faces"])
        mats = (gpt.TriMaterial * 4)()
        for m, (col, emi, mat) in zip(mats, [((0.8, 0.2, 0.2), (0, 0, 0), gpt.DIFF), ((1, 1, 1), (3, 3, 2.5), gpt.DIFF),
                                             ((0.9, 0.9, 0.9), (0, 0, 0), gpt.SPEC), ((0.7, 0.8, 0.9), (0, 0, 0), gpt.METAL)]):
            m.col[:] = col; m.emi[:] = emi; m.mat = mat
        face_mat = np.random.default_rng(4).integers(0, 4, n_faces).astype(np.int32)
        out["frames"] = {}
        for label, table in (("global_material", None), ("face_table", (mats, face_mat))):
            if table is not None:
                ctx.set_triangle_materials(*table)
            acc = torch.zeros(150, 200, 4, dtype=torch.float32, device="cuda")
            ctx.render(info, acc, 33, 0, 6)
            torch.cuda.synchronize()
            got = acc.cpu().numpy()
            want = scenes.oracle_render(sc, sph, info, 33, 0, 6, tri_materials=table[0] if table else None, face_material=table[1] if table else None)
            nd = int(np.any(got != want, axis=-1).sum())
            out["frames"][label] = {"pixels": 200 * 150, "differing": nd, "mean": float(got[..., :3].mean())}
            if nd > 2 or not np.isfinite(got).all():
                fails.append("frame (%s): %d pixels differ from the oracle" % (label, nd))
        ctx.set_triangle_materials(None, None)
        # ---- the entry point that allocates nothing walks the same tree: bit-identical to gpt_render
        acc_s = torch.zeros(150, 200, 4, dtype=torch.float32, device="cuda")
        acc_r = torch.zeros_like(acc_s)
        ctx.render_static(info, acc_s, 33, 6, 5)
        ctx.render(info, acc_r, 33, 6, 5)
        torch.cuda.synchronize()
        out["static_equals_render"] = bool(torch.equal(acc_s, acc_r)) and float(acc_r[..., 3].min()) == 5.0
        if not out["static_equals_render"]:
            fails.append("gpt_render_static differs from gpt_render on the rebuilt tree")
    else:
        W, H = 1920, 1080
        info = gpt.default_kernel_info(W, H)
        info.depth = 6
        acc = torch.zeros(H, W, 4, dtype=torch.float32, device="cuda")
        ref = torch.zeros_like(acc)
        keep.render(info, ref, 5, 0, 4)
        torch.cuda.synchronize()
        calls, worst, states = 0, 0, []
        if not ctx.trial_info()[0]:
            fails.append("no trial pending on a scene with a rebuilt tree")
        while calls < 16:
            acc.zero_()
            ctx.render(info, acc, 5, 0, 4)
            torch.cuda.synchronize()
            calls += 1
            worst = max(worst, int((acc != ref).any(dim=-1).sum().item()))
            states.append(ctx.rebuild_info()["state"])
            if not ctx.trial_info()[0] and calls >= 2:
                break
        # two more calls on whatever the trial chose
        for _ in range(2):
            acc.zero_()
            ctx.render(info, acc, 5, 0, 4)
            torch.cuda.synchronize()
            worst = max(worst, int((acc != ref).any(dim=-1).sum().item()))
        # ---- the megakernel's own trial: small calls (96 x 64 x 4 spp = 24 576 paths) until it has ended; frames equal to the uploaded tree's
        si = gpt.default_kernel_info(96, 64)
        si.depth = 5
        sref = torch.zeros(64, 96, 4, dtype=torch.float32, device="cuda")
        sacc = torch.zeros_like(sref)
        keep.render(si, sref, 7, 0, 4)
        mcalls, mworst = 0, 0
        while mcalls < 16:
            sacc.zero_()
            ctx.render(si, sacc, 7, 0, 4)
            torch.cuda.synchronize()
            mcalls += 1
            mworst = max(mworst, int((sacc != sref).any(dim=-1).sum().item()))
            if not ctx.megakernel_tree_info()["pending"]:
                break
        out["megakernel_trial"] = dict(ctx.megakernel_tree_info(), calls=mcalls, max_pixels_differing_from_uploaded_tree=mworst)
        if out["megakernel_trial"]["pending"] or mcalls != 9 or mworst > 2:
            fails.append("megakernel tree trial: %r" % (out["megakernel_trial"],))
        pend, ns = ctx.trial_info()
        out["trial"] = {"calls": calls, "pending_after": pend, "ns_per_ray_binary_wide_sorted": ns, "rebuilt": ctx.rebuild_info(), "trees": ctx.tree_info(),
                        "states_during": states, "max_pixels_differing_from_uploaded_tree": worst, "pixels": W * H}
        if pend or ns[0] <= 0.0 or out["trial"]["rebuilt"]["ns_per_ray"] <= 0.0:
            fails.append("the trial did not time both trees")
        if out["trial"]["rebuilt"]["state"] not in (0, 1) or out["trial"]["trees"]["pending"]:
            fails.append("rebuilt tree still on trial")
        if out["trial"]["trees"]["inner_nodes"]["split"] and out["trial"]["trees"]["ns_per_ray"]["split"] <= 0.0:
            fails.append("the split-leaf tree was built but not timed")
        if worst > max(2, W * H // 100000):
            fails.append("frames differ between the trees beyond t-ties: %d pixels" % worst)
    ctx.close(); keep.close()
    out["fails"], out["ok"] = fails, not fails
    print(json.dumps(out))
    return 0 if not fails else 1


if __name__ == "__main__":
    sys.exit(main())
