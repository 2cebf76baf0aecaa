#!/usr/bin/env python
"""One progressive frame of a scene bracketed by cudaProfilerStart/Stop, for a counter pass over EVERY kernel of the frame:

  ncu --profile-from-start off --clock-control none --csv --log-file gpurun_out/ncu_frame_<scene>.csv \
      --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum,\
lts__t_bytes.sum,l1tex__t_bytes.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,l1tex__t_sector_hit_rate.pct,lts__t_sector_hit_rate.pct \
      python profiles/tools/ncu_frame.py <scene> [spp]

profiles/tools/ncu_frame_summary.py turns the CSV into per-kernel-class totals per ray (profiles/r02_frame_counters.json), which
bench.py combines with its LIVE kernel times into the dram / l2 / issue-slot fractions of its roofline.
Scenes: bench (bench.py's workload) or any scene of tests/support/config_check.py.  Without ncu the script just renders."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import gpt_b200 as gpt  # noqa: E402
from support import config_check, procedural, scenes  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "bench"
    spp = int(sys.argv[2]) if len(sys.argv) > 2 else 16
    if name == "bench":
        import bench
        sc = bench.load_scene_arrays()
        sph = bench.bench_spheres(gpt)
        verts, faces, ref, W, H, mat = sc["verts"], sc["faces"], sc, 1920, 1080, gpt.DIFF
    else:
        verts, faces, ref, W, H, mat = config_check.make_scene(name, gpt, procedural, scenes)
        sph = gpt.default_spheres()
    info = gpt.default_kernel_info(W, H)
    info.depth = 8
    info.triCurrentMat = mat
    with gpt.Context(0, 0) as ctx:
        built = None
        if ref is not None:
            ctx.upload_scene(ref["nodes"], ref["tris"], ref["indices"], sph)
        else:
            built = ctx.build_bvh(verts, faces)
            ctx.upload_built(built, sph)
        acc = torch.zeros(H, W, 4, dtype=torch.float32, device="cuda")
        for w in range(16):                               # buffers + the per-scene trial (<= 12 calls)
            ctx.render(info, acc, 0, 1_000_000 + w * spp, spp)
            if w >= 2 and not ctx.trial_info()[0]:
                break
        torch.cuda.synchronize()
        ctx.reset_stats()
        torch.cuda.profiler.start()
        ctx.render(info, acc, 0, 0, spp)
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
        st = ctx.stats()
        out = {"scene": name, "width": W, "height": H, "spp": spp, "depth": 8, "triangles": int(len(faces)), "rays": int(st.rays), "paths": int(st.paths),
               "launches": int(st.kernel_launches), "tree_in_use": ctx.tree_info()["in_use"], "reorder": ctx.reorder_info()[0]}
        print(json.dumps(out))
        d = os.path.join(ROOT, "gpurun_out")
        if os.path.isdir(d):
            with open(os.path.join(d, "ncu_frame_%s.json" % name), "w") as fh:
                json.dump(out, fh)
        if built is not None:
            ctx.free_bvh(built)


if __name__ == "__main__":
    main()
