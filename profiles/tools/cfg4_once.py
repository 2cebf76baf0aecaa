#!/usr/bin/env python
"""One render of BASELINE configs[4] (synthetic 20 M triangles, 4K, depth 8, device-built LBVH) for ncu captures:
  ncu --set full -k regex:k_wf_extend_ww -s 1 -c 1 python profiles/tools/cfg4_once.py [n_tris] [spp]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import gpt_b200 as gpt  # noqa: E402
from support import procedural  # noqa: E402

n_tris = int(sys.argv[1]) if len(sys.argv) > 1 else 20_000_000
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 4
passes = int(sys.argv[3]) if len(sys.argv) > 3 else 2      # the first large calls of a scene are its tree / ray-order trial (gpt_trial_info)
v, f = procedural.synthetic_scene(n_tris, seed=1234)
info = gpt.default_kernel_info(3840, 2160)
info.depth = 8
with gpt.Context(0, gpt.FLAG_TIME_KERNELS) as ctx:
    b = ctx.build_bvh(v, f)
    ctx.upload_built(b, gpt.default_spheres())
    acc = torch.zeros(2160, 3840, 4, dtype=torch.float32, device="cuda")
    for i in range(passes):
        ctx.reset_stats()
        ctx.render(info, acc, 0, i * spp, spp)
        torch.cuda.synchronize()
        st, kt = ctx.stats(), ctx.kernel_times()
        print("pass %d: %d rays, %s, wide=%s" % (i, st.rays, {k: round(x[0], 2) for k, x in kt.items() if x[1]}, ctx.wide_info()[0] > 0))
    ctx.free_bvh(b)
