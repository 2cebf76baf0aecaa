#!/usr/bin/env python
"""Mints tests/golden/ref_frames.npz: frames rendered by the REFERENCE's own kernel (oracle/_ref/reftrace.cubin, built from the
reference sources in place by oracle/Makefile) for the cases in tests/support/ref_frames.py.  Needs a GPU:
    gpurun -- 'python tests/golden/make_ref_frames.py'      (writes gpurun_out/ref_frames.npz; copy it to tests/golden/)
The CPU test tests/test_oracle_ref_frames.py then pins the oracle (XORWOW replay, ORC_FLAG_REF_RNG) to these frames."""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, ROOT)
import gpt_b200 as gpt  # noqa: E402
from support import ref_frames, refgpu  # noqa: E402

refk = refgpu.RefGpuKernel(0)
out = {}
for case in ref_frames.CASES:
    out[ref_frames.key(case)] = ref_frames.reference_frame(gpt, torch, refgpu, refk, case)
refk.close()
out["meta"] = np.array([ref_frames.W, ref_frames.H], dtype=np.int32)
dst = os.path.join(ROOT, "gpurun_out")
os.makedirs(dst, exist_ok=True)
np.savez_compressed(os.path.join(dst, "ref_frames.npz"), **out)
print("wrote", len(out) - 1, "frames, device", torch.cuda.get_device_name(0))
