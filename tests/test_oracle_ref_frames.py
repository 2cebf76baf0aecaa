"""Pins the oracle to the reference: frames rendered by the reference's own CUDA kernel (tests/golden/ref_frames.npz, minted on a
B200 by tests/golden/make_ref_frames.py) against the oracle replaying the reference's XORWOW stream on the CPU.  Bit-exact
except where CUDA's powf (METAL, evaluated with the MUFU.RCP approximation unit) is an ulp from libm's -- see support/ref_frames.py."""
import os

import numpy as np
import pytest

import gpt_b200 as gpt
from support import ref_frames

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_frames.npz")


@pytest.fixture(scope="module")
def frames():
    assert os.path.exists(GOLD), "tests/golden/ref_frames.npz missing (mint it with tests/golden/make_ref_frames.py on a GPU box)"
    f = np.load(GOLD)
    assert tuple(f["meta"]) == (ref_frames.W, ref_frames.H)
    return f


@pytest.mark.parametrize("case", ref_frames.CASES, ids=ref_frames.key)
def test_oracle_reproduces_reference_kernel_frame(frames, case):
    ref = frames[ref_frames.key(case)]
    want = ref_frames.oracle_frame(gpt, case)
    r = ref_frames.compare(want, ref)
    assert ref.max() > 0.0, "reference frame is black: fixture broken"
    assert r["within_1e-5"] >= 1.0 - ref_frames.MAX_MISMATCH_FRACTION, r
    if case[1] != gpt.METAL and case[2] <= 5:
        # shallow paths away from powf: every pixel bit-identical
        assert r["mismatching_pixels"] == 0, r
    else:
        assert r["bit_identical"] >= 1.0 - ref_frames.MAX_MISMATCH_FRACTION, r


def test_argument_order_matters(frames):
    """the other evaluation order of the two curand_uniform arguments at tracer.cu:170 does NOT reproduce the frame -- the replay is
    a discriminating check, not a tautology"""
    from support import reflib, scenes
    case = ("cornell", 0, 4, 1, 0)
    sc = scenes.load_scene(case[0])
    acc = scenes.oracle_render(sc, ref_frames.make_spheres(gpt), ref_frames.make_info(gpt, case), ref_frames.seed_of(gpt, case), 0, 1,
                               flags=reflib.ORC_FLAG_REF_RNG | reflib.ORC_FLAG_REF_RNG_SWAP)
    r = ref_frames.compare(np.clip(acc[..., :3], 0.0, 1.0), frames[ref_frames.key(case)])
    assert r["bit_identical"] < 0.9, r
