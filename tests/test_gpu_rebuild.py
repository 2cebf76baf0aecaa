"""-m gpu: the rebuilt tree (gpt_rebuild_info) -- the triangle references of an uploaded reference-layout BVH under a tree of the
library's own device builder -- against the CPU oracle on the arrays as uploaded (tests/support/rebuild_check.py, one subprocess per
case: the policy is read from the environment once per process, and conftest.py pins GPT_REBUILD=0 for every other test)."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))


def run_check(scene, *extra, rebuild=None, split=None, timeout=900):
    env = {k: v for k, v in os.environ.items() if not k.startswith("GPT_")}
    if rebuild is not None:
        env["GPT_REBUILD"] = str(rebuild)
    if split is not None:
        env["GPT_SPLIT"] = str(split)
    if rebuild is not None or split is not None:
        env["GPT_SMALL_FRAME_PATHS"] = "0"          # small frames through the wavefront pipeline, like the other parity tests
    out = subprocess.run([sys.executable, os.path.join(HERE, "support", "rebuild_check.py"), scene, *extra],
                         env=env, capture_output=True, text=True, timeout=timeout)
    lines = [ln for ln in out.stdout.strip().splitlines() if ln.startswith("{")]
    assert lines, "rebuild_check.py printed no result (rc %d): %s" % (out.returncode, out.stderr[-3000:])
    r = json.loads(lines[-1])
    print(json.dumps(r))
    d = os.path.join(os.path.dirname(HERE), "gpurun_out")
    if os.path.isdir(d):
        with open(os.path.join(d, "rebuild_parity.jsonl"), "a") as fh:
            fh.write(json.dumps(r) + "\n")
    assert out.returncode == 0 and r["ok"], r["fails"]
    return r


@pytest.mark.parametrize("scene", ["bunny_cornell", "dragon_cornell", "gto_sixteen", "sphere_mesh"])
def test_rebuilt_tree_gives_the_oracle_hits_and_frames(torch_cuda, scene):
    """GPT_REBUILD=1: every call traverses the rebuilt tree; distances bit for bit, faces and frames equal but for t-ties"""
    r = run_check(scene, rebuild=1)
    assert r["rebuilt"]["state"] == 1 and r["rays"]["t_mismatch"] == 0 and r["primary"]["t_mismatch"] == 0


@pytest.mark.parametrize("scene", ["bunny_cornell", "dragon_cornell", "gto_sixteen", "sphere_mesh"])
def test_split_leaf_tree_gives_the_oracle_hits_and_frames(torch_cuda, scene):
    """GPT_SPLIT=1: every call traverses the uploaded tree with its leaves split (gpt_tree_info); same checks"""
    r = run_check(scene, split=1)
    assert r["trees"]["in_use"] == "split" and r["rays"]["t_mismatch"] == 0 and r["primary"]["t_mismatch"] == 0


def test_default_policy_times_the_trees_and_keeps_the_image(torch_cuda):
    """no GPT_* variable: the 5 000-triangle bench scene gets a rebuilt and a split-leaf tree at upload and a trial of 3 x 3 large calls"""
    r = run_check("bunny_cornell", "--trial")
    assert r["trial"]["calls"] == 9 and r["trial"]["rebuilt"]["build_ms"] > 0.0


def test_default_policy_with_the_wide_tree_as_fourth_candidate(torch_cuda):
    """100 k triangles on the reference's SBVH: uploaded binary tree, its 4-wide collapse, the rebuilt and the split-leaf tree, three calls each"""
    r = run_check("dragon_cornell", "--trial")
    assert r["trial"]["calls"] == 12
