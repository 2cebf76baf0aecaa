"""-m gpu: the scenes of BASELINE.json configs 2, 3 and 4 against the CPU oracle (tests/support/config_check.py, one subprocess per
scene with NO GPT_* variable in its environment, so the library runs the traversal plan it picks by itself there -- the per-scene
trial between the binary tree, the 4-wide tree and ray reordering included).  Bit-exact: t / face index / slot of 20 000 incoherent
rays, the 1080p primary-hit map, and 256 random pixels x 16 spp x depth 8 of the config's full-size frame (1080p or 4K)."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))


def run_check(scene, *extra, timeout=1500):
    env = {k: v for k, v in os.environ.items() if not k.startswith("GPT_")}
    out = subprocess.run([sys.executable, os.path.join(HERE, "support", "config_check.py"), scene, *extra],
                         env=env, capture_output=True, text=True, timeout=timeout)
    lines = [ln for ln in out.stdout.strip().splitlines() if ln.startswith("{")]
    assert lines, "config_check.py printed no result (rc %d): %s" % (out.returncode, out.stderr[-3000:])
    r = json.loads(lines[-1])
    print(json.dumps(r))
    # kept beside the other artefacts when the box has a gpurun_out/ (profiles/r02_config_parity.jsonl is a copy of it)
    d = os.path.join(os.path.dirname(HERE), "gpurun_out")
    if os.path.isdir(d):
        with open(os.path.join(d, "config_parity.jsonl"), "a") as fh:
            fh.write(json.dumps(r) + "\n")
    assert out.returncode == 0 and r["ok"], r["fails"]
    return r


def test_cfg2_dragon_on_reference_sbvh(torch_cuda):
    """BASELINE configs[2] stand-in (cornell_dragon.obj is a missing blob): Assets/dragon.obj + cornell.obj on the arrays the
    REFERENCE's SplitBVHBuilder/CudaBVH built (CudaBVH.cpp:121-270) -- the tree north_star's target is quoted on"""
    r = run_check("cfg2_dragon100k_sbvh")
    assert r["triangles"] == 100032 and r["pixels"]["size"] == [1920, 1080]
    assert r["rays"]["t_mismatch"] == 0 and r["primary"]["face_mismatch"] == 0


def test_cfg2_dragon_subdivided_1p6M_device_tree(torch_cuda):
    r = run_check("cfg2_dragon1p6M")
    assert r["triangles"] == 16 * 100000 + 32 * 16
    assert r["pixels"]["render_calls"] >= 9          # the trial ran: tree, then ray order (from 256 Ki triangles), three calls each


def test_cfg3_gto_sixteen_instanced_4k(torch_cuda):
    r = run_check("cfg3_gto16")
    assert r["triangles"] == 16 * 9791 and r["pixels"]["size"] == [3840, 2160]


def test_cfg4_synthetic_20M_4k(torch_cuda):
    """BASELINE configs[4]: 20 M triangles (2.3 GB of arrays), device-built tree of depth ~38, 4-wide tree, node_k = 16"""
    r = run_check("cfg4_20M", timeout=2400)
    assert r["triangles"] > 19_000_000 and r["pixels"]["size"] == [3840, 2160]
