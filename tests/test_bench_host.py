"""CPU-side checks of bench.py's host logic: both arms print the same `config` dict, the bound fractions are what their inputs say,
and the committed counter / byte files bench.py reads have the fields it reads."""
import json
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


class Args:
    width, height, depth, spp, scaling, frame_spp, gpus = 1920, 1080, 8, 16, "weak", 16, 1


def test_both_arms_describe_the_same_workload():
    a = Args()
    ours = bench.workload_config(a, 1)
    ref = bench.workload_config(a, max(1, a.gpus))
    assert ours == ref and ours["workload"] == bench.WORKLOAD and ours["frame_spp"] == 16
    a.scaling, a.frame_spp = "strong", 16
    s8 = bench.workload_config(a, 8)
    assert s8["frame_spp"] == 16 and s8["spp_per_gpu_per_step"] == 2 and s8["scaling"] == "strong"
    a.scaling = "weak"
    assert bench.workload_config(a, 8)["frame_spp"] == 128


def test_bound_fractions_arithmetic():
    cls = {"dram_bytes_per_ray": 40.0, "l2_bytes_per_ray": 100.0, "l1_bytes_per_ray": 300.0, "warp_inst_per_ray": 50.0,
           "lanes_per_instruction": 15.0, "issue_active_pct": 80.0, "l1_hit_pct": 82.0, "l2_hit_pct": 43.0}
    peaks = {"stream_l2_gbs": 30000.0, "stream_l1_gbs": 36000.0, "gather_l2_gbs": 8000.0, "gather_l1_gbs": 12000.0, "warp_inst_per_s": 1.1e12}
    b = bench.bound_fractions(cls, rays=1e9, ms=100.0, peaks=peaks, hbm=6500.0)          # 1e10 rays/s
    assert b["dram"]["achieved"] == pytest.approx(400.0) and b["dram"]["frac"] == pytest.approx(400.0 / 6500.0)
    assert b["l2"]["frac"] == pytest.approx(1000.0 / 30000.0) and b["l2"]["frac_of_random_gather_rate"] == pytest.approx(1000.0 / 8000.0)
    assert b["issue"]["frac"] == pytest.approx(0.5e12 / 1.1e12) and b["binding"] == "issue"
    assert all(v["frac"] <= 1.0 for k, v in b.items() if isinstance(v, dict) and "frac" in v)
    assert bench.bound_fractions(None, 1, 1, peaks, 1) is None


def test_committed_counter_files_have_what_bench_reads():
    ab = bench.load_algo_bytes("bench")
    assert 500 < ab["traversal_bytes_per_ray"] < 700 and ab["bytes_per_ray"] > ab["traversal_bytes_per_ray"]
    p = os.path.join(ROOT, "profiles", "r02_frame_counters.json")
    if os.path.exists(p):
        fc = json.load(open(p))
        for scene in ("bench", "cfg2_dragon100k_sbvh", "cfg2_dragon1p6M", "cfg4_20M"):
            for k in ("extend", "shade"):
                c = fc[scene]["classes"][k]
                for key in ("warp_inst_per_ray", "dram_bytes_per_ray", "l2_bytes_per_ray", "l1_bytes_per_ray", "lanes_per_instruction"):
                    assert c[key] > 0, (scene, k, key)
            assert bench.load_algo_bytes(scene) is not None
