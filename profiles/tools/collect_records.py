#!/usr/bin/env python
"""Copies the artefacts of a run_r02_m.sh call from gpurun_out/ into profiles/ (the committed records bench.py and the docs cite) and
regenerates the two ncu summaries.  usage: collect_records.py <call tag, e.g. m2>"""
import csv
import json
import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
G, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
tag = sys.argv[1]
dst = os.path.join(P, "r02_call_" + tag)
os.makedirs(dst, exist_ok=True)
for f in os.listdir(G):
    if f.endswith(".ncu-rep") or f.startswith("."):
        continue
    if f.startswith("ncu_frame_") and f.endswith(".csv"):
        shutil.copy(os.path.join(G, f), os.path.join(P, "r02_" + f))
        continue
    shutil.copy(os.path.join(G, f), os.path.join(dst, f))
for src, name in (("frame_counters.json", "r02_frame_counters.json"), ("algo_bytes.json", "r02_algorithmic_bytes.json"), ("bench.json", "r02_final_bench.json"),
                  ("bench_reference_arm.json", "r02_final_bench_reference_arm.json"), ("launches.csv", "r02_final_launches.csv"),
                  ("small_frames.jsonl", "r02_small_frames.jsonl"), ("config_parity.jsonl", "r02_config_parity.jsonl"),
                  ("refkernel_parity.jsonl", "r02_refkernel_parity.jsonl"), ("rebuild_parity.jsonl", "r02_rebuild_parity.jsonl"), ("pytest_gpu.log", "r02_pytest_gpu.log")):
    if os.path.exists(os.path.join(G, src)):
        shutil.copy(os.path.join(G, src), os.path.join(P, name))
# ---- --set full capture -> text
METRICS = """gpu__time_duration.sum dram__bytes_read.sum dram__bytes_write.sum gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed l1tex__t_sector_hit_rate.pct
lts__t_sector_hit_rate.pct lts__throughput.avg.pct_of_peak_sustained_elapsed l1tex__throughput.avg.pct_of_peak_sustained_elapsed sm__warps_active.avg.pct_of_peak_sustained_active
smsp__issue_active.avg.pct_of_peak_sustained_active smsp__thread_inst_executed_per_inst_executed.ratio smsp__inst_executed.sum launch__registers_per_thread launch__grid_size
launch__block_size smsp__warps_eligible.avg.per_cycle_active sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active
sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active sm__throughput.avg.pct_of_peak_sustained_elapsed
smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio smsp__average_warps_issue_stalled_wait_per_issue_active.ratio
smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio
smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio
smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio smsp__average_warps_issue_stalled_membar_per_issue_active.ratio
launch__shared_mem_per_block_static l1tex__data_pipe_lsu_wavefronts_mem_shared.sum l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum""".split()
raw = os.path.join(G, "r02_full_raw.csv")
if os.path.exists(raw):
    rows = list(csv.reader(open(raw)))
    hi = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
    hdr, units, data = rows[hi], rows[hi + 1], rows[hi + 2:]
    out = ["# ncu --set full --clock-control none --import-source on -k regex:k_wf_extend_ww|k_wf_shade -s 120 -c 2; command: python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-configs",
           "# final round-2 build, REBUILT tree in use (the 120 skipped launches are the trial's calls; call %s, profiles/r02_call_%s/r02_full_raw.csv); the two launches = one extend and one shade launch of a 1080p x 16 spp wave (33.2 M rays)" % (tag, tag)]
    kn = hdr.index("Kernel Name")
    for r in data:
        if not r or len(r) < len(hdr):
            continue
        out.append("== " + r[kn])
        for w in METRICS:
            if w in hdr:
                j = hdr.index(w)
                out.append("  %-90s %-16s %s" % (w, units[j], r[j]))
    open(os.path.join(P, "r02_final_ncu_summary.txt"), "w").write("\n".join(out) + "\n")
# ---- per-config counters -> text
d = json.load(open(os.path.join(P, "r02_frame_counters.json")))
bj = json.load(open(os.path.join(P, "r02_final_bench.json")))
pk, hbm = bj["roofline"]["peaks_measured_this_run"], bj["roofline"]["peak"]
o = ["# Per-config ncu summary of the final round-2 build: every launch of ONE frame profiled (ncu --profile-from-start off --clock-control none,",
     "# metrics gpu__time_duration, smsp__inst_executed, smsp__thread_inst_executed, dram__bytes_read+write, lts__t_bytes, l1tex__t_bytes, issue_active,",
     "# L1/L2 sector hit rates; profiles/tools/ncu_frame.py + ncu_frame_summary.py; raw launch lists: profiles/r02_ncu_frame_<config>.csv), after the",
     "# per-scene trial has ended (the tree in use is the one the library chose).  Times are ncu's (cold-cache, serialised launches): the fractions",
     "# bench.py reports use these PER-RAY counters with LIVE times.",
     "# Peaks measured in the bench run (profiles/r02_final_bench.json): HBM copy %.0f GB/s, L2 stream %.0f, L1 stream %.0f, issue %.4g warp inst/s;" % (hbm, pk["stream_l2_gbs"], pk["stream_l1_gbs"], pk["warp_inst_per_s"]),
     "# random 64-B gathers (2 x LDG.256): L1 %.0f, L2 %.0f, HBM %.0f GB/s.  Fractions below = per-class rate under ncu / those peaks." % (pk["gather_l1_gbs"], pk["gather_l2_gbs"], pk["gather_hbm_gbs"]), ""]
for name, rec in d.items():
    m = rec["meta"]
    o.append("== %s: %d triangles, %dx%d, %d spp, depth %d, %.1f M rays, %d launches, tree in use: %s, ray re-ordering: %s" % (name, m["triangles"], m["width"], m["height"], m["spp"], m["depth"], m["rays"] / 1e6, m["launches"], m["tree_in_use"], m["reorder"]))
    o.append("  %-10s %8s %9s %10s %9s %8s %8s %10s %10s %10s | %6s %6s %6s %6s" % ("class", "launches", "ms", "winst/ray", "lanes", "issue%", "L1hit%", "DRAM B/ray", "L2 B/ray", "L1 B/ray", "issue", "dram", "l2", "l1"))
    for cls, c in rec["classes"].items():
        s = c["ms"] * 1e-3
        fr = (c["warp_inst"] / s / pk["warp_inst_per_s"], c["dram_bytes"] / s / 1e9 / hbm, c["l2_bytes"] / s / 1e9 / pk["stream_l2_gbs"], c["l1_bytes"] / s / 1e9 / pk["stream_l1_gbs"])
        o.append("  %-10s %8d %9.3f %10.2f %9.2f %8.1f %8.1f %10.1f %10.1f %10.1f | %6.3f %6.3f %6.3f %6.3f" % (cls, c["launches"], c["ms"], c["warp_inst_per_ray"], c["lanes_per_instruction"], c["issue_active_pct"], c["l1_hit_pct"], c["dram_bytes_per_ray"], c["l2_bytes_per_ray"], c["l1_bytes_per_ray"], *fr))
        o.append("             kernels: " + ", ".join(c["kernels"]))
    o.append("")
open(os.path.join(P, "r02_configs_ncu_summary.txt"), "w").write("\n".join(o))
print("collected into", dst)
