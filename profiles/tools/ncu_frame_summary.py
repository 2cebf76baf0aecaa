#!/usr/bin/env python
"""gpurun_out/ncu_frame_<scene>.csv (+ .json) -> per-kernel-class counters of ONE frame, per ray:
   python profiles/tools/ncu_frame_summary.py gpurun_out/ncu_frame_bench.csv [...]  > profiles/r02_frame_counters.json
Classes: extend (k_wf_extend_ww*), shade (k_wf_shade*), other (collect / accumulate / sort).  For each: launches, ms under ncu,
warp instructions, thread instructions (-> lanes per instruction), DRAM / L2 / L1 bytes, time-weighted issue-active %, hit rates."""
import csv
import json
import os
import sys


def klass(name):
    if "k_wf_extend" in name:
        return "extend"
    if "k_wf_shade" in name:
        return "shade"
    return "other"


def fnum(s):
    return float(str(s).replace(",", ""))


UNIT = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "ms": 1.0, "us": 1e-3, "ns": 1e-6, "s": 1e3}   # usecond/msecond handled below


def scale(unit):
    u = unit.strip()
    if u in ("usecond", "us"):
        return 1e-3
    if u in ("msecond", "ms"):
        return 1.0
    if u in ("nsecond", "ns"):
        return 1e-6
    if u in ("second", "s"):
        return 1e3
    return UNIT.get(u, 1.0)


def summarise(path):
    rows = []
    with open(path) as fh:
        lines = [ln for ln in fh if ln.startswith('"')]
    rd = csv.DictReader(lines)
    per = {}
    for r in rd:
        kid = r["ID"]
        d = per.setdefault(kid, {"name": r["Kernel Name"]})
        d[r["Metric Name"]] = fnum(r["Metric Value"]) * scale(r["Metric Unit"])
    meta = {}
    mp = path[:-4] + ".json"
    if os.path.exists(mp):
        meta = json.load(open(mp))
    out = {"meta": meta, "classes": {}}
    for kid, d in per.items():
        c = out["classes"].setdefault(klass(d["name"]), {"launches": 0, "ms": 0.0, "warp_inst": 0.0, "thread_inst": 0.0, "dram_bytes": 0.0,
                                                         "l2_bytes": 0.0, "l1_bytes": 0.0, "_issue_w": 0.0, "_l1hit_w": 0.0, "_l2hit_w": 0.0, "kernels": set()})
        ms = d.get("gpu__time_duration.sum", 0.0)
        c["launches"] += 1
        c["ms"] += ms
        c["warp_inst"] += d.get("smsp__inst_executed.sum", 0.0)
        c["thread_inst"] += d.get("smsp__thread_inst_executed.sum", 0.0)
        c["dram_bytes"] += d.get("dram__bytes_read.sum", 0.0) + d.get("dram__bytes_write.sum", 0.0)
        c["l2_bytes"] += d.get("lts__t_bytes.sum", 0.0)
        c["l1_bytes"] += d.get("l1tex__t_bytes.sum", 0.0)
        c["_issue_w"] += ms * d.get("smsp__issue_active.avg.pct_of_peak_sustained_active", 0.0)
        c["_l1hit_w"] += ms * d.get("l1tex__t_sector_hit_rate.pct", 0.0)
        c["_l2hit_w"] += ms * d.get("lts__t_sector_hit_rate.pct", 0.0)
        c["kernels"].add(d["name"].split("(")[0])
    rays = float(meta.get("rays", 0)) or None
    for c in out["classes"].values():
        ms = c["ms"] or 1.0
        c["issue_active_pct"] = c.pop("_issue_w") / ms
        c["l1_hit_pct"] = c.pop("_l1hit_w") / ms
        c["l2_hit_pct"] = c.pop("_l2hit_w") / ms
        c["lanes_per_instruction"] = c["thread_inst"] / c["warp_inst"] if c["warp_inst"] else None
        c["kernels"] = sorted(c["kernels"])
        if rays:
            c["warp_inst_per_ray"] = c["warp_inst"] / rays
            c["dram_bytes_per_ray"] = c["dram_bytes"] / rays
            c["l2_bytes_per_ray"] = c["l2_bytes"] / rays
            c["l1_bytes_per_ray"] = c["l1_bytes"] / rays
    return out


def main():
    res = {}
    for p in sys.argv[1:]:
        s = summarise(p)
        res[s["meta"].get("scene", os.path.basename(p))] = s
    json.dump(res, sys.stdout, indent=1)
    print()


if __name__ == "__main__":
    main()
