#!/usr/bin/env python
"""ncu report -> the per-kernel summary text and the traffic / issue JSON bench.py reads.
usage: ncu_summary.py report.ncu-rep out_summary.txt out_traffic.json "<header note>"   (runs `ncu -i ... --page raw --csv` here)"""
import csv
import io
import json
import subprocess
import sys

rep, out_txt, out_json, note = sys.argv[1], sys.argv[2], sys.argv[3], sys.argv[4] if len(sys.argv) > 4 else ""
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
METRICS = ["dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
           "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
           "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
           "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
           "smsp__inst_executed.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
           "smsp__warps_eligible.avg.per_cycle_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
           "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
           "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
           "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio", "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio"]
ix = {h: i for i, h in enumerate(hdr)}
SCALE = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "ms": 1.0, "us": 1e-3, "s": 1e3, "ns": 1e-6}
txt, js = [], {}
for r in rows[2:]:
    name = r[ix["Kernel Name"]]
    short = name.split("(")[0].split("<")[0].replace("void ", "").strip()
    txt.append("== %s  (%s)" % (short, note))
    txt.append("  " + name)
    val = {}
    for m in METRICS:
        if m not in ix:
            continue
        u, v = units[ix[m]], r[ix[m]].replace(",", "")
        txt.append("  %-90s %-16s %s" % (m, u, v))
        try:
            val[m] = float(v) * (SCALE.get(u, 1.0) if m.startswith("dram__bytes") or m == "gpu__time_duration.sum" else 1.0)
        except ValueError:
            pass
    txt.append("")
    js[short] = {"dram_bytes_per_launch": val.get("dram__bytes_read.sum", 0) + val.get("dram__bytes_write.sum", 0),
                 "dram_read": val.get("dram__bytes_read.sum"), "dram_write": val.get("dram__bytes_write.sum"),
                 "issue_active_pct": val.get("smsp__issue_active.avg.pct_of_peak_sustained_active"),
                 "lanes_per_instruction": val.get("smsp__thread_inst_executed_per_inst_executed.ratio"),
                 "warp_instructions": val.get("smsp__inst_executed.sum"), "ms_under_ncu": val.get("gpu__time_duration.sum"),
                 "registers": val.get("launch__registers_per_thread"), "grid": val.get("launch__grid_size"),
                 "source": "%s (%s)" % (out_txt, note)}
open(out_txt, "w").write("\n".join(txt))
json.dump(js, open(out_json, "w"), indent=1)
print("\n".join(txt))
