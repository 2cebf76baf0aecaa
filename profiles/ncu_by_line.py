#!/usr/bin/env python
"""Joins an ncu SASS source page (ncu -i X.ncu-rep --page source --csv) with nvdisasm -g line info of the
same cubin and aggregates per source line: warp instructions, thread instructions, SIMT efficiency and
stall samples.  Usage: ncu_by_line.py src.csv dis.txt kernel_mangled_substring [top_n]"""
import csv
import re
import sys
from collections import defaultdict

src_csv, dis_txt, kname = sys.argv[1], sys.argv[2], sys.argv[3]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
addr2line = {}
cur = None
infunc = False
for line in open(dis_txt):
    if line.startswith(".text."):
        infunc = kname in line
        continue
    if not infunc:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', line)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
    if m:
        addr2line[int(m.group(1), 16)] = (cur, m.group(2).strip())
rows = list(csv.reader(open(src_csv)))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
agg = defaultdict(lambda: [0, 0, 0, 0])
tot = [0, 0, 0]
base = None
for r in rows[2:]:
    if len(r) < len(hdr) or r[ix["Address"]] == "Address":     # section header of a further launch in the report
        base = None
        continue
    a = int(r[ix["Address"]], 16) if r[ix["Address"]].startswith("0x") else int(r[ix["Address"]])
    if base is None:
        base = a
    off = a - base
    ln = addr2line.get(off, (("?", 0), ""))[0]
    wi = int(r[ix["Instructions Executed"]] or 0)
    ti = int(r[ix["Thread Instructions Executed"]] or 0)
    ss = int(r[ix["Warp Stall Sampling (All Samples)"]] or 0)
    g = agg[ln]
    g[0] += wi; g[1] += ti; g[2] += ss; g[3] += 1
    tot[0] += wi; tot[1] += ti; tot[2] += ss
print("total warp inst %d  thread inst %d  avg threads/inst %.2f  samples %d" % (tot[0], tot[1], tot[1] / max(1, tot[0]), tot[2]))
print("%-28s %7s %7s %6s %7s %5s" % ("file:line", "warp%", "thr%", "thr/wi", "stall%", "sass"))
for ln, g in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print("%-28s %6.2f%% %6.2f%% %6.1f %6.2f%% %5d" % ("%s:%d" % ln, 100 * g[0] / tot[0], 100 * g[1] / tot[1], g[1] / max(1, g[0]), 100 * g[2] / max(1, tot[2]), g[3]))
