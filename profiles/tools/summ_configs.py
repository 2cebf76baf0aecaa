#!/usr/bin/env python
"""stdin: configs_bench.py JSON lines -> one short line per measurement."""
import json
import sys

for line in sys.stdin:
    try:
        d = json.loads(line)
    except Exception:
        continue
    print("  %-50s %-26s %7.2f ms %6.0f Mr/s ext %.2f shade %.2f" % (d["config"][:50], d["tree"][:26], d["ms_per_frame"], d["Mrays_per_s"],
                                                                 d["kernel_class_ms"].get("extend", 0), d["kernel_class_ms"].get("shade", 0)))
