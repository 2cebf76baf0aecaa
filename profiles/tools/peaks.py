#!/usr/bin/env python
"""The measured peaks bench.py's roofline fractions are stated against: the micro-benchmarks live in the
library (g.p.u-pathtracer_b200/csrc/gpt_peaks.cu, C entry gpt_measure_peaks) so that bench.py can run them in the SAME process and
on the same clocks as the timed region; this script just prints them.
   python profiles/tools/peaks.py  ->  {"gather_l1_gbs": ..., "gather_l2_gbs": ..., "gather_hbm_gbs": ..., "warp_inst_per_s": ...,
                                        "stream_l1_gbs": ..., "stream_l2_gbs": ...}
gather_*: random 64-byte fetches in the BVH node fetch's own load shape (2 x LDG.E.256) from an L1- / L2- / HBM-resident working set;
stream_*: coalesced 16-byte loads sweeping an L1- / L2-resident working set (the ceiling of the level); warp_inst_per_s: eight
independent FFMA chains per thread with every scheduler busy."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import gpt_b200 as gpt  # noqa: E402

with gpt.Context(int(sys.argv[1]) if len(sys.argv) > 1 else 0) as ctx:
    print(json.dumps(ctx.measure_peaks()))
