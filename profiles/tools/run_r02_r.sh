#!/bin/bash
# round 2, GPU call R (1 GPU): compat and FAST shade kernels side by side under ncu (default build and the pipelined variant)
mkdir -p gpurun_out
M=gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active,smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio,smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio,smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread,launch__occupancy_limit_registers,dram__bytes_read.sum,dram__bytes_write.sum,l1tex__t_sector_hit_rate.pct,lts__t_sector_hit_rate.pct
cp g.p.u-pathtracer_b200/libgpt_b200.so /tmp/default.so
for f in /tmp/default.so variants/pipe_pf1.so; do
  cp $f g.p.u-pathtracer_b200/libgpt_b200.so; n=$(basename $f .so)
  GPT_REBUILD=1 timeout 600 ncu --clock-control none --csv --log-file gpurun_out/ncu_shade_$n.csv --metrics $M -k regex:k_wf_shade -s 41 -c 90 python profiles/tools/shade_ab.py > gpurun_out/ncu_shade_$n.log 2>&1
  python - $n <<'PY'
import csv, collections, sys
rows=list(csv.reader(open('gpurun_out/ncu_shade_%s.csv' % sys.argv[1])))
hi=[i for i,r in enumerate(rows) if r and r[0]=='ID'][0]
h=rows[hi]; kn=h.index('Kernel Name'); mn=h.index('Metric Name'); mv=h.index('Metric Value'); idc=h.index('ID')
d=collections.OrderedDict()
for r in rows[hi+1:]:
    if len(r)<=mv: continue
    d.setdefault((int(r[idc]), r[kn][:22]), {})[r[mn]]=r[mv]
for k,v in d.items():
    if k[0] in (2, 3, 4, 42, 43, 44): print(sys.argv[1], k, {a.split('.')[0][-30:]:b for a,b in v.items()})
PY
done
cp /tmp/default.so g.p.u-pathtracer_b200/libgpt_b200.so
