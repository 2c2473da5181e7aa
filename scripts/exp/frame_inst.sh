#!/bin/bash
# experiment helper: warp instructions / issue utilisation / time of every k_frame_fused class launch of one C3 solve
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum \
  --clock-control none -k regex:k_frame_fused --launch-skip 18 --launch-count 9 --csv python scripts/exp/frame_clocks.py 262144 2>/dev/null | python -c "
import csv,sys,collections
rows=[r for r in csv.reader(sys.stdin) if len(r)>14 and r[0].isdigit()]
d=collections.defaultdict(dict)
for r in rows: d[r[0]][r[12]]=float(r[14].replace(',',''))
T=sum(v['gpu__time_duration.sum'] for v in d.values()); I=sum(v['smsp__inst_executed.sum'] for v in d.values()); W=sum(v['l1tex__data_pipe_lsu_wavefronts_mem_shared.sum'] for v in d.values())
print('launches',len(d),'time_us %.0f'%(T/1e3),'inst/structure %.0f'%(I/262144),'smem wavefronts/structure %.0f'%(W/262144),'issue%% (time-weighted) %.1f'%(sum(v['gpu__time_duration.sum']*v['smsp__issue_active.avg.pct_of_peak_sustained_active'] for v in d.values())/T))
"
