#!/bin/bash
# experiment: per-kernel times / occupancy of the three-kernel frame path (one solve of 262 144 frames)
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread,launch__shared_mem_per_block_dynamic,launch__grid_size,dram__bytes_read.sum,dram__bytes_write.sum \
  --clock-control none -k regex:k_frame --launch-skip ${SKIP:-22} --launch-count ${COUNT:-11} --csv python scripts/exp/frame_clocks.py 262144 2>/dev/null | python -c "
import csv,sys,collections
rows=[r for r in csv.reader(sys.stdin) if len(r)>14 and r[0].isdigit()]
d=collections.OrderedDict()
for r in rows: d.setdefault(r[0],{'k':r[4][:40]})[r[12]]=float(r[14].replace(',',''))
T=0
for i,v in d.items():
    T+=v['gpu__time_duration.sum']
    print('%-42s %8.1f us  grid %5d smem %6d regs %3d  warps %4.1f%% issue %4.1f%%  inst %6.1fM  dram r/w %6.1f / %6.1f MB'%(v['k'],v['gpu__time_duration.sum']/1e3,v['launch__grid_size'],v['launch__shared_mem_per_block_dynamic'],v['launch__registers_per_thread'],v['sm__warps_active.avg.pct_of_peak_sustained_active'],v['smsp__issue_active.avg.pct_of_peak_sustained_active'],v['smsp__inst_executed.sum']/1e6,v['dram__bytes_read.sum']/1e6,v['dram__bytes_write.sum']/1e6))
print('total %.0f us'%(T/1e3))
"
