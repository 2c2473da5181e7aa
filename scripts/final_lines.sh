#!/bin/bash
# one-GPU evidence pass: bench lines of every config + launch list and ncu --set full of the C3 kernel (profiles/ naming: pass TAG)
TAG=${1:-r02n}
mkdir -p gpurun_out
python bench.py > gpurun_out/${TAG}_bench_c2.json 2> gpurun_out/${TAG}_bench_c2.err
python bench.py --impl reference --steps 3 > gpurun_out/${TAG}_bench_c2_reference_arm.json 2>> gpurun_out/${TAG}_bench_c2.err
python bench.py --config c3 > gpurun_out/${TAG}_bench_c3.json 2> gpurun_out/${TAG}_bench_c3.err
python bench.py --config c4 > gpurun_out/${TAG}_bench_c4.json 2> gpurun_out/${TAG}_bench_c4.err
python bench.py --config c5 > gpurun_out/${TAG}_bench_c5.json 2> gpurun_out/${TAG}_bench_c5.err
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,launch__grid_size,launch__shared_mem_per_block_dynamic,launch__registers_per_thread \
  --clock-control none -k regex:k_frame_fused --launch-skip 18 --launch-count 9 --csv --log-file gpurun_out/${TAG}_launches_c3_frame_classes.csv \
  python scripts/exp/frame_clocks.py 262144 > /dev/null 2>&1
ncu --set full --import-source on --clock-control none -k regex:k_frame_fused --launch-skip 18 --launch-count 9 -o gpurun_out/${TAG}_frame \
  python scripts/exp/frame_clocks.py 262144 > /dev/null 2>&1
ls -la gpurun_out | tail -12
