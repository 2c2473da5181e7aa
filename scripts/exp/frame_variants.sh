#!/bin/bash
# experiment helper: run every libobgpu variant under variants/ through the C3 timing + instruction count
for so in variants/*.so; do
  cp $so openbeam_b200/libobgpu.so
  echo "== $so"
  python scripts/exp/frame_clocks.py 262144 | head -1
  bash scripts/exp/frame_inst.sh
done
