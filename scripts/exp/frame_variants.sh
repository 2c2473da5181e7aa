#!/bin/bash
# experiment helper: time every libobgpu variant under variants/ on 262 144 C3 frames (the default library is restored by a rebuild)
for so in variants/*.so; do
  cp $so openbeam_b200/libobgpu.so
  echo "== $so"
  python scripts/exp/frame_clocks.py 262144 | head -1
done
