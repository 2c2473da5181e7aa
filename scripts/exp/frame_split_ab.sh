#!/bin/bash
# experiment: fused kernel against the three-kernel path on 262 144 C3 frames, and bitwise comparison of their results
for v in 0 1; do OBGPU_FRAME_SPLIT=$v python scripts/exp/frame_clocks.py 262144 | head -1; done
python - <<'PY'
import os, numpy as np
from openbeam_b200 import obgpu, workloads
out = {}
for v in ("0", "1"):
    os.environ["OBGPU_FRAME_SPLIT"] = v
    ctx = obgpu.Context(0); ctx.upload(workloads.c3_batch(0, 20000), device=True); ctx.symbolic(); ctx.solve_static()
    out[v] = [ctx.download(w).copy() for w in (obgpu.RES_UF, obgpu.RES_FB, obgpu.RES_STRESS, obgpu.RES_FF, obgpu.RES_INFO)]
    print(v, ctx.kernel_info()); ctx.close()
for n, x, y in zip(("U_f", "F_b", "stress", "F_f", "info"), out["0"], out["1"]):
    d = np.abs(x - y).max() / max(np.abs(x).max(), 1e-300)
    print(n, "identical" if np.array_equal(x, y) else "max rel diff %.3g" % d)
PY
