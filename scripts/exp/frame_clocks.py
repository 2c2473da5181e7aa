"""Experiment: per-stage cycle shares of k_frame_fused.  Needs libobgpu.so built with -DOBGPU_FRAME_CLOCKS
(make -C openbeam_b200/csrc NVFLAGS+=-DOBGPU_FRAME_CLOCKS after touching kernels_frame.cu)."""
import ctypes, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
from openbeam_b200 import obgpu, workloads

S = int(sys.argv[1]) if len(sys.argv) > 1 else 262144
ctx = obgpu.Context(0)
ctx.upload(workloads.c3_batch(0, S))
ctx.symbolic()
lib = obgpu.lib()
buf = (ctypes.c_ulonglong * 8)()
ms = []
have = hasattr(lib, "obgpu_debug_frame_clocks")
for it in range(6):
    ctx.solve_static(); ctx.synchronize()
    t = ctx.timings()
    ms.append(t["assembleProblem"] + t["solveStatic.2_solve"] + t["postProcCalcStress"])
    if have:
        lib.obgpu_debug_frame_clocks(buf, 1)
print("structures/s %.4g  (ms %s)" % (S / (min(ms[2:]) * 1e-3), ["%.3f" % m for m in ms]))
if not have:
    sys.exit(0)
v = np.array(list(buf), float)
names = ["zero+nodal loads", "A: element compute", "A: scatter", "B: factor+fwd", "B: back subst + U store", "C: stress+reactions"]
print(json.dumps({n: round(x / v.sum(), 4) for n, x in zip(names, v)}), "cycles/structure", v.sum() / S)
