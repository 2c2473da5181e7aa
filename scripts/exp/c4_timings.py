"""Experiment: phase timings (ctx.timings) of repeated C4 solves."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from openbeam_b200 import obgpu, workloads
n = int(sys.argv[1]) if len(sys.argv) > 1 else 817
ctx = obgpu.Context(0)
ctx.upload(workloads.grid_frame(n, n)); ctx.symbolic()
import time
for i in range(4):
    t0 = time.perf_counter()
    ctx.solve_static(method=obgpu.PCG_MG, rtol=1e-10)
    t1 = time.perf_counter()
    ctx.synchronize()
    t2 = time.perf_counter()
    t = ctx.timings()
    print("wall: call %.2f ms, sync %.2f ms;" % ((t1 - t0) * 1e3, (t2 - t1) * 1e3),
          json.dumps({k: round(v, 3) for k, v in t.items() if k in ("assembleProblem", "solveStatic.2_solve")}), ctx.pcg_info()["iters"])
