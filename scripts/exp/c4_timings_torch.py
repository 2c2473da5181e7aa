"""Experiment: does an initialised torch (the bench harness imports it) change the host-driven PCG loop's time?"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
mode = sys.argv[1] if len(sys.argv) > 1 else "plain"
if mode != "plain":
    import torch
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    torch.zeros(1, device=dev)
from openbeam_b200 import obgpu, workloads
ctx = obgpu.Context(0)
ctx.upload(workloads.grid_frame(817, 817)); ctx.symbolic()
if mode == "stream":
    ext = torch.cuda.ExternalStream(ctx.stream(), device=dev)
    cm = torch.cuda.stream(ext)
    cm.__enter__()
for i in range(4):
    t0 = time.perf_counter()
    ctx.solve_static(method=obgpu.PCG_MG, rtol=1e-10); ctx.synchronize()
    print(mode, "wall %.2f ms" % ((time.perf_counter() - t0) * 1e3), ctx.timings()["solveStatic.2_solve"])
