import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from openbeam_b200 import obgpu, workloads
ctx = obgpu.Context(0)
ctx.upload(workloads.grid_frame(817, 817)); ctx.symbolic()
ctx.solve_static(method=obgpu.PCG_MG, rtol=1e-10)
for w in (obgpu.PROF_MG_SPMV, obgpu.PROF_SPMV):
    print(os.environ.get("OBGPU_MG_FP32"), w, ctx.profile_kernel(w, 50))
