"""Large-lattice run (BASELINE configs C4/C5 shape): timings of symbolic / assembly / PCG at a given grid size."""
import argparse
import json
import sys
import time
import os

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from openbeam_b200 import obgpu, workloads

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=817)
ap.add_argument("--structures", type=int, default=1)
ap.add_argument("--rtol", type=float, default=1e-10)
ap.add_argument("--max-iter", type=int, default=0)
ap.add_argument("--method", default="mg", choices=["mg", "jacobi"])
ap.add_argument("--repeat", type=int, default=1, help="solve this many times, report the last (first includes plan + warm-up)")
ap.add_argument("--check", action="store_true", help="compare U_f with scipy splu on the oracle's K (CPU, slow)")
a = ap.parse_args()
t0 = time.time()
sts = [workloads.grid_frame(a.n, a.n, seed_index=i) for i in range(a.structures)]
t_gen = time.time() - t0
ctx = obgpu.Context(0)
t0 = time.time(); ctx.upload(sts); t_up = time.time() - t0
t0 = time.time(); ctx.symbolic(); t_sym = time.time() - t0
z = ctx.sizes()
t0 = time.time(); ctx.assemble(); ctx.synchronize(); t_asm_first = time.time() - t0
ctx.assemble(); ctx.synchronize()
t_asm = ctx.timings()["assembleProblem"]
method = obgpu.PCG_MG if a.method == "mg" else obgpu.PCG_JACOBI
solve_first = None
for _ in range(a.repeat):
    t0 = time.time()
    try:
        ctx.solve_static(method=method, rtol=a.rtol, max_iter=a.max_iter)
        status = "converged"
    except obgpu.ObgpuError as e:
        status = str(e)
    ctx.synchronize()
    t_solve = time.time() - t0
    if solve_first is None:
        solve_first = t_solve
t = ctx.timings()
out = dict(n=a.n, structures=a.structures, sizes=z, gen_s=t_gen, upload_s=t_up, symbolic_s=t_sym,
           assemble_first_s=t_asm_first, assemble_ms=t_asm, solve_s=t_solve, solve_ms_events=t["solveStatic.2_solve"],
           solve_first_s=solve_first, method=a.method, rtol=a.rtol, pcg=ctx.pcg_info(), kernel=ctx.kernel_info(),
           pcg_iters=t["pcg_iters"], launches=t["launches"], status=status)
if a.check:
    import scipy.sparse as sp
    import scipy.sparse.linalg as spla
    from oracle.pyoracle import OracleProblem
    U = ctx.download(obgpu.RES_UF)[0]
    fo = ctx.offsets(obgpu.OFF_FREE)
    o = OracleProblem(sts[0]).assembleProblem(skip_subK=True)
    outer, inner, val = o.K_csc_upper()
    n = len(outer) - 1
    Ku = sp.csc_matrix((val, inner, outer), shape=(n, n))
    K = (Ku + sp.triu(Ku, 1).T).tocsc()
    f = o.free_dof_indices
    Uref = spla.splu(K[f][:, f].tocsc()).solve(o.F_f)
    out["relerr_vs_splu"] = float(np.abs(U[fo[0]:fo[1]] - Uref).max() / np.abs(Uref).max())
print(json.dumps(out))
