"""CPU baselines beside the GPU numbers of configs C3 and C4, timed on the box's host cores.
Test/bench infrastructure: this script (like bench.py's cpu_baseline leg) is allowed to call oracle/.

C3: the oracle port in faithful mode (dense-structured K_ff extraction + dense lower Cholesky, one
    solveStatic + postProcCalcStress per structure), 1 thread and one Python thread per core (the ctypes calls
    release the GIL; the reference itself is single-threaded).
C4: (a) the largest faithful size that completes in a bounded time (the reference's O(n^2) sub-matrix extraction makes
    full size infeasible), (b) a clearly NON-reference sparse baseline at a given grid size: oracle sparse assembly +
    scipy splu."""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from openbeam_b200 import workloads
from oracle.pyoracle import OracleProblem

ap = argparse.ArgumentParser()
ap.add_argument("--c3", type=int, default=4096, help="C3 structures in the sample")
ap.add_argument("--faithful-n", type=int, nargs="*", default=[12, 20, 28], help="grid sizes of the faithful C4 runs")
ap.add_argument("--splu-n", type=int, default=0, help="grid size of the non-reference splu run (0 = skip)")
a = ap.parse_args()
cores = os.cpu_count() or 1
out = {"cores": cores}

def solve(p):
    p.solveStatic().postProcCalcStress()


def c3_worker(rng):
    """One process per core: builds its slice of the sample (untimed), then times its solves."""
    probs = [OracleProblem(workloads.c3_frame(i)) for i in range(*rng)]
    t0 = time.perf_counter()
    for p in probs:
        solve(p)
    return len(probs), time.perf_counter() - t0, float(np.sum([p.sizes()["n"] for p in probs]))


import multiprocessing as mp

n1, t1, d1 = c3_worker((0, min(1024, a.c3)))
per = (a.c3 + cores - 1) // cores
with mp.get_context("fork").Pool(cores) as pool:
    res = pool.map(c3_worker, [(k * per, min(a.c3, (k + 1) * per)) for k in range(cores) if k * per < a.c3])
out["c3"] = {"sample": a.c3, "structures_per_s_1core": n1 / t1,
             "structures_per_s_all_cores": sum(r[0] for r in res) / max(r[1] for r in res),
             "mean_dofs": sum(r[2] for r in res) / sum(r[0] for r in res),
             "kind": "port (oracle, faithful mode), one process per core, solve calls only"}

out["c4_faithful"] = []
for n in a.faithful_n:
    st = workloads.grid_frame(n, n)
    p = OracleProblem(st)
    t0 = time.perf_counter()
    solve(p)
    dt = time.perf_counter() - t0
    out["c4_faithful"].append({"grid": n, "n_dof": p.sizes()["n"], "seconds": dt, "dof_per_s": p.sizes()["n"] / dt})
    if dt > 60:
        break

if a.splu_n:
    import scipy.sparse as sp
    import scipy.sparse.linalg as spla
    st = workloads.grid_frame(a.splu_n, a.splu_n)
    t0 = time.perf_counter()
    o = OracleProblem(st).assembleProblem(skip_subK=True)
    t_asm = time.perf_counter() - t0
    outer, inner, val = o.K_csc_upper()
    n = len(outer) - 1
    t0 = time.perf_counter()
    Ku = sp.csc_matrix((val, inner, outer), shape=(n, n))
    K = (Ku + sp.triu(Ku, 1).T).tocsc()
    f = o.free_dof_indices
    Kff = K[f][:, f].tocsc()
    t_ext = time.perf_counter() - t0
    t0 = time.perf_counter()
    U = spla.splu(Kff).solve(o.F_f)
    t_lu = time.perf_counter() - t0
    out["c4_sparse_nonreference"] = {"grid": a.splu_n, "n_dof": n, "oracle_sparse_assembly_s": t_asm,
                                     "extract_s": t_ext, "splu_factor_solve_s": t_lu,
                                     "dof_per_s": n / (t_asm + t_ext + t_lu), "umax": float(np.abs(U).max())}
print(json.dumps(out))
