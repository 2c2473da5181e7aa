"""Config C3 throughput (BASELINE configs[2] shape, one GPU's share): S random small frames through the
warp-per-structure fused route (k_frame_fused).  Distinct structures are generated from the counter RNG and tiled
to S (host-side generation is Python; the kernel sees S independent systems either way).  Prints one JSON line:
structures/s, DOF/s, algorithmic GB/s (in 16N + 33E + 12 n_b + 12 #loads, out 8n + 8 n_b + 96E)."""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from openbeam_b200 import obgpu, workloads

ap = argparse.ArgumentParser()
ap.add_argument("--structures", type=int, default=1 << 20)
ap.add_argument("--distinct", type=int, default=4096)
ap.add_argument("--steps", type=int, default=5)
ap.add_argument("--check", type=int, default=64, help="structures compared with the oracle")
a = ap.parse_args()

t0 = time.time()
base = [workloads.c3_frame(i) for i in range(a.distinct)]
reps = (a.structures + a.distinct - 1) // a.distinct
sts = (base * reps)[:a.structures]
t_gen = time.time() - t0
ctx = obgpu.Context(0)
t0 = time.time(); ctx.upload(sts); t_up = time.time() - t0
t0 = time.time(); ctx.symbolic(); t_sym = time.time() - t0
z = ctx.sizes()
ms = []
for _ in range(a.steps + 2):
    ctx.solve_static()
    ctx.synchronize()
    t = ctx.timings()
    ms.append(t["assembleProblem"] + t["solveStatic.2_solve"] + t["postProcCalcStress"])
ms = ms[2:]
best, med = min(ms), float(np.median(ms))
loads = z["NLtot"] + z["ELtot"]
alg_bytes = 16 * z["Ntot"] + 33 * z["Etot"] + 12 * z["nb_tot"] + 12 * loads + 8 * z["n_tot"] + 8 * z["nb_tot"] + 96 * z["Etot"]
out = dict(structures=a.structures, distinct=a.distinct, sizes=z, gen_s=t_gen, upload_s=t_up, symbolic_s=t_sym,
           ms_per_step_median=med, ms_per_step_best=best, structures_per_s=a.structures / (med * 1e-3),
           dof_per_s=z["n_tot"] / (med * 1e-3), algorithmic_bytes=alg_bytes, algorithmic_GBps=alg_bytes / (med * 1e-3) / 1e9,
           launches=t["launches"])
if a.check:
    from oracle.pyoracle import OracleProblem
    U = ctx.download(obgpu.RES_UF)[0]
    Fb = ctx.download(obgpu.RES_FB)[0]
    fo, bo = ctx.offsets(obgpu.OFF_FREE), ctx.offsets(obgpu.OFF_BOUNDED)
    worst = 0.0
    for s in range(a.check):
        o = OracleProblem(sts[s]).solveStatic()
        worst = max(worst, float(np.abs(U[fo[s]:fo[s + 1]] - o.U_f).max() / np.abs(o.U_f).max()),
                    float(np.abs(Fb[bo[s]:bo[s + 1]] - o.F_b).max() / np.abs(o.F_b).max()))
    out["max_relerr_vs_oracle"] = worst
print(json.dumps(out))
