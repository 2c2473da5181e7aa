"""Regenerates tests/golden/reference/*.json from THE REFERENCE ITSELF: oracle/_ref/libob_ref.so is libopenbeam's own,
unmodified hot-path sources (/root/reference/libopenbeam/src/*.cpp) compiled against the header stand-ins of
oracle/ref_shim/ (oracle/Makefile.ref) and driven through the reference's public API (oracle/ref_harness.cpp).
/root/reference does not exist on the GPU box, so the vectors are committed (hex floats, bit-exact):
    python tests/golden/make_golden_reference.py
What is frozen per case: DOF list, free/bounded partition, dof_index inverse list, the dense-structured K_ff (lower),
K_bf, K_bb (upper) of BuildProblemInfo, U_b, F_f, U_f, F_b and the element end forces.
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path[:0] = [os.path.dirname(os.path.dirname(HERE)), os.path.dirname(HERE)]
import fixtures  # noqa: E402
from openbeam_b200 import workloads  # noqa: E402
from oracle import pyoracle  # noqa: E402
from oracle.pyoracle import OracleProblem  # noqa: E402

OUT = os.path.join(HERE, "reference")
hx = lambda a: [float(x).hex() for x in np.asarray(a).ravel()]


def record(p, with_matrices=True):
    g = dict(problem_DoFs=p.problem_DoFs.tolist(), free_dof_indices=p.free_dof_indices.tolist(),
             bounded_dof_indices=p.bounded_dof_indices.tolist(), dof_index=p.geti(pyoracle.I_DOF_INDEX).tolist(),
             U_b=hx(p.U_b), F_f=hx(p.F_f), U_f=hx(p.U_f), F_b=hx(p.F_b), stress=hx(p.element_stress))
    if with_matrices:
        g.update(K_ff=hx(p.getf(pyoracle.F_KFF)), K_bf=hx(p.getf(pyoracle.F_KBF)), K_bb=hx(p.getf(pyoracle.F_KBB)))
    return g


def cases():
    for name, make in fixtures.ALL.items():
        yield name, make()
    for i in (0, 1, 2, 3, 5, 8, 13, 21):       # config C3 structures (random frames, mixed RR/RA/AR/AA)
        yield "c3_frame_%d" % i, workloads.c3_frame(i)
    yield "grid_frame_7x6", workloads.grid_frame(7, 6)   # the C4/C5 family at a size the dense reference handles


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    for name, st in cases():
        p = OracleProblem(st, reference=True).solveStatic().postProcCalcStress()
        json.dump(record(p, with_matrices=st.n_nodes <= 16), open(os.path.join(OUT, name + ".json"), "w"), indent=0)
        print(name, p.sizes())
    # config C2: demo3_frame under the first 48 randomized load cases (and 16 with per-case sections)
    st = workloads.demo3_frame()
    for tag, per_case in (("c2_cases", False), ("c2b_cases", True)):
        L = 16 if per_case else 48
        nl, q, eai = workloads.c2_cases(0, L, per_case)
        _, U, F, S = OracleProblem(st, reference=True).run_cases(L, nl, q, eai)
        json.dump(dict(first=0, count=L, U_f=hx(U), F_b=hx(F), stress=hx(S)), open(os.path.join(OUT, tag + ".json"), "w"), indent=0)
        print(tag, L)
