"""Regenerates tests/golden/*.json from the CPU oracle (oracle/).  The reference itself cannot be
built or imported in this image (MRPT + Eigen missing), so these vectors freeze the ORACLE's
output (hex floats, bit-exact) after it was pinned by closed forms and independent numerics
(tests/test_oracle.py).  Run:  python tests/golden/make_golden.py
"""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path[:0] = [os.path.dirname(os.path.dirname(HERE)), os.path.dirname(HERE)]
import fixtures  # noqa: E402
from oracle.pyoracle import OracleProblem  # noqa: E402

for name, make in fixtures.ALL.items():
    p = OracleProblem(make())
    p.solveStatic().postProcCalcStress()
    outer, inner, val = p.K_csc_upper()
    hx = lambda a: [float(x).hex() for x in a]
    g = dict(sizes=p.sizes(), problem_DoFs=p.problem_DoFs.tolist(),
             free_dof_indices=p.free_dof_indices.tolist(), bounded_dof_indices=p.bounded_dof_indices.tolist(),
             K_outer=outer.tolist(), K_inner=inner.tolist(), K_val=hx(val), U_b=hx(p.U_b), F_f=hx(p.F_f),
             U_f=hx(p.U_f), F_b=hx(p.F_b), stress=hx(p.element_stress.ravel()),
             U_f_dec=p.U_f.tolist(), F_b_dec=p.F_b.tolist())
    json.dump(g, open(os.path.join(HERE, name + ".json"), "w"), indent=0)
    print(name, p.sizes())
