"""GPU results against vectors produced by THE REFERENCE'S OWN CODE (tests/golden/reference/*.json: libopenbeam's
unmodified sources compiled into oracle/_ref, see tests/golden/make_golden_reference.py).  1e-9 relative per result vector
(north_star); DOF numbering and partition bit-exact.  Each case runs on the route a user would get for it."""
import json
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import make_golden_reference as mgr
from openbeam_b200 import obgpu, workloads
from openbeam_b200.problem import CStructureProblem, StaticSolveProblemInfo, StressInfo

pytestmark = pytest.mark.gpu
REFDIR = os.path.join(HERE, "golden", "reference")
unhex = lambda v: np.array([float.fromhex(x) for x in v])
RTOL = 1e-9
CASES = dict(mgr.cases())


def relerr(a, b):
    a, b = np.asarray(a, float).ravel(), np.asarray(b, float).ravel()
    assert a.shape == b.shape
    if a.size == 0:
        return 0.0
    den = np.max(np.abs(b))
    return float(np.max(np.abs(a - b)) / den) if den > 0 else float(np.max(np.abs(a)))


@pytest.fixture(scope="module")
def ctx():
    c = obgpu.Context(0)
    yield c
    c.close()


@pytest.mark.parametrize("name", sorted(CASES))
def test_structure_matches_reference(ctx, name):
    g = json.load(open(os.path.join(REFDIR, name + ".json")))
    p = CStructureProblem(CASES[name], ctx=ctx)
    info, sinfo = StaticSolveProblemInfo(), StressInfo()
    p.solveStatic(info)
    p.postProcCalcStress(sinfo, info)
    assert p.getProblemDoFs().tolist() == g["problem_DoFs"]
    assert info.build_info.free_dof_indices.tolist() == g["free_dof_indices"]
    assert info.build_info.bounded_dof_indices.tolist() == g["bounded_dof_indices"]
    assert relerr(info.build_info.F_f, unhex(g["F_f"])) <= RTOL
    assert relerr(info.U_f, unhex(g["U_f"])) <= RTOL
    assert relerr(info.F_b, unhex(g["F_b"])) <= RTOL
    assert relerr(sinfo.element_stress, unhex(g["stress"])) <= RTOL


def test_grid_frame_multigrid_matches_reference(ctx):
    g = json.load(open(os.path.join(REFDIR, "grid_frame_7x6.json")))
    ctx.upload(CASES["grid_frame_7x6"])
    for method in (obgpu.PCG_MG, obgpu.PCG_JACOBI):
        ctx.solve_static(method=method, rtol=1e-12)
        assert relerr(ctx.download(obgpu.RES_UF)[0], unhex(g["U_f"])) <= RTOL
        assert relerr(ctx.download(obgpu.RES_FB)[0], unhex(g["F_b"])) <= RTOL
        assert relerr(ctx.download(obgpu.RES_STRESS)[0], unhex(g["stress"])) <= RTOL


@pytest.mark.parametrize("tag,per_case", [("c2_cases", False), ("c2b_cases", True)])
def test_c2_cases_match_reference(ctx, tag, per_case):
    g = json.load(open(os.path.join(REFDIR, tag + ".json")))
    L = g["count"]
    nl, q, eai = workloads.c2_cases(g["first"], L, per_case)
    ctx.upload(workloads.demo3_frame(), L, nl, q, eai)
    ctx.solve_static()
    assert relerr(ctx.download(obgpu.RES_UF), unhex(g["U_f"])) <= RTOL
    assert relerr(ctx.download(obgpu.RES_FB), unhex(g["F_b"])) <= RTOL
    assert relerr(ctx.download(obgpu.RES_STRESS), unhex(g["stress"])) <= RTOL
