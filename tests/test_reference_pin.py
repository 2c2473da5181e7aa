"""Pins the oracle (oracle/, the CPU restatement) to the REFERENCE ITSELF.

Two ties, both mechanical:
  1. committed vectors tests/golden/reference/*.json, produced by libopenbeam's own unmodified sources compiled into
     oracle/_ref/libob_ref.so (tests/golden/make_golden_reference.py) — checked everywhere;
  2. the live library, where it is built (this container builds it from /root/reference; elsewhere the prebuilt file
     that travelled with the snapshot is used, else the live part is skipped).
Bar: DOF list / partition / inverse list bit-exact; the dense-structured K_ff, K_bf, K_bb, U_b and F_f bit-exact (no
arithmetic differs between the two); U_f, F_b and element end forces within 1e-12 relative (the factorisation is the
only place where operation order may differ)."""
import json
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import fixtures
import make_golden_reference as mgr
from openbeam_b200 import workloads
from oracle import pyoracle
from oracle.pyoracle import OracleProblem

REFDIR = os.path.join(HERE, "golden", "reference")
unhex = lambda v: np.array([float.fromhex(x) for x in v])
TOL = 1e-12


def relerr(a, b):
    a, b = np.asarray(a, float).ravel(), np.asarray(b, float).ravel()
    assert a.shape == b.shape
    if a.size == 0:
        return 0.0
    den = np.max(np.abs(b))
    return float(np.max(np.abs(a - b)) / den) if den > 0 else float(np.max(np.abs(a)))


CASES = dict(mgr.cases())


@pytest.mark.parametrize("name", sorted(CASES))
def test_oracle_matches_reference_vectors(name):
    g = json.load(open(os.path.join(REFDIR, name + ".json")))
    p = OracleProblem(CASES[name]).solveStatic().postProcCalcStress()
    assert p.problem_DoFs.tolist() == g["problem_DoFs"]
    assert p.free_dof_indices.tolist() == g["free_dof_indices"]
    assert p.bounded_dof_indices.tolist() == g["bounded_dof_indices"]
    assert p.geti(pyoracle.I_DOF_INDEX).tolist() == g["dof_index"]
    assert np.array_equal(p.U_b, unhex(g["U_b"])) and np.array_equal(p.F_f, unhex(g["F_f"]))
    if "K_ff" in g:
        assert np.array_equal(p.getf(pyoracle.F_KFF), unhex(g["K_ff"]))
        assert np.array_equal(p.getf(pyoracle.F_KBF), unhex(g["K_bf"]))
        assert np.array_equal(p.getf(pyoracle.F_KBB), unhex(g["K_bb"]))
    assert relerr(p.U_f, unhex(g["U_f"])) <= TOL
    assert relerr(p.F_b, unhex(g["F_b"])) <= TOL
    assert relerr(p.element_stress, unhex(g["stress"])) <= TOL


@pytest.mark.parametrize("tag,per_case", [("c2_cases", False), ("c2b_cases", True)])
def test_oracle_c2_cases_match_reference_vectors(tag, per_case):
    g = json.load(open(os.path.join(REFDIR, tag + ".json")))
    L = g["count"]
    nl, q, eai = workloads.c2_cases(g["first"], L, per_case)
    _, U, F, S = OracleProblem(workloads.demo3_frame()).run_cases(L, nl, q, eai)
    assert relerr(U, unhex(g["U_f"])) <= TOL and relerr(F, unhex(g["F_b"])) <= TOL and relerr(S, unhex(g["stress"])) <= TOL


needs_ref = pytest.mark.skipif(not pyoracle.ref_available(), reason="oracle/_ref/libob_ref.so not built and /root/reference absent")


@needs_ref
def test_reference_vectors_are_current():
    """the committed vectors are what the reference build produces now"""
    for name in ("demo3_frame", "springs_and_slider", "c3_frame_5"):
        g = json.load(open(os.path.join(REFDIR, name + ".json")))
        p = OracleProblem(CASES[name], reference=True).solveStatic().postProcCalcStress()
        assert np.array_equal(p.U_f, unhex(g["U_f"])) and np.array_equal(p.element_stress.ravel(), unhex(g["stress"]))


@needs_ref
def test_oracle_matches_live_reference_on_random_frames():
    """200 random C3 frames (RR/RA/AR/AA mix): the restatement against the reference's own code"""
    worst = 0.0
    for i in range(100, 300):
        st = workloads.c3_frame(i)
        a = OracleProblem(st).solveStatic().postProcCalcStress()
        b = OracleProblem(st, reference=True).solveStatic().postProcCalcStress()
        assert np.array_equal(a.problem_DoFs, b.problem_DoFs) and np.array_equal(a.free_dof_indices, b.free_dof_indices)
        assert np.array_equal(a.bounded_dof_indices, b.bounded_dof_indices)
        assert np.array_equal(a.F_f, b.F_f)
        if i < 120:
            assert np.array_equal(a.getf(pyoracle.F_KFF), b.getf(pyoracle.F_KFF))
            assert np.array_equal(a.getf(pyoracle.F_KBF), b.getf(pyoracle.F_KBF))
            assert np.array_equal(a.getf(pyoracle.F_ELEM_KGLOBAL), b.getf(pyoracle.F_ELEM_KGLOBAL))
        worst = max(worst, relerr(a.U_f, b.U_f), relerr(a.F_b, b.F_b), relerr(a.element_stress, b.element_stress))
    assert worst <= 1e-10   # frames up to cond ~1e8: the two Cholesky orders differ by cond * eps


@needs_ref
def test_reference_raises_on_mechanism():
    """solver_static.cpp:98-105: the reference's own message, through the compiled reference"""
    st = fixtures.ALL["demo1"]()
    st.con_node, st.con_dof, st.con_val = st.con_node[:1], st.con_dof[:1], st.con_val[:1]
    with pytest.raises(pyoracle.OracleError) as e:
        OracleProblem(st, reference=True).solveStatic()
    assert "Ill-conditioned matrix. Too few restrictions?" in str(e.value)
    with pytest.raises(pyoracle.OracleError) as e2:
        OracleProblem(st).solveStatic()
    assert "Ill-conditioned matrix. Too few restrictions?" in str(e2.value)
