"""GPU parity tests of the thread-per-system fused path (BASELINE configs C1, C2): libobgpu through the C ABI
vs the CPU oracle on the same inputs.  Tolerance: 1e-9 relative (norm-wise, per result vector) as north_star
states; integer artefacts bit-exact."""
import json
import os

import numpy as np
import pytest

import fixtures
from openbeam_b200 import obgpu, workloads
from openbeam_b200.problem import (CStructureProblem, StaticSolveProblemInfo, StaticSolverAlgorithm,
                                   StaticSolverOptions, StressInfo)
from oracle.pyoracle import OracleProblem

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
RTOL = 1e-9


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


@pytest.mark.parametrize("name", fixtures.PLAIN)
def test_fixture_parity(ctx, name):
    st = fixtures.ALL[name]()
    p = CStructureProblem(st, ctx=ctx)
    info, sinfo = StaticSolveProblemInfo(), StressInfo()
    p.solveStatic(info)
    p.postProcCalcStress(sinfo, info)
    o = OracleProblem(st).solveStatic().postProcCalcStress()
    assert np.array_equal(info.build_info.free_dof_indices, o.free_dof_indices)
    assert np.array_equal(info.build_info.bounded_dof_indices, o.bounded_dof_indices)
    assert np.array_equal(p.getProblemDoFs(), o.problem_DoFs)
    assert relerr(info.build_info.F_f, o.F_f) <= RTOL
    assert relerr(info.U_f, o.U_f) <= RTOL
    assert relerr(info.F_b, o.F_b) <= RTOL
    assert relerr(sinfo.element_stress, o.element_stress) <= RTOL
    g = json.load(open(os.path.join(ROOT, "tests", "golden", name + ".json")))
    unhex = lambda v: np.array([float.fromhex(x) for x in v])
    assert relerr(info.U_f, unhex(g["U_f"])) <= RTOL
    assert relerr(info.F_b, unhex(g["F_b"])) <= RTOL
    assert relerr(sinfo.element_stress, unhex(g["stress"])) <= RTOL


def test_demo1_closed_form(ctx):
    """statically determinate truss."""
    p = CStructureProblem(fixtures.demo1(), ctx=ctx)
    info, s = StaticSolveProblemInfo(), StressInfo()
    p.solveStatic(info)
    p.postProcCalcStress(s)
    assert info.build_info.free_dof_indices.tolist() == [2, 4, 5]
    assert np.allclose(info.U_f, [1.25e-5, 2.5225424859373689e-4, 5.0e-5], rtol=1e-12)
    assert np.allclose(info.F_b, [-1000, -2000, 2000], rtol=1e-12)
    assert np.allclose(s.element_stress[:, 1, 0], [1000, -1000 * np.sqrt(5), 2000], rtol=1e-12)


def _oracle_iterative(st):
    """StaticSolverOptions::nonLinearIterative through the oracle — and through the reference's own code (oracle/_ref)
    when it is there: 40 passes on moved poses, poses restored, then postProcCalcStress on the SAME problem object
    (lengths from the restored poses, orientations and fixed-end terms from the last pass; solver_static.cpp:139-193)."""
    from oracle import pyoracle
    o = OracleProblem(st).solveStatic(nonLinearIterative=True).postProcCalcStress()
    if pyoracle.ref_available():
        r = OracleProblem(st, reference=True).solveStatic(nonLinearIterative=True).postProcCalcStress()
        assert relerr(o.U_f, r.U_f) <= 1e-12 and relerr(o.F_b, r.F_b) <= 1e-12
        assert relerr(o.element_stress, r.element_stress) <= 1e-12
        return r
    return o


# (structures whose DOF SET changes when the nodes move — an axis-aligned spring gaining a second translation — are left
# out: the reference keys constraints and loads by DOF index, which goes stale when its own numbering changes mid-loop)
@pytest.mark.parametrize("name", ["demo3_frame", "demo1", "inclined_mixed", "demo2"])
def test_non_linear_iterative(ctx, name):
    """StaticSolverOptions::nonLinearIterative: 40 assemble+solve passes with the node poses moved by each pass's U_f,
    results of the last pass; afterwards the problem solves on its original poses again."""
    st = fixtures.ALL[name]()
    p = CStructureProblem(st, ctx=ctx)
    info, s = StaticSolveProblemInfo(), StressInfo()
    p.solveStatic(info, StaticSolverOptions(nonLinearIterative=True))
    p.postProcCalcStress(s)
    o = _oracle_iterative(st)
    assert relerr(info.U_f, o.U_f) <= RTOL and relerr(info.F_b, o.F_b) <= RTOL
    assert relerr(s.element_stress, o.element_stress) <= RTOL
    lin = OracleProblem(st).solveStatic()
    assert relerr(info.U_f, lin.U_f) > 1e-9        # the passes did move the nodes
    assert p.structure is st
    p.solveStatic(info)
    assert relerr(info.U_f, lin.U_f) <= RTOL and relerr(info.F_b, lin.F_b) <= RTOL


def test_non_linear_iterative_divergence_raises_like_the_reference(ctx):
    """ex2021_01 under nonLinearIterative: the moved geometry loses positive definiteness in the reference (its own code
    raises 'Ill-conditioned matrix'); the GPU path raises the same message and restores the poses."""
    st = fixtures.ALL["ex2021_01"]()
    with pytest.raises(Exception, match="Ill-conditioned matrix. Too few restrictions"):
        _oracle_iterative(st)
    p = CStructureProblem(st, ctx=ctx)
    with pytest.raises(RuntimeError, match="Ill-conditioned matrix. Too few restrictions"):
        p.solveStatic(StaticSolveProblemInfo(), StaticSolverOptions(nonLinearIterative=True))
    assert p.structure is st


def test_non_linear_iterative_singular_restores_poses(ctx):
    """A mechanism throws out of the first pass with the reference's message (solver_static.cpp:98-105) and leaves the
    problem on its original poses."""
    st = fixtures.mechanism()
    p = CStructureProblem(st, ctx=ctx)
    with pytest.raises(RuntimeError, match="Ill-conditioned matrix. Too few restrictions"):
        p.solveStatic(StaticSolveProblemInfo(), StaticSolverOptions(nonLinearIterative=True))
    assert p.structure is st


def test_singular_raises_reference_message(ctx):
    p = CStructureProblem(fixtures.mechanism(), ctx=ctx)
    with pytest.raises(RuntimeError, match="Ill-conditioned matrix. Too few restrictions"):
        p.solveStatic(StaticSolveProblemInfo())


def test_cpu_algorithms_refused(ctx):
    p = CStructureProblem(fixtures.demo1(), ctx=ctx)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        p.solveStatic(StaticSolveProblemInfo(), StaticSolverOptions(algorithm=StaticSolverAlgorithm.Dense_LLT))


@pytest.mark.parametrize("per_case_section", [False, True])
def test_c2_sample_parity(ctx, per_case_section):
    """Config C2: demo3_frame x load cases; 10^4-case sample against the oracle's per-case solveStatic."""
    st = workloads.demo3_frame()
    L = 10000
    nl, q, eai = workloads.c2_cases(123456, L, per_case_section)
    ctx.upload(st, L, nl, q, eai)
    ctx.solve_static()
    U, Fb, S, Ff = (ctx.download(w) for w in (obgpu.RES_UF, obgpu.RES_FB, obgpu.RES_STRESS, obgpu.RES_FF))
    assert not ctx.download(obgpu.RES_INFO).any()
    _, oU, oF, oS = OracleProblem(st).run_cases(L, nl, q, eai, n_threads=4)
    for c in range(0, L, 97):
        assert relerr(U[c], oU[c]) <= RTOL and relerr(Fb[c], oF[c]) <= RTOL and relerr(S[c], oS[c]) <= RTOL
    assert relerr(U, oU) <= RTOL and relerr(Fb, oF) <= RTOL and relerr(S, oS) <= RTOL
    # host-buffer sweep (the e2e entry point) gives the same bits as upload + solve + download
    U2, F2, S2 = np.empty_like(U), np.empty_like(Fb), np.empty_like(S)
    info = np.empty(L, np.int32)
    ctx.solve_cases_host(L, nl, q, eai, U2, F2, S2, info)
    assert np.array_equal(U, U2) and np.array_equal(Fb, F2) and np.array_equal(S, S2) and not info.any()
    assert Ff.shape == U.shape


@pytest.mark.parametrize("L", [1, 31, 33, 1003, 10007])
def test_ragged_case_counts_and_skip_ff(ctx, L):
    """Case counts that are not a multiple of the warp / CTA width (the bulk-copy record stores fall back to plain
    stores for a ragged tail), and OBGPU_SOLVE_SKIP_FF: same bits for every other result, F_f unavailable."""
    st = workloads.demo3_frame()
    nl, q, _ = workloads.c2_cases(777, L)
    _, oU, oF, oS = OracleProblem(st).run_cases(L, nl, q, None, n_threads=2)
    ctx.upload(st, L, nl, q)
    ctx.solve_static()
    U, Fb, S, Ff = (ctx.download(w) for w in (obgpu.RES_UF, obgpu.RES_FB, obgpu.RES_STRESS, obgpu.RES_FF))
    assert not ctx.download(obgpu.RES_INFO).any()
    assert relerr(U, oU) <= RTOL and relerr(Fb, oF) <= RTOL and relerr(S, oS) <= RTOL
    # F_f by linearity from the oracle's base problem (zero nodal loads, q = 9810): the six nodal-load slots are the
    # free DOFs of nodes 1 and 2 in order
    base = OracleProblem(st).solveStatic()
    assert relerr(Ff, nl + (q / 9810.0) * np.asarray(base.F_f)[None, :]) <= RTOL
    ctx.solve_static(keep_Ff=False)
    assert np.array_equal(U, ctx.download(obgpu.RES_UF)) and np.array_equal(Fb, ctx.download(obgpu.RES_FB))
    assert np.array_equal(S, ctx.download(obgpu.RES_STRESS))
    with pytest.raises(obgpu.ObgpuError, match="F_f was not kept"):
        ctx.download(obgpu.RES_FF)


@pytest.mark.parametrize("name", ["load_kinds", "springs_and_slider"])
def test_case_sweeps_with_load_kinds_and_springs(ctx, name):
    """Load-case sweeps (per-case first load parameter, nodal load values and sections) over structures with
    triangular / concentrated / temperature loads and spring / slider elements, against the oracle case by case."""
    st = fixtures.ALL[name]()
    L = 777
    rng = np.random.default_rng(99)
    NL, EL, E = len(st.nload_val), len(st.eload_q), st.n_elems
    nl = st.nload_val[None, :] * rng.uniform(0.5, 1.5, (L, NL)) if NL else None
    q = st.eload_q[None, :] * rng.uniform(0.5, 1.5, (L, EL)) if EL else None
    eai = (st.elem_EAI.reshape(1, -1) * rng.uniform(0.8, 1.2, (L, 3 * E)))
    ctx.upload(st, L, nl, q, eai)
    ctx.solve_static()
    assert not ctx.download(obgpu.RES_INFO).any()
    U, Fb, S = (ctx.download(w) for w in (obgpu.RES_UF, obgpu.RES_FB, obgpu.RES_STRESS))
    _, oU, oF, oS = OracleProblem(st).run_cases(L, nl, q, eai, n_threads=2)
    assert relerr(U, oU) <= RTOL and relerr(Fb, oF) <= RTOL and relerr(S, oS) <= RTOL
    for c in range(0, L, 111):
        assert relerr(U[c], oU[c]) <= RTOL and relerr(S[c], oS[c]) <= RTOL


def test_bound_summaries_match_separate_kernel(ctx):
    """obgpu_bind_summaries: the summaries the solve kernel writes from registers are bit-identical to the separate
    k_summaries pass, and a later unbound solve leaves the bound buffer alone."""
    import torch
    st = workloads.demo3_frame()
    L = 5003
    nl, q, _ = workloads.c2_cases(4242, L)
    ctx.upload(st, L, nl, q)
    ctx.solve_static()
    ref = torch.empty((L, 4), dtype=torch.float64, device="cuda:0")
    ctx.summaries_device(ref.data_ptr())
    ctx.synchronize()
    fused = torch.full((L, 4), -1.0, dtype=torch.float64, device="cuda:0")
    ctx.bind_summaries(fused.data_ptr())
    try:
        ctx.solve_static(keep_Ff=False)
        assert ctx.timings()["launches"] == (1 if ctx.kernel_info().startswith("k_tiny_jit") else 2)
        ctx.summaries_device(fused.data_ptr())
        ctx.synchronize()
        assert torch.equal(ref, fused)
        U, Fb = ctx.download(obgpu.RES_UF), ctx.download(obgpu.RES_FB)
        h = fused.cpu().numpy()
        assert np.array_equal(h[:, 0], np.abs(U).max(axis=1))
        assert np.array_equal(h[:, 1], Fb[:, 0] + Fb[:, 3]) and np.array_equal(h[:, 2], Fb[:, 1] + Fb[:, 4])
        assert not h[:, 3].any()
    finally:
        ctx.bind_summaries(0)
    fused.fill_(-2.0)
    ctx.solve_static()
    ctx.synchronize()
    assert bool((fused == -2.0).all())


def test_c2_full_size_equilibrium(ctx):
    """All 10^6 cases: global equilibrium  sum(F_b) + sum(F_ext) = 0  per case (size-independent property),
    and linearity: solution of (a + b) = solution(a) + solution(b)."""
    st = workloads.demo3_frame()
    L = 1_000_000
    nl, q, _ = workloads.c2_cases(0, L)
    ctx.upload(st, L, nl, q)
    ctx.solve_static()
    Fb = ctx.download(obgpu.RES_FB)
    assert not ctx.download(obgpu.RES_INFO).any()
    # bounded DOFs: node0 (DX,DY,RZ), node3 (DX,DY,RZ); beam length 4 m, UDL q down
    fx = Fb[:, 0] + Fb[:, 3] + nl[:, 0] + nl[:, 3]
    fy = Fb[:, 1] + Fb[:, 4] + nl[:, 1] + nl[:, 4] - 4.0 * q[:, 0]
    scale = np.abs(Fb).max()
    assert np.abs(fx).max() <= 1e-9 * scale and np.abs(fy).max() <= 1e-9 * scale
    # moment equilibrium about node 0: nodes at (0,0),(0,2),(4,2),(4,0); UDL resultant 4q at x = 2
    mz = (Fb[:, 2] + Fb[:, 5] + nl[:, 2] + nl[:, 5] + 4.0 * Fb[:, 4] + 4.0 * nl[:, 4] - 2.0 * nl[:, 0]
          - 2.0 * nl[:, 3] - 2.0 * 4.0 * q[:, 0])
    assert np.abs(mz).max() <= 1e-9 * scale * 4
    U = ctx.download(obgpu.RES_UF)
    n = 1000
    ctx.update_cases(n, np.ascontiguousarray(nl[:n] + nl[n:2 * n]), np.ascontiguousarray(q[:n] + q[n:2 * n]))
    ctx.solve_static()
    assert relerr(ctx.download(obgpu.RES_UF), U[:n] + U[n:2 * n]) <= 1e-9


def test_timings_and_launch_count(ctx):
    st = workloads.demo3_frame()
    nl, q, _ = workloads.c2_cases(0, 4096)
    ctx.upload(st, 4096, nl, q)
    ctx.solve_static()
    ctx.synchronize()
    t = ctx.timings()
    assert t["launches"] == 1 and t["solveStatic.2_solve"] > 0


def _compact_of(full):
    """[.., 2, 6] FaceStress records -> the compact (N, Vy, Mz) ones; also checks Vz, Mx, My are exactly zero"""
    full = np.asarray(full)
    assert not full[..., 2:5].any()
    return full[..., [0, 1, 5]]


@pytest.mark.parametrize("L", [1, 31, 4096 + 17])
@pytest.mark.parametrize("per_case_section", [False, True])
def test_compact_stress_bit_identical(ctx, L, per_case_section):
    """OBGPU_SOLVE_COMPACT_STRESS: the (N, Vy, Mz) records are the same bits as those columns of the six-component
    record, on the device-resident route and through the host-buffer sweep."""
    st = workloads.demo3_frame()
    nl, q, eai = workloads.c2_cases(5, L, per_case_section)
    ctx.upload(st, L, nl, q, eai)
    ctx.solve_static()
    full = ctx.download(obgpu.RES_STRESS)
    U0, F0 = ctx.download(obgpu.RES_UF), ctx.download(obgpu.RES_FB)
    ctx.solve_static(compact_stress=True, keep_Ff=False)
    comp = ctx.download(obgpu.RES_STRESS_COMPACT)
    assert comp.shape == (L, 3, 2, 3)
    assert np.array_equal(comp, _compact_of(full))
    assert np.array_equal(ctx.download(obgpu.RES_UF), U0) and np.array_equal(ctx.download(obgpu.RES_FB), F0)
    if ctx.kernel_info().startswith("k_tiny_jit"):
        with pytest.raises(obgpu.ObgpuError):
            ctx.download(obgpu.RES_STRESS)      # the six-component record was not produced
    else:                                       # precompiled kernel: six-component record + compaction pass
        assert np.array_equal(ctx.download(obgpu.RES_STRESS), full)
    hU, hF, hS, hi = np.zeros((L, 6)), np.zeros((L, 6)), np.zeros((L, 3, 2, 3)), np.zeros(L, np.int32)
    ctx.solve_cases_host(L, nl, q, eai, hU, hF, hS, hi, compact_stress=True)
    assert np.array_equal(hS, comp) and np.array_equal(hU, U0) and np.array_equal(hF, F0) and not hi.any()


def test_compact_stress_other_routes(ctx):
    """the warp-per-structure and sparse routes compact their six-component record in one more pass"""
    sts = [workloads.c3_frame(i) for i in range(40)]
    ctx.upload(sts)
    ctx.solve_static()
    full = ctx.download(obgpu.RES_STRESS)
    ctx.solve_static(compact_stress=True)
    assert np.array_equal(ctx.download(obgpu.RES_STRESS_COMPACT), _compact_of(full))
    st = workloads.grid_frame(20, 17)
    ctx.upload(st)
    ctx.solve_static(method=obgpu.PCG_MG, rtol=1e-11)
    full = ctx.download(obgpu.RES_STRESS)
    ctx.solve_static(method=obgpu.PCG_MG, rtol=1e-11, compact_stress=True)
    assert np.array_equal(ctx.download(obgpu.RES_STRESS_COMPACT), _compact_of(full))


def test_stress_after_pcg_without_stress_keeps_the_route(ctx):
    """obgpu_stress() after a with_stress=0 PCG solve re-runs the SAME route (it used to switch to LDLt)"""
    st = workloads.grid_frame(30, 25)
    ctx.upload(st)
    ctx.solve_static(method=obgpu.PCG_MG, with_stress=False, rtol=1e-11)
    U = ctx.download(obgpu.RES_UF)
    ctx.stress()
    assert "k_mg_spmv" in ctx.kernel_info()
    assert np.array_equal(ctx.download(obgpu.RES_UF), U)
    od = OracleProblem(st).solveStatic().postProcCalcStress()
    assert relerr(ctx.download(obgpu.RES_STRESS)[0], od.element_stress) <= RTOL


@pytest.mark.skipif(os.environ.get("OBGPU_TINY_JIT") == "0", reason="already inside the fallback run")
def test_precompiled_fallback_kernel_passes_the_same_tests():
    """Without libnvrtc / libcuda (or with OBGPU_TINY_JIT=0) the thread-per-case route runs the precompiled k_tiny_fused,
    and structures with non-uniform loads go to the warp-per-structure kernel.  The switch is read once per process, so the
    tiny-route tests and the reference vectors are re-run in a child process with it set."""
    import subprocess
    import sys
    env = dict(os.environ, OBGPU_TINY_JIT="0")
    r = subprocess.run([sys.executable, "-m", "pytest", "-q", "-x", "-m", "gpu", "-p", "no:cacheprovider",
                        os.path.join(os.path.dirname(__file__), "test_gpu_tiny.py"),
                        os.path.join(os.path.dirname(__file__), "test_gpu_reference_vectors.py")],
                       env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-1000:]
