"""GPU parity tests of the sparse route (BASELINE configs C4/C5): K1+K4 global assembly (values of K in the
reference's CSC order), K5 BC gather, Jacobi-PCG, reactions and stress, against the CPU oracle and scipy."""
import numpy as np
import pytest
import scipy.sparse as sp
import scipy.sparse.linalg as spla

import fixtures
from openbeam_b200 import obgpu, workloads
from oracle.pyoracle import OracleProblem

pytestmark = pytest.mark.gpu
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


def test_assemble_matches_oracle(ctx):
    """obgpu_assemble: values of K (upper CSC == lower CSR, explicit zeros kept) and F_f, structure by structure."""
    sts = [fixtures.ALL[k]() for k in fixtures.PLAIN] + [workloads.c3_frame(i) for i in range(50)] + \
          [workloads.grid_frame(9, 7), workloads.grid_frame(5, 6, etype=3)]
    ctx.upload(sts)
    ctx.symbolic()
    ctx.assemble()
    kval = ctx.download(obgpu.RES_KVAL)[0]
    Ff = ctx.download(obgpu.RES_FF)[0]
    zo, fo = ctx.offsets(obgpu.OFF_NNZ), ctx.offsets(obgpu.OFF_FREE)
    for s, st in enumerate(sts):
        o = OracleProblem(st).assembleProblem(skip_subK=True)
        _, _, oval = o.K_csc_upper()
        mine = kval[zo[s]:zo[s + 1]]
        assert relerr(mine, oval) <= 1e-13, s
        assert np.array_equal(mine == 0, oval == 0)   # structural zeros stay exact zeros
        assert relerr(Ff[fo[s]:fo[s + 1]], o.F_f) <= 1e-13, s


def test_assemble_per_case_sections(ctx):
    st = workloads.demo3_frame()
    L = 7
    nl, q, eai = workloads.c2_cases(5, L, True)
    ctx.upload(st, L, nl, q, eai)
    ctx.assemble()
    kval = ctx.download(obgpu.RES_KVAL, lk=L)
    o = OracleProblem(st)
    for c in range(L):
        o.set_elem_EAI(eai[c])
        o.assembleProblem(skip_subK=True)
        assert relerr(kval[c], o.K_csc_upper()[2]) <= 1e-13


def oracle_sparse_solve(st):
    """Independent numerics: oracle's K (sparse) -> scipy splu on K_ff."""
    o = OracleProblem(st).assembleProblem(skip_subK=True)
    outer, inner, val = o.K_csc_upper()
    n = len(outer) - 1
    Ku = sp.csc_matrix((val, inner, outer), shape=(n, n))
    K = (Ku + sp.triu(Ku, 1).T).tocsc()
    f, b = o.free_dof_indices, o.bounded_dof_indices
    U = spla.splu(K[f][:, f].tocsc()).solve(o.F_f)
    return o, K, U, f, b


@pytest.mark.parametrize("nx,ny", [(12, 10), (40, 35)])
def test_pcg_grid(ctx, nx, ny):
    st = workloads.grid_frame(nx, ny)
    ctx.upload(st)
    ctx.solve_static(method=obgpu.PCG_JACOBI, rtol=1e-12)
    U, Fb, S = (ctx.download(w)[0] for w in (obgpu.RES_UF, obgpu.RES_FB, obgpu.RES_STRESS))
    assert not ctx.download(obgpu.RES_INFO).any()
    t = ctx.timings()
    assert t["pcg_iters"] > 10
    o, K, Uref, f, b = oracle_sparse_solve(st)
    assert relerr(U, Uref) <= RTOL
    if nx * ny <= 200:  # dense oracle end to end (reactions, stress)
        od = OracleProblem(st).solveStatic().postProcCalcStress()
        assert relerr(U, od.U_f) <= RTOL and relerr(Fb, od.F_b) <= RTOL and relerr(S, od.element_stress) <= RTOL
    # equilibrium: sum of vertical reactions = total UDL; horizontal reactions balance the lateral loads
    fb = Fb.reshape(-1, 3)
    assert abs(fb[:, 1].sum() - 9810.0 * len(st.eload_q)) <= 1e-9 * np.abs(fb[:, 1]).sum()
    assert abs(fb[:, 0].sum() + st.nload_val.sum()) <= 1e-9 * max(np.abs(fb[:, 0]).sum(), 1.0)


def test_pcg_batch_of_grids(ctx):
    """Config C5 shape: several large-ish structures in one batch, each solved by its own PCG."""
    sts = [workloads.grid_frame(14, 11, seed_index=i, section=(2.1e11, workloads.IPE[5 + i][0], workloads.IPE[5 + i][1]))
           for i in range(3)]
    ctx.upload(sts)
    ctx.solve_static(method=obgpu.PCG_JACOBI, rtol=1e-12)
    U, Fb, S = (ctx.download(w)[0] for w in (obgpu.RES_UF, obgpu.RES_FB, obgpu.RES_STRESS))
    fo, bo = ctx.offsets(obgpu.OFF_FREE), ctx.offsets(obgpu.OFF_BOUNDED)
    eo = np.concatenate([[0], np.cumsum([s.n_elems for s in sts])])
    for s, st in enumerate(sts):
        od = OracleProblem(st).solveStatic().postProcCalcStress()
        assert relerr(U[fo[s]:fo[s + 1]], od.U_f) <= RTOL
        assert relerr(Fb[bo[s]:bo[s + 1]], od.F_b) <= RTOL
        assert relerr(S[eo[s]:eo[s + 1]], od.element_stress) <= RTOL


def test_pcg_matches_batched_ldlt(ctx):
    sts = [workloads.c3_frame(i) for i in range(10)]
    ctx.upload(sts)
    ctx.solve_static()
    U1, F1 = ctx.download(obgpu.RES_UF), ctx.download(obgpu.RES_FB)
    ctx.solve_static(method=obgpu.PCG_JACOBI, rtol=1e-12, max_iter=20000)
    U2, F2 = ctx.download(obgpu.RES_UF), ctx.download(obgpu.RES_FB)
    assert relerr(U2, U1) <= RTOL and relerr(F2, F1) <= RTOL


@pytest.mark.parametrize("nx,ny", [(12, 10), (40, 35), (97, 61)])
def test_mg_pcg_grid(ctx, nx, ny):
    """OBGPU_PCG_MG: aggregation-multigrid preconditioned CG (1, 3 and 5 levels here) against scipy splu on the
    oracle's K_ff; the stopping test is on the true residual."""
    st = workloads.grid_frame(nx, ny)
    ctx.upload(st)
    ctx.solve_static(method=obgpu.PCG_MG, rtol=1e-11)
    U, Fb, S = (ctx.download(w)[0] for w in (obgpu.RES_UF, obgpu.RES_FB, obgpu.RES_STRESS))
    assert not ctx.download(obgpu.RES_INFO).any()
    pi = ctx.pcg_info()
    assert pi["true_relres"] <= 1e-11 and pi["iters"] <= 120, pi
    o, K, Uref, f, b = oracle_sparse_solve(st)
    assert relerr(U, Uref) <= RTOL
    # reactions from the independent solution: F_b = K_bf U_f - equivalent loads on the bounded DOFs
    fb = Fb.reshape(-1, 3)
    assert abs(fb[:, 1].sum() - 9810.0 * len(st.eload_q)) <= 1e-9 * np.abs(fb[:, 1]).sum()
    assert abs(fb[:, 0].sum() + st.nload_val.sum()) <= 1e-9 * max(np.abs(fb[:, 0]).sum(), 1.0)
    if nx * ny <= 200:
        od = OracleProblem(st).solveStatic().postProcCalcStress()
        assert relerr(U, od.U_f) <= RTOL and relerr(Fb, od.F_b) <= RTOL and relerr(S, od.element_stress) <= RTOL
    # far fewer iterations than Jacobi on the same system
    if nx * ny > 1000:
        ctx.solve_static(method=obgpu.PCG_JACOBI, rtol=1e-11)
        assert ctx.pcg_info()["iters"] > 4 * pi["iters"]
        assert relerr(ctx.download(obgpu.RES_UF)[0], Uref) <= RTOL


def test_mg_pcg_batch_of_grids(ctx):
    """C5 shape: structures of different sizes and sections in one batch (different hierarchy depths), each solved by
    its own multigrid-PCG on its own stream."""
    dims = [(14, 11), (50, 41), (33, 60), (20, 20)]
    sts = [workloads.grid_frame(nx, ny, seed_index=i, section=(2.1e11, workloads.IPE[5 + i][0], workloads.IPE[5 + i][1]))
           for i, (nx, ny) in enumerate(dims)]
    ctx.upload(sts)
    ctx.solve_static(method=obgpu.PCG_MG, rtol=1e-11)
    U, Fb = (ctx.download(w)[0] for w in (obgpu.RES_UF, obgpu.RES_FB))
    fo, bo = ctx.offsets(obgpu.OFF_FREE), ctx.offsets(obgpu.OFF_BOUNDED)
    for s, st in enumerate(sts):
        o, K, Uref, f, b = oracle_sparse_solve(st)
        assert relerr(U[fo[s]:fo[s + 1]], Uref) <= RTOL
    assert ctx.pcg_info()["true_relres"] <= 1e-11


def test_mg_pcg_matches_batched_ldlt(ctx):
    """mixed RR/RA/AR/AA frames (nodes with 2 and 3 free DOFs): multigrid-PCG against the batched LDLt route"""
    sts = [workloads.c3_frame(i) for i in range(10)]
    ctx.upload(sts)
    ctx.solve_static()
    U1, F1 = ctx.download(obgpu.RES_UF), ctx.download(obgpu.RES_FB)
    ctx.solve_static(method=obgpu.PCG_MG, rtol=1e-12)
    U2, F2 = ctx.download(obgpu.RES_UF), ctx.download(obgpu.RES_FB)
    assert relerr(U2, U1) <= RTOL and relerr(F2, F1) <= RTOL


def test_mg_pcg_case_sweep(ctx):
    """per-case sections: K changes per case, so the Galerkin hierarchy is rebuilt per case"""
    st = workloads.grid_frame(45, 38)
    E = st.n_elems
    eai = np.tile(st.elem_EAI.reshape(1, E, 3), (2, 1, 1)).copy()
    eai[1, :, 1] *= 1.7
    eai[1, :, 2] *= 0.6
    ctx.upload(st, 2, None, None, eai.reshape(2, -1))
    ctx.solve_static(method=obgpu.PCG_MG, rtol=1e-11)
    U = ctx.download(obgpu.RES_UF)
    for c in range(2):
        st_c = workloads.grid_frame(45, 38, section=tuple(eai[c, 0]))
        o, K, Uref, f, b = oracle_sparse_solve(st_c)
        assert relerr(U[c], Uref) <= RTOL


def test_mg_pcg_c4_scale_parity(ctx):
    """One step below config C4's size (400 x 400 grid, 478 800 free DOF): U_f against scipy splu within 1e-9, at the
    tolerance bench.py uses.  The full-size figure (817 x 817) is recorded by scripts/run_lattice.py --check."""
    st = workloads.grid_frame(400, 400)
    ctx.upload(st)
    ctx.solve_static(method=obgpu.PCG_MG, rtol=1e-10)
    U = ctx.download(obgpu.RES_UF)[0]
    o, K, Uref, f, b = oracle_sparse_solve(st)
    assert relerr(U, Uref) <= RTOL
    assert ctx.pcg_info()["true_relres"] <= 1e-10


def test_pcg_not_converged_reported(ctx):
    st = workloads.grid_frame(30, 30)
    ctx.upload(st)
    with pytest.raises(obgpu.ObgpuError) as e:
        ctx.solve_static(method=obgpu.PCG_JACOBI, rtol=1e-13, max_iter=5)
    assert e.value.status == obgpu.ENOTCONVERGED


@pytest.mark.parametrize("name", ["demo3_frame", "demo2", "springs_and_slider", "inclined_mixed", "demo3_frame_tilted_slider"])
def test_sub_matrices_match_oracle(ctx, name):
    """BuildProblemInfo::K_ff / K_bf / K_bb (CFiniteElementProblem.h:86-91): dense-structured views with the reference's
    storage — K_ff lower triangle only, K_bb upper triangle only (quirk Q1) — and the true-sparse K_ff CSR."""
    import fixtures
    st = fixtures.ALL[name]()
    ctx.upload(st)
    ctx.assemble()
    (kff,), (kbf,), (kbb,) = ctx.sub_matrices()
    o = OracleProblem(st).assembleProblem()
    s = o.sizes()
    from oracle import pyoracle
    off = o.getf(pyoracle.F_KFF).reshape(s["n_f"], s["n_f"]).T
    obf = o.getf(pyoracle.F_KBF).reshape(s["n_f"], s["n_b"]).T
    obb = o.getf(pyoracle.F_KBB).reshape(s["n_b"], s["n_b"]).T
    assert relerr(kff, off) <= 1e-13 and relerr(kbf, obf) <= 1e-13 and relerr(kbb, obb) <= 1e-13
    assert not np.triu(kff, 1).any() and not np.tril(kbb, -1).any()        # storage halves exactly as the reference's
    assert np.array_equal(kff != 0, off != 0) and np.array_equal(kbb != 0, obb != 0)
    # true-sparse K_ff: CSR (lower) over free indices
    rp, col = ctx.symbolic_array(obgpu.SYM_KFF_ROWPTR), ctx.symbolic_array(obgpu.SYM_KFF_COL)
    val = ctx.download(obgpu.RES_KFF_CSR_VAL)
    dense = np.zeros_like(kff)
    for r in range(s["n_f"]):
        for p in range(rp[r], rp[r + 1]):
            dense[r, col[p]] = val[p]
    assert np.array_equal(dense, kff)
    # the mirror fills the reference's field names
    from openbeam_b200.problem import BuildProblemInfo, CStructureProblem
    bi = BuildProblemInfo()
    CStructureProblem(st, ctx=ctx).assembleProblem(bi)
    assert np.array_equal(bi.K_ff, kff) and np.array_equal(bi.K_bf, kbf) and np.array_equal(bi.K_bb, kbb)


@pytest.mark.parametrize("method", [obgpu.PCG_MG, obgpu.PCG_JACOBI])
def test_pcg_rotated_frames_and_settlements(ctx, method):
    """Nodal rotated frames (assemble.cpp:144-179, 349-435) and non-zero prescribed displacements (solver_static.cpp:37-43,
    196-198, with the reference's upper-only K_bb, quirk Q1) on the sparse route: every fixture + the settlement cases of
    the batched-LDLt tests in one batch, each structure solved by its own PCG, against the oracle."""
    import fixtures
    from test_gpu_frame import settlement_cases
    sts = [fixtures.ALL[k]() for k in fixtures.ALL] + settlement_cases()
    ctx.upload(sts)
    ctx.solve_static(method=method, rtol=1e-13, max_iter=5000)
    U, Fb, S = (ctx.download(w)[0] for w in (obgpu.RES_UF, obgpu.RES_FB, obgpu.RES_STRESS))
    fo, bo = ctx.offsets(obgpu.OFF_FREE), ctx.offsets(obgpu.OFF_BOUNDED)
    eo = np.concatenate([[0], np.cumsum([s.n_elems for s in sts])])
    for s, st in enumerate(sts):
        od = OracleProblem(st).solveStatic().postProcCalcStress()
        assert relerr(U[fo[s]:fo[s + 1]], od.U_f) <= RTOL, s
        assert relerr(Fb[bo[s]:bo[s + 1]], od.F_b) <= RTOL, s
        assert relerr(S[eo[s]:eo[s + 1]], od.element_stress) <= RTOL, s


def test_mg_fp32_matrix_copies_do_not_change_the_answer(ctx, monkeypatch):
    """The V-cycle streams FP32 copies of its block-SELL matrices (OBGPU_MG_FP32, default on; vectors, products and the
    PCG itself stay FP64).  A preconditioner's precision must not show in the result: both variants stop on the same true
    residual test, take the same number of iterations (to within one check interval) and agree with the oracle."""
    st = workloads.grid_frame(34, 29)
    o = OracleProblem(st).solveStatic()
    got = {}
    for flag in ("0", "1"):
        monkeypatch.setenv("OBGPU_MG_FP32", flag)
        ctx.upload(st)
        ctx.solve_static(method=obgpu.PCG_MG, rtol=1e-12)
        got[flag] = (ctx.download(obgpu.RES_UF)[0].copy(), ctx.download(obgpu.RES_FB)[0].copy(), ctx.pcg_info())
        assert relerr(got[flag][0], o.U_f) <= RTOL and relerr(got[flag][1], o.F_b) <= RTOL
    assert abs(got["0"][2]["iters"] - got["1"][2]["iters"]) <= 2
    assert relerr(got["1"][0], got["0"][0]) <= RTOL


def test_summaries_of_large_structures(ctx):
    """obgpu_summaries_device on the sparse route's few, large structures (one CTA per system, fixed-order reductions)
    against numpy on the downloaded vectors."""
    import torch
    dims = [(50, 41), (33, 60)]
    sts = [workloads.grid_frame(nx, ny, seed_index=i) for i, (nx, ny) in enumerate(dims)]
    ctx.upload(sts)
    ctx.solve_static(method=obgpu.PCG_MG, rtol=1e-11)
    assert ctx.sizes()["nf_tot"] // 2 >= 4096       # the large-system kernel
    out = torch.zeros((2, 4), dtype=torch.float64, device="cuda")
    ctx.summaries_device(out.data_ptr())
    ctx.synchronize()
    got = out.cpu().numpy()
    U, Fb = (ctx.download(w)[0] for w in (obgpu.RES_UF, obgpu.RES_FB))
    fo, bo = ctx.offsets(obgpu.OFF_FREE), ctx.offsets(obgpu.OFF_BOUNDED)
    for s in range(2):
        u, f = U[fo[s]:fo[s + 1]], Fb[bo[s]:bo[s + 1]]
        assert got[s, 0] == np.abs(u).max() and got[s, 3] == 0
        # bounded DOFs of these frames: (DX, DY, RZ) per base node
        assert abs(got[s, 1] - f[0::3].sum()) <= 1e-9 * np.abs(f[0::3]).sum()
        assert abs(got[s, 2] - f[1::3].sum()) <= 1e-9 * np.abs(f[1::3]).sum()
