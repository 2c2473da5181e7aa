"""GPU parity tests of the warp-per-structure fused path (BASELINE config C3: batches of different small frames):
libobgpu through the C ABI vs the CPU oracle, structure by structure.  1e-9 relative per result vector."""
import numpy as np
import pytest

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


def check_batch(ctx, sts, rtol=RTOL):
    ctx.upload(sts)
    ctx.symbolic()
    ctx.solve_static()
    U, Fb, S, Ff = (ctx.download(w)[0] for w in (obgpu.RES_UF, obgpu.RES_FB, obgpu.RES_STRESS, obgpu.RES_FF))
    info = ctx.download(obgpu.RES_INFO)[0]
    fo, bo = ctx.offsets(obgpu.OFF_FREE), ctx.offsets(obgpu.OFF_BOUNDED)
    eo = np.concatenate([[0], np.cumsum([s.n_elems for s in sts])])
    assert not info.any()
    worst = 0.0
    for s, st in enumerate(sts):
        o = OracleProblem(st).solveStatic().postProcCalcStress()
        errs = (relerr(Ff[fo[s]:fo[s + 1]], o.F_f), relerr(U[fo[s]:fo[s + 1]], o.U_f),
                relerr(Fb[bo[s]:bo[s + 1]], o.F_b), relerr(S[eo[s]:eo[s + 1]], o.element_stress))
        assert max(errs) <= rtol, (s, errs)
        worst = max(worst, max(errs))
    return worst


def test_fixture_batch(ctx):
    """All plain fixtures as ONE batch: forces the general warp-per-structure kernel (S > 1)."""
    check_batch(ctx, [fixtures.ALL[k]() for k in fixtures.PLAIN])
    assert ctx.timings()["launches"] == 1


def settlement_cases():
    """Non-zero prescribed displacements (support settlement / imposed rotation), with and without nodal rotated
    frames, incl. a support whose DOFs are only partly constrained (K_bb upper-triangle quirk Q1 is exercised by
    nodes with several prescribed DOFs)."""
    from openbeam_b200.structure import build_structure
    sec = fixtures.IPE200
    L, H = 4.0, 2.0
    el = [("BEAM2D_RR", 0, 1, "IPE200"), ("BEAM2D_RR", 1, 2, "IPE200"), ("BEAM2D_RR", 2, 3, "IPE200")]
    out = []
    for rot in (None, 30.0, -20.0):
        n3 = (L, 0) if rot is None else (L, 0, rot)
        nodes = [(0, 0), (0, H), (L, H), n3]
        cons = [(0, "DX", 1e-3), (0, "DY", -2e-3), (0, "RZ", 5e-4), (3, "DX", 0.0), (3, "DY", -5e-3), (3, "RZ", -1e-3)]
        out.append(build_structure(nodes, el, sec, cons, [(1, "DX", 5e3)], [(1, 9810.0, fixtures.DOWN)]))
        cons = [(0, "DXDYRZ"), (3, "DY", -4e-3)]  # slider with a settlement
        out.append(build_structure(nodes, el, sec, cons, [(2, "DY", -1e4)], [(1, 9810.0, fixtures.DOWN)]))
    # propped cantilever, prescribed end rotation and a mid support settlement
    nodes = [(0, 0), (2.0, 0), (4.5, 0, 15.0)]
    el = [("BEAM2D_RR", 0, 1, "IPE200"), ("BEAM2D_RA", 1, 2, "IPE200")]
    cons = [(0, "DXDYRZ"), (1, "DY", -3e-3), (2, "DY", 1e-3)]
    out.append(build_structure(nodes, el, sec, cons, [], [(0, 5e3, fixtures.DOWN), (1, 5e3, fixtures.DOWN)]))
    return out


def test_rotated_frames_and_settlements(ctx):
    """nodal rotated frames (assemble.cpp:144-179,349-435) and U_b != 0 (solver_static.cpp:37-43,196-198)
    on the GPU, against the oracle; the tilted reference fixtures ride in the same batch as plain ones."""
    sts = [fixtures.ALL[k]() for k in fixtures.ALL] + settlement_cases()
    check_batch(ctx, sts)
    # U_b as BuildProblemInfo holds it (nodal coordinates)
    Ub = ctx.download(obgpu.RES_UB)
    bo = ctx.offsets(obgpu.OFF_BOUNDED)
    for s, st in enumerate(sts):
        o = OracleProblem(st).solveStatic()
        assert relerr(Ub[bo[s]:bo[s + 1]], o.U_b) <= 1e-15
    # a single structure with U_b != 0 must not take the thread-per-case kernel's U_b = 0 shortcut
    check_batch(ctx, [settlement_cases()[0]])


def test_random_frames(ctx):
    """Config C3 sample: 300 random frames (10-60 elements, mixed RR/RA/AR/AA beams)."""
    check_batch(ctx, [workloads.c3_frame(i) for i in range(300)])


def test_single_larger_structure(ctx):
    """S = 1 but n_f > 12: the thread-per-system kernel does not apply, the banded kernel must."""
    st = workloads.grid_frame(6, 9)
    check_batch(ctx, [st])
    st = workloads.grid_frame(5, 7, etype=3)  # AA grid: pin-jointed, only DX/DY DOFs... a mechanism without bracing
    ctx.upload([st])
    ctx.solve_static()
    assert ctx.download(obgpu.RES_INFO).any()  # unbraced pin-jointed grid is singular -> info > 0


def test_singular_member_flagged(ctx):
    sts = [fixtures.demo1(), fixtures.mechanism(), fixtures.demo3_frame()]
    ctx.upload(sts)
    ctx.solve_static()
    info = ctx.download(obgpu.RES_INFO)[0]
    assert info[0] == 0 and info[1] > 0 and info[2] == 0
    # the healthy neighbours are still right
    U = ctx.download(obgpu.RES_UF)[0]
    fo = ctx.offsets(obgpu.OFF_FREE)
    assert relerr(U[fo[2]:fo[3]], OracleProblem(sts[2]).solveStatic().U_f) <= RTOL


def test_batch_with_load_cases(ctx):
    """S structures x L load cases: system (c, s)."""
    sts = [workloads.c3_frame(i) for i in range(40, 48)]
    pk = obgpu.pack_batch(sts)
    L = 5
    rng = np.random.default_rng(7)
    nl = pk["nload_val"][None, :] * rng.uniform(0.5, 1.5, (L, pk["nload_val"].size))
    q = pk["eload_q"][None, :] * rng.uniform(0.5, 1.5, (L, pk["eload_q"].size))
    ctx.upload(pk, L, nl, q)
    ctx.solve_static()
    U, Fb = ctx.download(obgpu.RES_UF), ctx.download(obgpu.RES_FB)
    fo, bo = ctx.offsets(obgpu.OFF_FREE), ctx.offsets(obgpu.OFF_BOUNDED)
    no = np.concatenate([[0], np.cumsum([len(s.nload_val) for s in sts])])
    qo = np.concatenate([[0], np.cumsum([len(s.eload_q) for s in sts])])
    for s, st in enumerate(sts):
        o = OracleProblem(st)
        for c in (0, L - 1):
            o.set_load_values(nl[c, no[s]:no[s + 1]], q[c, qo[s]:qo[s + 1]])
            o.solveStatic()
            assert relerr(U[c, fo[s]:fo[s + 1]], o.U_f) <= RTOL
            assert relerr(Fb[c, bo[s]:bo[s + 1]], o.F_b) <= RTOL


def test_summaries(ctx):
    import torch
    sts = [workloads.c3_frame(i) for i in range(64)]
    ctx.upload(sts)
    ctx.solve_static()
    out = torch.empty((64, 4), dtype=torch.float64, device="cuda:0")
    ctx.summaries_device(out.data_ptr())
    ctx.synchronize()
    out = out.cpu().numpy()
    U, Fb = ctx.download(obgpu.RES_UF)[0], ctx.download(obgpu.RES_FB)[0]
    fo, bo = ctx.offsets(obgpu.OFF_FREE), ctx.offsets(obgpu.OFF_BOUNDED)
    for s, st in enumerate(sts):
        assert out[s, 0] == np.abs(U[fo[s]:fo[s + 1]]).max()
        fb = Fb[bo[s]:bo[s + 1]].reshape(-1, 3)  # bases are fixed: (DX, DY, RZ) per base node
        assert np.isclose(out[s, 1], fb[:, 0].sum(), rtol=1e-12) and np.isclose(out[s, 2], fb[:, 1].sum(), rtol=1e-12)
        # global equilibrium: sum of horizontal reactions = -(sum of wind loads)
        assert abs(out[s, 1] + st.nload_val.sum()) <= 1e-9 * max(1.0, np.abs(fb[:, 0]).max())
        assert out[s, 3] == 0


def test_three_kernel_path_is_bitwise_the_fused_kernel(monkeypatch):
    """kernels_frame.cu: k_frame_pre -> k_frame_fused<split> -> k_frame_post (the default for >= 4096 structures, one load
    case, identity node frames, U_b = 0) forms every sum in the fused kernel's order: identical bits, both symbolic modes."""
    from openbeam_b200 import workloads
    batch = workloads.c3_batch(300, 3000)
    for device_mode in (True, False):
        got = {}
        for split in ("0", "1"):
            monkeypatch.setenv("OBGPU_FRAME_SPLIT", split)
            monkeypatch.setenv("OBGPU_FRAME_SPLIT_MIN", "1")
            ctx = obgpu.Context(0)
            ctx.upload(batch, device=device_mode)
            ctx.symbolic()
            ctx.solve_static()
            assert ("k_frame_pre" in ctx.kernel_info()) == (split == "1")
            got[split] = [ctx.download(w).copy() for w in (obgpu.RES_UF, obgpu.RES_FB, obgpu.RES_STRESS, obgpu.RES_FF, obgpu.RES_INFO)]
            ctx.solve_static(compact_stress=True)
            got[split].append(ctx.download(obgpu.RES_STRESS_COMPACT).copy())
            ctx.close()
        for x, y in zip(got["0"], got["1"]):
            assert np.array_equal(x, y)
