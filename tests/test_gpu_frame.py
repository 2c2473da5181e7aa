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
