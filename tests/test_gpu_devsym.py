"""Device batch mode (obgpu_batch_upload_device + the symbolic phase as device kernels, kernels_symbolic.cu) against the
host symbolic phase and the oracle: same offsets, same bits in every result."""
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
    den = np.max(np.abs(b)) if b.size else 0.0
    return float(np.max(np.abs(a - b)) / den) if den > 0 else (float(np.max(np.abs(a))) if a.size else 0.0)


@pytest.fixture(scope="module")
def ctx():
    c = obgpu.Context(0)
    yield c
    c.close()


def run(ctx, batch, device, **kw):
    ctx.upload(batch, device=device, **kw)
    ctx.symbolic()
    ctx.solve_static()
    res = {k: ctx.download(w) for k, w in (("U", obgpu.RES_UF), ("Fb", obgpu.RES_FB), ("S", obgpu.RES_STRESS),
                                           ("Ff", obgpu.RES_FF), ("Ub", obgpu.RES_UB))}
    res["info"] = ctx.download(obgpu.RES_INFO)
    res["off"] = [ctx.offsets(w) for w in (obgpu.OFF_DOF, obgpu.OFF_FREE, obgpu.OFF_BOUNDED)]
    res["sizes"] = ctx.sizes()
    return res


def test_random_frames_device_equals_host(ctx, monkeypatch):
    batch = workloads.c3_batch(7000, 3000)
    h_rcm = run(ctx, batch, False)     # host mode renumbers the free DOFs of wide-band structures (reverse Cuthill-McKee)
    monkeypatch.setenv("OBGPU_FRAME_NO_RCM", "1")
    h = run(ctx, batch, False)         # same elimination order as the device mode: same bits
    d = run(ctx, batch, True)
    for k in ("U", "Fb", "S"):
        assert relerr(h_rcm[k], d[k]) <= 1e-10, k
    for a, b in zip(h["off"], d["off"]):
        assert np.array_equal(a, b)
    for k in ("U", "Fb", "S", "Ff", "Ub", "info"):
        assert np.array_equal(h[k], d[k]), k
    assert h["sizes"]["nf_tot"] == d["sizes"]["nf_tot"] and h["sizes"]["n_tot"] == d["sizes"]["n_tot"]
    # and against the oracle on a sample
    fo, bo = d["off"][1], d["off"][2]
    for s in range(0, 3000, 97):
        o = OracleProblem(workloads.c3_frame(7000 + s)).solveStatic()
        assert relerr(d["U"][0, fo[s]:fo[s + 1]], o.U_f) <= RTOL and relerr(d["Fb"][0, bo[s]:bo[s + 1]], o.F_b) <= RTOL


def test_fixtures_device_equals_host(ctx, monkeypatch):
    """every element type, load kind, rotated nodal frames and non-zero prescribed displacements"""
    monkeypatch.setenv("OBGPU_FRAME_NO_RCM", "1")
    sts = [make() for make in fixtures.ALL.values()]
    h = run(ctx, sts, False)
    d = run(ctx, sts, True)
    for a, b in zip(h["off"], d["off"]):
        assert np.array_equal(a, b)
    assert np.array_equal(h["info"], d["info"])
    for k in ("U", "Fb", "S", "Ff", "Ub"):   # cos/sin of a rotated frame may differ in the last bit between host and device
        assert relerr(d[k], h[k]) <= 1e-13, k


def test_device_batch_with_load_cases_and_summaries(ctx, monkeypatch):
    import torch
    monkeypatch.setenv("OBGPU_FRAME_NO_RCM", "1")
    batch = workloads.c3_batch(100, 64)
    L = 3
    rng = np.random.default_rng(3)
    nl = rng.uniform(0, 2e4, (L, len(batch["nload_val"])))
    q = rng.uniform(5e3, 3e4, (L, len(batch["eload_q"])))
    h = run(ctx, batch, False, n_cases=L, case_nload_val=nl, case_eload_q=q)
    d = run(ctx, batch, True, n_cases=L, case_nload_val=nl, case_eload_q=q)
    for k in ("U", "Fb", "S"):
        assert np.array_equal(h[k], d[k]), k
    out = torch.zeros((L * 64, 4), dtype=torch.float64, device="cuda:0")
    ctx.summaries_device(out.data_ptr())
    ctx.synchronize()
    fo = d["off"][1]
    got = out.cpu().numpy().reshape(L, 64, 4)
    for c in range(L):
        for s in (0, 17, 63):
            assert got[c, s, 0] == np.abs(d["U"][c, fo[s]:fo[s + 1]]).max()


def test_device_batch_materialises_for_host_queries_and_pcg(ctx):
    batch = workloads.c3_batch(0, 20)
    ctx.upload(batch, device=True)
    ctx.symbolic()
    ctx.solve_static()
    U1 = ctx.download(obgpu.RES_UF)
    free_d = ctx.symbolic_array(obgpu.SYM_FREE)     # brings the batch to the host, runs the host phase
    ctx.upload(batch)
    ctx.symbolic()
    assert np.array_equal(free_d, ctx.symbolic_array(obgpu.SYM_FREE))
    ctx.upload(batch, device=True)
    ctx.solve_static(method=obgpu.PCG_MG, rtol=1e-12)   # the sparse route plans on the host
    assert relerr(ctx.download(obgpu.RES_UF), U1) <= RTOL


def test_device_batch_refuses_malformed_structure(ctx):
    batch = workloads.c3_batch(0, 50)
    bad = dict(batch)
    en = batch["elem_nodes"].copy()
    e0 = batch["elem_off"][31]
    en[2 * e0 + 1] = 1000          # node index out of range in structure 31
    bad["elem_nodes"] = en
    ctx.upload(bad, device=True)
    with pytest.raises(obgpu.ObgpuError) as e:
        ctx.symbolic()
    assert e.value.status == obgpu.EINVAL and "structure 31" in str(e.value) and "out of range" in str(e.value)
    bad = dict(batch)
    t = batch["elem_type"].copy()
    t[batch["elem_off"][7]] = 77
    bad["elem_type"] = t
    ctx.upload(bad, device=True)
    with pytest.raises(obgpu.ObgpuError) as e:
        ctx.symbolic()
    assert "unknown element type" in str(e.value)


def test_batch_pipeline_matches_one_context():
    """openbeam_b200/pipeline.py: sub-batches dealt to several contexts on concurrent host threads give, concatenated, the
    results of one context over the whole list (bit for bit: independent structures, same kernel)."""
    from openbeam_b200.pipeline import BatchPipeline
    S, nsub = 6000, 5
    ctx = obgpu.Context(0)
    ctx.upload(workloads.c3_batch(100, S), device=True)
    ctx.symbolic()
    ctx.solve_static(compact_stress=True)
    U, Fb, St = ctx.download(obgpu.RES_UF), ctx.download(obgpu.RES_FB), ctx.download(obgpu.RES_STRESS_COMPACT)
    ctx.close()
    per = S // nsub
    subs = [workloads.c3_batch(100 + i * per, per) for i in range(nsub)]
    pipe = BatchPipeline(0, workers=3)
    res = pipe.run(subs, offsets=True)
    res = pipe.run(subs, res, offsets=True)   # second pass reuses the result buffers
    pipe.close()
    assert np.array_equal(np.concatenate([r["U_f"][0] for r in res]), U[0])
    assert np.array_equal(np.concatenate([r["F_b"][0] for r in res]), Fb[0])
    assert np.array_equal(np.concatenate([r["stress"][0] for r in res]), St[0])
    assert all(not r["info"].any() for r in res) and all(len(r["free_off"]) == per + 1 for r in res)
