"""Host logic of openbeam_b200/pipeline.py on the CPU: a stub context records what the workers call.  (The GPU parity of
the pipeline is tests/test_gpu_devsym.py::test_batch_pipeline_matches_one_context.)"""
import threading
import time

import numpy as np
import pytest

from openbeam_b200 import obgpu, pipeline


class StubCtx:
    """Stands in for obgpu.Context: results derive from the batch so that misrouted buffers show."""
    log, lock, active = [], threading.Lock(), {"upload": 0, "solve": 0, "download": 0}
    peak = {"upload": 0, "solve": 0, "download": 0}

    def __init__(self, device):
        self.device, self.batch, self.closed = device, None, False

    def _enter(self, what):
        with StubCtx.lock:
            StubCtx.active[what] += 1
            StubCtx.peak[what] = max(StubCtx.peak[what], StubCtx.active[what])

    def _leave(self, what):
        with StubCtx.lock:
            StubCtx.active[what] -= 1

    def upload(self, batch, device=False):
        assert device
        self._enter("upload"); time.sleep(0.002); self._leave("upload")
        self.batch = batch
        with StubCtx.lock:
            StubCtx.log.append(("upload", id(self), batch["tag"]))

    def symbolic(self):
        pass

    def solve_static(self, compact_stress=False):
        if self.batch.get("fail"):
            raise obgpu.ObgpuError(obgpu.ESINGULAR, "stub failure")
        self._enter("solve"); time.sleep(0.002); self._leave("solve")
        self.compact = compact_stress

    def synchronize(self):
        pass

    def sizes(self):
        n = self.batch["S"]
        return {"S": n, "nf_tot": 3 * n, "nb_tot": 2 * n, "Etot": 4 * n}

    def download(self, which, out):
        self._enter("download"); time.sleep(0.001); self._leave("download")
        out[...] = self.batch["tag"] * 10 + which
        return out

    def offsets(self, which):
        return np.arange(self.batch["S"] + 1) * (3 if which == obgpu.OFF_FREE else 2)

    def close(self):
        self.closed = True


@pytest.fixture
def no_pinned(monkeypatch):
    monkeypatch.setattr(obgpu, "pinned_empty", lambda shape, dtype=np.float64: np.empty(shape, dtype))
    StubCtx.log.clear()
    for k in StubCtx.peak:
        StubCtx.peak[k] = 0


def test_every_sub_batch_lands_in_its_own_result(no_pinned):
    pipe = pipeline.BatchPipeline(0, workers=3, context=StubCtx)
    subs = [{"S": 5 + i, "tag": i} for i in range(8)]
    res = pipe.run(subs, offsets=True)
    assert len(res) == 8
    for i, r in enumerate(res):
        assert r["U_f"].shape == (1, 3 * (5 + i)) and (r["U_f"] == i * 10 + obgpu.RES_UF).all()
        assert r["F_b"].shape == (1, 2 * (5 + i)) and (r["F_b"] == i * 10 + obgpu.RES_FB).all()
        assert r["stress"].shape == (1, 4 * (5 + i), 2, 3) and (r["stress"] == i * 10 + obgpu.RES_STRESS_COMPACT).all()
        assert r["info"].dtype == np.int32 and r["info"].shape == (1, 5 + i)
        assert len(r["free_off"]) == 6 + i
    # dealt round-robin: worker w takes sub-batches w, w + 3, ... in order
    by_ctx = {}
    for what, cid, tag in StubCtx.log:
        by_ctx.setdefault(cid, []).append(tag)
    assert sorted(map(tuple, by_ctx.values())) == [(0, 3, 6), (1, 4, 7), (2, 5)]
    # one sub-batch at a time per resource
    assert StubCtx.peak == {"upload": 1, "solve": 1, "download": 1}
    # second pass reuses the result arrays, full records on request
    keep = [r["U_f"] for r in res]
    res2 = pipe.run(subs, res, compact_stress=False)
    assert all(a is b["U_f"] for a, b in zip(keep, res2))
    assert res2[0]["stress"].shape == (1, 20, 2, 6)
    pipe.close()
    assert pipe.ctxs == []


def test_errors_reach_the_caller_and_lengths_are_checked(no_pinned):
    pipe = pipeline.BatchPipeline(0, workers=2, context=StubCtx)
    with pytest.raises(ValueError):
        pipe.run([{"S": 1, "tag": 0}], results=[{}, {}])
    with pytest.raises(obgpu.ObgpuError, match="stub failure"):
        pipe.run([{"S": 2, "tag": 0}, {"S": 2, "tag": 1, "fail": True}, {"S": 2, "tag": 2}])
    pipe.close()
