"""Host-side pipeline for jobs of MANY small structures (BASELINE config C3) whose inputs and results live in host memory.

One obgpu context works through upload -> symbolic -> solve -> download strictly in sequence, so a host-to-host job is
bound by the SUM of its two PCIe directions plus the kernels.  `BatchPipeline` cuts the job into sub-batches and deals
them to a few contexts (one host thread and one CUDA stream each, all on the same GPU): while one context copies its
results out, another copies its inputs in and a third runs its kernels, so the job approaches max(H2D, D2H, kernels).
Everything goes through the C-ABI calls a single context uses (obgpu_batch_upload_device, obgpu_symbolic,
obgpu_solve_static, obgpu_download); ctypes releases the GIL for the duration of every call.

The structures of different sub-batches are independent problems (the reference solves one structure per process,
apps/ob-solve/ob-solve_main.cpp), so cutting the list anywhere changes no result.
"""
import threading
import time

import numpy as np

from . import obgpu


class BatchPipeline:
    def __init__(self, device=0, workers=3, copy_slots=1, context=obgpu.Context):
        self.device = device
        self.ctxs = [context(device) for _ in range(max(1, int(workers)))]   # `context`: the ctx class (tests pass a stub)
        # one sub-batch at a time per resource (H2D copy engine, SMs, D2H copy engine): the workers then fall into a
        # staggered rhythm in which the three resources are busy at once.  Without the locks all workers upload together,
        # compute together and download together, and the two PCIe directions never overlap.
        self._h2d, self._sm, self._d2h = threading.Semaphore(copy_slots), threading.Lock(), threading.Semaphore(copy_slots)

    def close(self):
        for c in self.ctxs:
            c.close()
        self.ctxs = []

    @staticmethod
    def _buf(res, key, shape, dtype):
        a = res.get(key)
        if a is None or a.shape != tuple(shape) or a.dtype != np.dtype(dtype):
            res[key] = obgpu.pinned_empty(shape, dtype)
        return res[key]

    def _one(self, ctx, batch, res, compact_stress, offsets):
        tr = res.get("trace")  # optional: a list that receives (stage, t_begin, t_end) wall-clock triples
        with self._h2d:
            t0 = time.perf_counter()
            ctx.upload(batch, device=True)
            ctx.synchronize()
        t1 = time.perf_counter()
        ctx.symbolic()  # small kernels and host round trips: latency, not throughput — runs beside the other stages
        with self._sm:
            t2 = time.perf_counter()
            ctx.solve_static(compact_stress=compact_stress)
            ctx.synchronize()
        z = ctx.sizes()
        bufs = [(obgpu.RES_UF, self._buf(res, "U_f", (1, z["nf_tot"]), np.float64)),
                (obgpu.RES_FB, self._buf(res, "F_b", (1, z["nb_tot"]), np.float64)),
                (obgpu.RES_STRESS_COMPACT, self._buf(res, "stress", (1, z["Etot"], 2, 3), np.float64)) if compact_stress
                else (obgpu.RES_STRESS, self._buf(res, "stress", (1, z["Etot"], 2, 6), np.float64)),
                (obgpu.RES_INFO, self._buf(res, "info", (1, z["S"]), np.int32))]
        with self._d2h:
            t3 = time.perf_counter()
            for which, out in bufs:
                ctx.download(which, out)
        if offsets:
            res["free_off"] = ctx.offsets(obgpu.OFF_FREE)
            res["bnd_off"] = ctx.offsets(obgpu.OFF_BOUNDED)
        if tr is not None:
            tr[:] = [("upload", t0, t1), ("symbolic", t1, t2), ("solve", t2, t3), ("download", t3, time.perf_counter())]

    def run(self, batches, results=None, compact_stress=True, offsets=False):
        """Solve every packed sub-batch of `batches` (dicts as obgpu.pack_batch / workloads.c3_batch build them; pinned
        arrays make the copies asynchronous).  `results`: the list an earlier call returned — its pinned result arrays
        are reused where the shapes still fit.  Returns one dict per sub-batch: U_f [1][nf_tot], F_b [1][nb_tot],
        stress [1][Etot][2][3] (or [..][6]), info [1][S], and with offsets=True free_off / bnd_off [S+1]."""
        if results is None:
            results = [dict() for _ in batches]
        if len(results) != len(batches):
            raise ValueError("results must have one entry per sub-batch")
        errors = []

        def work(w):
            try:
                for i in range(w, len(batches), len(self.ctxs)):
                    self._one(self.ctxs[w], batches[i], results[i], compact_stress, offsets)
            except BaseException as e:  # re-raised in the caller's thread
                errors.append(e)

        threads = [threading.Thread(target=work, args=(w,)) for w in range(min(len(self.ctxs), len(batches)))]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        if errors:
            raise errors[0]
        return results
