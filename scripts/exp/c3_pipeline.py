"""Experiment: C3 host-to-host throughput of BatchPipeline for several (workers, sub-batches) against one context."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
from openbeam_b200 import obgpu, workloads
from openbeam_b200.pipeline import BatchPipeline

S = int(sys.argv[1]) if len(sys.argv) > 1 else 524288
pin = lambda shape, dt: obgpu.pinned_empty(shape, dt)
for workers, nsub in ((1, 1), (4, 8), (5, 8), (5, 10), (6, 12), (5, 16), (8, 16)):
    per = S // nsub
    batches = [workloads.c3_batch(i * per, per, pinned=pin) for i in range(nsub)]
    pipe = BatchPipeline(0, workers)
    res = None
    for _ in range(2):
        res = pipe.run(batches, res)
    t0 = time.perf_counter()
    K = 5
    for _ in range(K):
        res = pipe.run(batches, res)
    dt = (time.perf_counter() - t0) / K
    for r in res:
        r["trace"] = []
    t0 = time.perf_counter()
    pipe.run(batches, res)
    if (workers, nsub) == (5, 8):
        for i, r in enumerate(res):
            print("   sub %2d worker %d: " % (i, i % workers) + "  ".join("%s %.1f-%.1f" % (n, (a - t0) * 1e3, (b - t0) * 1e3) for n, a, b in r["trace"]))
    print("workers %d sub-batches %2d: %.1f ms per %d structures = %.2f M structures/s" % (workers, nsub, dt * 1e3, per * nsub, per * nsub / dt / 1e6), flush=True)
    pipe.close()
    del batches, res
