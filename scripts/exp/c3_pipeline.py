"""Experiment: C3 host-to-host throughput of BatchPipeline for several (workers, sub-batches) against one context."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
from openbeam_b200 import obgpu, workloads
from openbeam_b200.pipeline import BatchPipeline

S = int(sys.argv[1]) if len(sys.argv) > 1 else 524288
pin = lambda shape, dt: obgpu.pinned_empty(shape, dt)
K = 5
for workers, nsub, slots in ((1, 1, 1), (4, 8, 1), (4, 8, 2), (6, 12, 2), (6, 6, 2), (8, 8, 2), (4, 4, 1)):
    per = S // nsub
    batches = [workloads.c3_batch(i * per, per, pinned=pin) for i in range(nsub)]
    pipe = BatchPipeline(0, workers, copy_slots=slots)
    res = None
    for _ in range(2):
        res = pipe.run(batches, res)
    t0 = time.perf_counter()
    pipe.run(batches * K, res * K)
    dt = (time.perf_counter() - t0) / K
    print("workers %d sub-batches %2d copy slots %d, %d steps streamed: %.1f ms per %d structures = %.2f M structures/s" % (workers, nsub, slots, K, dt * 1e3, per * nsub, per * nsub / dt / 1e6), flush=True)
    pipe.close()
    del batches, res
