"""Pinned-memory PCIe ceiling with 1, 2, 4, 8 GPUs copying at the same time (the end-to-end leg of bench.py is
PCIe-bound; this is the number it should be compared with).  One process, one stream pair per GPU, `chunk` bytes per
cudaMemcpyAsync, H2D and D2H together like obgpu_solve_cases_host (56 B in / 244 B out per C2 case).
    python scripts/copy_ceiling.py [--mb 512] [--chunk-mb 32]
Prints one JSON line per GPU count."""
import argparse
import json
import time

import torch

ap = argparse.ArgumentParser()
ap.add_argument("--mb", type=int, default=512)
ap.add_argument("--chunk-mb", type=int, default=32)
a = ap.parse_args()
n_gpu = torch.cuda.device_count()
nbytes, chunk = a.mb << 20, a.chunk_mb << 20
host_in = [torch.empty(nbytes, dtype=torch.uint8).pin_memory() for _ in range(n_gpu)]
host_out = [torch.empty(nbytes, dtype=torch.uint8).pin_memory() for _ in range(n_gpu)]
dev_in = [torch.empty(nbytes, dtype=torch.uint8, device="cuda:%d" % g) for g in range(n_gpu)]
dev_out = [torch.empty(nbytes, dtype=torch.uint8, device="cuda:%d" % g) for g in range(n_gpu)]
s_in = [torch.cuda.Stream(device=g) for g in range(n_gpu)]
s_out = [torch.cuda.Stream(device=g) for g in range(n_gpu)]


def run(gpus, h2d, d2h):
    for g in gpus:
        torch.cuda.synchronize(g)
    t0 = time.perf_counter()
    for off in range(0, nbytes, chunk):
        for g in gpus:
            if h2d:
                with torch.cuda.stream(s_in[g]):
                    dev_in[g][off:off + chunk].copy_(host_in[g][off:off + chunk], non_blocking=True)
            if d2h:
                with torch.cuda.stream(s_out[g]):
                    host_out[g][off:off + chunk].copy_(dev_out[g][off:off + chunk], non_blocking=True)
    for g in gpus:
        torch.cuda.synchronize(g)
    return time.perf_counter() - t0


for n in (1, 2, 4, 8):
    if n > n_gpu:
        break
    gpus = list(range(n))
    out = {"gpus": n, "mb_per_gpu_per_direction": a.mb, "chunk_mb": a.chunk_mb}
    for name, h2d, d2h in (("h2d", True, False), ("d2h", False, True), ("both", True, True)):
        run(gpus, h2d, d2h)
        dt = min(run(gpus, h2d, d2h) for _ in range(3))
        out[name + "_GBps_total"] = n * nbytes * (int(h2d) + int(d2h)) / dt / 1e9
    # the C2 end-to-end leg moves 56 B in and 244 B out per case: its ceiling is set by the D2H direction
    out["c2_e2e_ceiling_Msolves_per_s"] = out["d2h_GBps_total"] * 1e9 / 244 / 1e6
    print(json.dumps(out), flush=True)
