"""bench.py leg for BASELINE.json config C3: "parametric sweep: 4M random small 2D frames (10-60 elements each, mixed
RR/RA/AA ends) sharded across 8 x B200" — 524 288 DIFFERENT frames per GPU (= 4M / 8), weak scaling.  Part of the bench
harness, not of the product package (its cpu_baseline / --impl reference legs execute oracle/).

A *step* = one pass over the rank's frames: per structure the complete solveStatic + postProcCalcStress (element
stiffness, assembly into the banded K_ff, BC elimination, LDL^T, solve, reactions, element end forces) in the
warp-per-structure kernel, one launch per size class.
  value  the batch and its symbolic products resident in HBM (uploaded and numbered once, before the timed region)
  e2e    per step, from pinned HOST buffers to pinned HOST buffers: the rank's frames cut into 8 sub-batches and dealt to
         4 contexts (openbeam_b200/pipeline.py); per sub-batch obgpu_batch_upload_device (H2D) + obgpu_symbolic (DOF
         numbering, partition, slots as device kernels) + obgpu_solve_static + download of U_f, F_b, compact stresses,
         info — the H2D copy of one sub-batch overlaps the kernels of another and the D2H copy of a third
"""
from __future__ import annotations

import os
import time

import numpy as np

from openbeam_b200 import obgpu, sharding, workloads
from openbeam_b200.pipeline import BatchPipeline

# dram__bytes_read + dram__bytes_write of the three-kernel path, all launches of one solve of 262 144 frames together:
# k_frame_pre 545 + 1 668 MB, the nine k_frame_fused<split> launches 1 940 + 165 MB, k_frame_post 884 + 677 MB.  4.7 x the
# algorithmic bytes: the 240-byte element records make a round trip through HBM, and gathers of 16-byte node records fetch
# whole sectors — on a path that uses 3 % of the HBM bandwidth
TRAFFIC_PER_STRUCTURE = 5879e6 / 262144
TRAFFIC_SOURCE = "profiles/README.md r02o table + profiles/r02o_ncu_full_k_frame_split_summary.txt (262 144 frames, scaled per structure)"
E2E_SUB, E2E_WORKERS = 8, 4   # sub-batches per step and contexts of the end-to-end leg (scripts/exp/c3_pipeline.py)

METRIC = "static solves/s (FP64; full assemble+factor+solve+stress per load case)"
UNIT = "solves/s"
PER_GPU = 524_288


def config_dict(args, world, z=None):
    S = args.structures or PER_GPU
    d = {
        "workload": "C3: %d different random 2D frames per GPU (b in 1..5 bays x s in 1..6 storeys, 10-60 elements, beams "
                    "RR .6 / RA .15 / AR .15 / AA .1, random IPE sections, UDL on every beam + wind loads; counter RNG, "
                    "rank r owns frames [r*S, (r+1)*S)); each structure = full assembleProblem+solveStatic(LDLt)+"
                    "postProcCalcStress" % S,
        "structures_per_gpu": S,
        "parallelism": "independent structures sharded over %d GPU(s), no data-path collective; one final NCCL all-gather "
                       "of the 32-byte per-structure summaries inside the timed region" % world,
        "stress_record": "value leg: six-component FaceStress; e2e leg: compact (N, Vy, Mz) records",
    }
    if z:
        d.update({"n_dof_total_per_gpu": z["n_tot"], "n_free_total_per_gpu": z["nf_tot"], "elements_per_gpu": z["Etot"],
                  "l2_policy": "no flush: the batch (%.0f MB in, %.0f MB out) exceeds the 126 MB L2"
                               % (alg_bytes_in(z) / 1e6, alg_bytes_out(z) / 1e6)})
    return d


def alg_bytes_in(z):    # 16N + 33E + 12 n_b + 12 #loads
    return 16 * z["Ntot"] + 33 * z["Etot"] + 12 * z["nb_tot"] + 12 * (z["NLtot"] + z["ELtot"])


def alg_bytes_out(z):   # 8n + 8 n_b + 96E
    return 8 * z["n_tot"] + 8 * z["nb_tot"] + 96 * z["Etot"]


def flops_estimate(batch):
    """FP64 flops of the banded LDL^T + solves + element work, from the generator's (bays, storeys): n_f = 3(b+1)s,
    half bandwidth 3(b+2)-1; trailing update hb(hb+1) per column, two band solves 4 hb per row, ~300 per element."""
    nN = np.diff(batch["node_off"]).astype(np.float64)
    nE = np.diff(batch["elem_off"]).astype(np.float64)
    nC = np.diff(batch["con_off"]).astype(np.float64)
    b1 = nC / 3.0                      # b + 1
    nf = 3.0 * (nN - b1)
    hb = 3.0 * (b1 + 1.0) - 1.0
    return float(np.sum(nf * hb * (hb + 1.0) + 4.0 * nf * hb + 300.0 * nE))


def ours(args, rank, local_rank, world, emit, ClockSampler, peak_info):
    import torch
    import torch.distributed as dist

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — libobgpu has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        if "NCCL_DEBUG_FILE" not in os.environ:
            os.environ["NCCL_DEBUG_FILE"] = "/dev/stderr"
        dist.init_process_group("nccl", device_id=dev)
    K, W = args.steps, args.warmup
    S = args.structures or PER_GPU
    ctx = obgpu.Context(local_rank)
    t0 = time.perf_counter()
    batch = workloads.c3_batch(rank * S, S, pinned=lambda shape, dt: obgpu.pinned_empty(shape, dt))
    t_gen = time.perf_counter() - t0
    t0 = time.perf_counter(); ctx.upload(batch, device=True); t_upload = time.perf_counter() - t0
    t0 = time.perf_counter(); ctx.symbolic(); t_sym = time.perf_counter() - t0
    z = ctx.sizes()
    ext = torch.cuda.ExternalStream(ctx.stream(), device=dev)
    summ = torch.empty((S, 4), dtype=torch.float64, device=dev)
    gathered = torch.empty((world * S, 4), dtype=torch.float64, device=dev) if world > 1 else None

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    flops = flops_estimate(batch)
    in_bytes = int(sum(v.nbytes for v in batch.values() if isinstance(v, np.ndarray)))
    ctx._keep = None
    del batch            # the value leg works on the device-resident batch: give the pinned host copy back

    with torch.cuda.stream(ext):
        for _ in range(W):
            ctx.solve_static()
        barrier()
        sampler = ClockSampler(local_rank) if rank == 0 else None
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if world > 1:
            dist.all_reduce(torch.zeros(1, device=dev))
        e0.record(ext)
        for _ in range(K):
            ctx.solve_static()
        ctx.summaries_device(summ.data_ptr())
        if world > 1:
            sharding.gather_summaries(summ, gathered)   # the job's one exchange
        e1.record(ext)
        barrier()
        ms_total = e0.elapsed_time(e1)
        kern_ms = []
        for _ in range(min(K, 5)):
            ctx.solve_static()
            ctx.synchronize()
            kern_ms.append(ctx.timings()["solveStatic.2_solve"])
        launches_per_step = int(ctx.timings()["launches"])
        kernel_name = ctx.kernel_info()
        if ctx.download(obgpu.RES_INFO).any():
            raise SystemExit("bench.py: a structure reported a non-positive pivot")
        U_res = ctx.download(obgpu.RES_UF)
        barrier()
        # ---- end to end: host buffers in, host buffers out, every step ----
        per = S // E2E_SUB
        subs = [workloads.c3_batch(rank * S + i * per, per if i < E2E_SUB - 1 else S - per * (E2E_SUB - 1),
                                   pinned=lambda shape, dt: obgpu.pinned_empty(shape, dt)) for i in range(E2E_SUB)]
        pipe = BatchPipeline(local_rank, E2E_WORKERS)
        res = None
        for _ in range(2):
            res = pipe.run(subs, res)
        barrier()
        # the K steps are streamed: step k+1's first uploads run while step k's last results leave (no drain between steps).
        # Sub-batch i of every step lands in the same result buffers; E2E_SUB is a multiple of E2E_WORKERS, so those
        # reuses are sequential on one worker.
        assert E2E_SUB % E2E_WORKERS == 0
        t0 = time.perf_counter()
        pipe.run(subs * K, res * K)
        torch.cuda.synchronize(dev)
        e2e_s = time.perf_counter() - t0
        out_bytes = int(sum(r[k].nbytes for r in res for k in ("U_f", "F_b", "stress", "info")))
        e2e_same = bool(np.array_equal(np.concatenate([r["U_f"][0] for r in res]), U_res[0]))
        if any(r["info"].any() for r in res):
            raise SystemExit("bench.py: a structure of the end-to-end leg reported a non-positive pivot")
        pipe.close()
        t_last = ctx.timings()
        barrier()
        clocks = sampler.stop() if sampler else None
    tt = torch.tensor([ms_total, e2e_s, float(np.mean(kern_ms))], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    ms_total, e2e_s, k_ms = tt.tolist()
    if rank == 0:
        peak, peak_src = peak_info()
        value = world * S * K / (ms_total * 1e-3)
        alg = alg_bytes_in(z) + alg_bytes_out(z)
        achieved = alg / (k_ms * 1e-3) / 1e9
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms_total / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic", "config": config_dict(args, world, z),
            "dof_per_s": value * z["n_tot"] / S,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": TRAFFIC_PER_STRUCTURE * S if "k_frame_pre" in kernel_name else None,
                         "traffic_source": TRAFFIC_SOURCE if "k_frame_pre" in kernel_name else None,
                         "kernel": kernel_name, "kernel_ms": k_ms,
                         "algorithmic_bytes_per_launch": alg,
                         "note": "all size-class launches of one step together (%d launches); this kernel is bound by "
                                 "instruction issue and shared-memory latency, not by HBM (DESIGN.md 4)" % launches_per_step,
                         "fp64": {"flops_per_step_estimate": flops, "achieved_tflops": flops / (k_ms * 1e-3) / 1e12,
                                  "peak_tflops": 36.6, "peak_source": "profiles/r02f_fp64_peak.jsonl (measured DFMA peak)",
                                  "frac": flops / (k_ms * 1e-3) / 1e12 / 36.6},
                         "peak_source": peak_src},
            "e2e": {"value": world * S * K / e2e_s, "unit": UNIT, "h2d_bytes_per_step": in_bytes,
                    "d2h_bytes_per_step": out_bytes,
                    "api": "openbeam_b200.pipeline.BatchPipeline: %d sub-batches over %d contexts; per sub-batch "
                           "obgpu_batch_upload_device (pinned host arrays) + obgpu_symbolic (device kernels) + "
                           "obgpu_solve_static + obgpu_download of U_f, F_b, compact stress, info into pinned host "
                           "buffers; the K steps streamed back to back (every step copies all its inputs in and all its "
                           "results out)" % (E2E_SUB, E2E_WORKERS),
                    "U_f_identical_to_resident_leg": e2e_same, "symbolic_ms_whole_batch": t_last["symbolic"]},
            "gpu_launches": int(K * launches_per_step),
            "once_s": {"generate_inputs_host": t_gen, "upload": t_upload, "symbolic_device": t_sym},
            "clocks": clocks,
        }
        if not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(args)
        emit(line)
    ctx.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


def cpu_sample(n, threads):
    """the reference's CPU path (oracle/_ref = libopenbeam's own sources when built, else the oracle port) on the first n
    frames of the workload, one structure at a time per host thread"""
    from oracle import pyoracle
    use_ref = pyoracle.ref_available()
    batch = workloads.c3_batch(0, n)
    dt, _ = pyoracle.run_batch(batch, threads, use_ref)
    return dt, ("reference" if use_ref else "port")


def cpu_baseline(args):
    cores = os.cpu_count() or 1
    n = 40_000
    dt, kind = cpu_sample(n, cores)
    n1 = 5_000
    dt1, _ = cpu_sample(n1, 1)
    return {"value": n / dt, "unit": UNIT, "cores": cores, "kind": kind, "value_1core": n1 / dt1,
            "sample": "the first %d frames of the same workload on %d host threads (and %d on 1 thread): per structure a "
                      "full problem set-up + solveStatic + postProcCalcStress; kind 'reference' = libopenbeam's own unmodified "
                      "sources (oracle/_ref)" % (n, cores, n1)}


def reference(args, rank, world, emit):
    if rank != 0:
        return 0
    cores = os.cpu_count() or 1
    n = 20_000
    kind = "port"
    for _ in range(min(args.warmup, 1)):
        cpu_sample(2000, cores)
    tot = 0.0
    for _ in range(args.steps):
        dt, kind = cpu_sample(n, cores)
        tot += dt
    v = n * args.steps / tot
    emit({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * tot / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config_dict(args, world),
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": kind,
                         "sample": "each step: the first %d frames of the workload, %d host threads, per structure a full "
                                   "problem set-up + solveStatic + postProcCalcStress" % (n, cores)},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "kind 'reference' = libopenbeam's own unmodified hot-path sources compiled against header stand-ins for "
                "Eigen/MRPT (oracle/Makefile.ref); 'port' = the restated oracle",
    })
    return 0
