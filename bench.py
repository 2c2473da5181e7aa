#!/usr/bin/env python
"""bench.py — static solves/s of the openbeam hot path on B200 (BASELINE.json metric).

Workload at N=1 (BASELINE.json configs[1]): demo3_frame.yaml x 1,000,000 randomized load cases; a *step* is one
pass over the whole batch: per case the full  assembleProblem -> solveStatic -> postProcCalcStress  pipeline
(element stiffness, assembly, BC elimination, LDL^T factor + solve, reactions, element end forces), i.e. what
one ob-solve run does per structure.  Weak scaling: every rank owns its own 1M cases (case index offset by
rank), no data-path collective; the job's one exchange is the final all-gather of the per-case summaries (N > 1):
the last step's kernel stores them into every rank's gathered array over NVLink (symmetric memory), else NCCL.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config c2|c3|c4|c5] [--cases L]

--config selects the BASELINE.json configuration (default c2 = configs[1], the one the metric is quoted on):
  c3  random small frames (10-60 elements), 524 288 per GPU (= 4M / 8), warp-per-structure fused route
  c4  817 x 817 grid frame, 2.0M DOF, one structure on one GPU: assembly + multigrid-PCG + stress  (N = 1 only)
  c5  408 x 408 grid frames (0.5M DOF each), 8 per GPU (= 64 / 8), their PCGs concurrent on 8 streams

Prints ONE JSON line on rank 0 (contract in the task statement).  `--impl reference` times the CPU oracle
(restatement of the reference's algorithm; the reference itself cannot be built here) on all host
cores.  The oracle is only ever the baseline / checker here — the product path is libobgpu.so.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "static solves/s (FP64; full assemble+factor+solve+stress per load case)"
UNIT = "solves/s"
N_DOF = 12           # m_problem_DoFs.size() of demo3_frame (n_f = 6, n_b = 6)
ALG_BYTES_FULL = 440      # 8*(7 + n_f + n_b + 12E) with the reference's six-component FaceStress record
ALG_BYTES_COMPACT = 296   # 8*(7 + n_f + n_b + 6E): (N, Vy, Mz) per face — Vz, Mx, My are structurally zero in 2D
ALG_BYTES_PER_CASE = ALG_BYTES_COMPACT
QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", dest="steps_given", type=int, default=None)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cases", type=int, default=1_000_000, help="load cases per GPU per step")
    ap.add_argument("--per-case-section", action="store_true", help="variant C2b: per-case A, Iz sweep")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--gather", default="root", choices=["root", "all"],
                    help="c2, N > 1: the job's one exchange — a gather of the summaries to rank 0 (north_star) or an all-gather")
    ap.add_argument("--full-stress", action="store_true",
                    help="c2: six-component stress records (440 B/case, the r01 configuration) instead of compact ones")
    ap.add_argument("--config", default="c2", choices=["c2", "c3", "c4", "c5"])
    ap.add_argument("--grid", type=int, default=0, help="c4/c5: nodes per side (default 817 / 408)")
    ap.add_argument("--structures", type=int, default=0, help="c3/c5: structures per GPU (default 524288 / 8)")
    ap.add_argument("--rtol", type=float, default=1e-10, help="c4/c5: PCG tolerance on the true residual")
    ap.add_argument("--pcg", default="mg", choices=["mg", "jacobi"], help="c4/c5: preconditioner")
    ap.add_argument("--check", action="store_true", help="c4/c5: also compare U_f with scipy splu (CPU, minutes)")
    a = ap.parse_args()
    if a.steps_given is None:
        a.steps = {"c2": 20, "c3": 20, "c4": 10, "c5": 5}[a.config]
    else:
        a.steps = a.steps_given
    return a


def config_dict(args, world):
    global ALG_BYTES_PER_CASE
    ALG_BYTES_PER_CASE = ALG_BYTES_FULL if args.full_stress else ALG_BYTES_COMPACT
    return {
        "stress_record": "six-component FaceStress [E][2][6] (reference layout)" if args.full_stress else
                         "compact [E][2][3] = N, Vy, Mz per face (Vz, Mx, My are structurally zero in 2D; the host mirrors "
                         "expand it into the reference's FaceStress)",
        "workload": "C2: demo3_frame.yaml x %d randomized load cases per GPU (q on beam + Fx,Fy,Mz on 2 nodes%s); "
                    "each case = full assembleProblem+solveStatic(LDLt)+postProcCalcStress" %
                    (args.cases, "; per-case section sweep" if args.per_case_section else ""),
        "cases_per_gpu": args.cases, "n_dof": N_DOF, "n_free": 6, "n_bounded": 6, "elements": 3,
        "parallelism": "independent load cases sharded over %d GPU(s), no data-path collective; one final %s of the "
                       "32-byte per-case summaries of the last step inside the timed region (peer stores of the last "
                       "step's kernel over NVLink + device barrier; NCCL when symmetric memory is unavailable)"
                       % (world, "gather to rank 0" if args.gather == "root" else "all-gather"),
        "l2_policy": "no flush: per-step inputs+outputs (%.0f MB) exceed the 126 MB L2" %
                     (args.cases * (ALG_BYTES_PER_CASE + 48 + 36) / 1e6),
    }


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled every 50 ms DURING the timed region."""

    def __init__(self, gpu_index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(gpu_index), "--query-gpu=" + QUERY,
                                       "--format=csv,noheader,nounits", "-lms", "50"], stdout=self.f,
                                      stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, smax, power, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f:
            c = [x.strip() for x in line.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1])); smax.append(float(c[2])); power.append(float(c[3]))
            except ValueError:
                continue
            for k, nm in enumerate(names):
                if c[5 + k].lower().startswith("active"):
                    reasons.add(nm)
        self.f.close()
        os.unlink(self.f.name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(smax)), "power_w_max": float(max(power)),
                "samples": len(sm), "reasons": sorted(reasons)}


def cpu_kind():
    """("reference", True) when oracle/_ref/libob_ref.so — libopenbeam's own unmodified sources, oracle/Makefile.ref — is
    there (it travels with the snapshot), else ("port", False): the restated oracle."""
    from oracle import pyoracle
    try:
        ok = pyoracle.ref_available()
    except Exception:  # noqa: BLE001
        ok = False
    return ("reference", True) if ok else ("port", False)


CPU_NOTE = {"reference": "libopenbeam's own unmodified hot-path sources (oracle/_ref, built by oracle/Makefile.ref against "
                         "header stand-ins for Eigen/MRPT, which are not in this image), driven through its public API",
            "port": "oracle port of the reference's algorithm (oracle/_ref not present)"}


def cpu_run(args, n_cases, threads, repeats=1):
    """Times the reference's CPU path (oracle/_ref when built, else the oracle port) on `n_cases` of the same workload with
    `threads` host threads (one solveStatic + postProcCalcStress per case, cases dealt to the threads).  Only this function
    (cpu_baseline / --impl reference) touches oracle/."""
    from openbeam_b200 import workloads
    from oracle.pyoracle import OracleProblem
    st = workloads.demo3_frame()
    nl, q, eai = workloads.c2_cases(0, n_cases, args.per_case_section)
    o = OracleProblem(st, reference=cpu_kind()[1])
    best = None
    for _ in range(repeats):
        dt, *_ = o.run_cases(n_cases, nl, q, eai, n_threads=threads, want_outputs=True)
        best = dt if best is None else min(best, dt)
    return best


def reference_arm(args, rank, world):
    if rank != 0:
        return 0
    cores = os.cpu_count() or 1
    # one step = a bounded sample of the workload: 250k of the 1M cases (~0.1-2 s on all cores)
    sample = min(args.cases, 250_000)
    from openbeam_b200 import workloads
    from oracle.pyoracle import OracleProblem
    st = workloads.demo3_frame()
    nl, q, eai = workloads.c2_cases(0, sample, args.per_case_section)
    kind, use_ref = cpu_kind()
    o = OracleProblem(st, reference=use_ref)
    for _ in range(args.warmup):
        o.run_cases(sample, nl, q, eai, n_threads=cores)
    total = 0.0
    for _ in range(args.steps):
        dt, *_ = o.run_cases(sample, nl, q, eai, n_threads=cores)
        total += dt
    v = sample * args.steps / total
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config_dict(args, world),
        "dof_per_s": v * N_DOF,
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": kind,
                         "sample": "%d of the %d cases per step, one solveStatic+postProcCalcStress per case, "
                                   "%d host threads (the reference is single-threaded; harness-level threads); %s"
                                   % (sample, args.cases, cores, CPU_NOTE[kind])},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": CPU_NOTE[kind],
    }
    emit(line)
    return 0


def bind_to_gpu_numa_node(index):
    """Pins this process to the CPUs NVML reports as local to GPU `index`, so that the pinned host buffers of the e2e
    leg land on the memory node next to the GPU's PCIe root (one process per GPU: without this all ranks allocate on
    whatever node they were started on and share one socket's memory and PCIe path).  Best effort."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        cpus = {64 * w + b for w, m in enumerate(words) for b in range(64) if (m >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
    except Exception:  # noqa: BLE001
        pass


def ours_arm(args, rank, local_rank, world):
    import torch
    import torch.distributed as dist
    from openbeam_b200 import obgpu, sharding, workloads

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — libobgpu has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        # NCCL writes its banner ("NCCL version ...") to STDOUT when NCCL_DEBUG is VERSION or higher; stdout carries
        # exactly one JSON line, so its log goes to stderr
        if os.environ.get("NCCL_DEBUG", "").upper() in ("VERSION", "") or "NCCL_DEBUG_FILE" not in os.environ:
            os.environ["NCCL_DEBUG_FILE"] = "/dev/stderr"
        dist.init_process_group("nccl", device_id=dev)
    L, K, W = args.cases, args.steps, args.warmup
    bind_to_gpu_numa_node(local_rank)   # before the pinned buffers are allocated (first touch decides their node)
    ctx = obgpu.Context(local_rank)
    st = workloads.demo3_frame()
    # pinned host buffers: this rank's cases [rank*L, (rank+1)*L)
    h_nl, h_q = obgpu.pinned_empty((L, 6)), obgpu.pinned_empty((L, 1))
    h_eai = obgpu.pinned_empty((L, 9)) if args.per_case_section else None
    workloads.c2_cases(sharding.weak_range(L, rank)[0], L, args.per_case_section, h_nl, h_q, h_eai)
    compact = not args.full_stress
    config_dict(args, world)   # fixes ALG_BYTES_PER_CASE for the record layout
    h_U, h_Fb, h_S = obgpu.pinned_empty((L, 6)), obgpu.pinned_empty((L, 6)), obgpu.pinned_empty((L, 3, 2, 3 if compact else 6))
    h_info = obgpu.pinned_empty((L,), np.int32)

    ctx.upload(st, L, h_nl, h_q, h_eai)   # inputs resident in HBM from here on
    ctx.symbolic()
    ext = torch.cuda.ExternalStream(ctx.stream(), device=dev)
    summ = torch.empty((L, 4), dtype=torch.float64, device=dev)
    gathered = torch.empty((world * L, 4), dtype=torch.float64, device=dev) if world > 1 else None
    peer = None
    if world > 1 and not os.environ.get("OBGPU_BENCH_NCCL_GATHER"):
        try:   # symmetric memory: the last step's kernel stores its summaries straight into every rank's gathered array
            peer = sharding.PeerGather(L, dev)
        except Exception as e:  # noqa: BLE001 - any failure here means "use the NCCL all-gather"
            print("symmetric memory unavailable (%s): final gather through NCCL" % e, file=sys.stderr)
    ctx.bind_summaries(summ.data_ptr())      # the solve kernel writes the per-case summaries itself

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def step():
        ctx.solve_static(keep_Ff=False, compact_stress=compact)   # 1 fused kernel: K1+K2+K4+K5+K6+K10 per case + summaries
        ctx.summaries_device(summ.data_ptr())    # per-case {max|U|, sum Fb_x, sum Fb_y, info}: no launch when the
                                                 # solve kernel already wrote them (a 2nd kernel on the fallback path)

    root_only = args.gather == "root"

    def final_gather():
        # the only collective of the job (north_star: "a final NCCL gather of reaction and displacement
        # summaries"): once, after the last step; the solves themselves exchange nothing
        if world > 1:
            if root_only:
                dist.gather(summ, list(gathered.view(world, L, 4).unbind(0)) if rank == 0 else None, dst=0)
            else:
                sharding.gather_summaries(summ, gathered)

    # peer-store destinations of the last step's kernel: this rank's block inside rank 0's gathered array only (gather), or
    # inside every rank's (all-gather)
    peer_slots = None if peer is None else ([peer.slots[0]] if root_only else peer.slots)

    def run_steps(n):
        """n steps and the job's one exchange.  With symmetric memory the LAST step's kernel also stores its summaries
        into every rank's gathered array (peer stores over NVLink overlapped with its own arithmetic), followed by a
        device-side barrier; otherwise one NCCL all-gather after the last step."""
        if peer is None:
            for _ in range(n):
                step()
            final_gather()
            return
        for _ in range(n - 1):
            step()
        ctx.bind_summary_peers(peer_slots)
        step()
        ctx.bind_summary_peers([])
        peer.fence()

    with torch.cuda.stream(ext):
        run_steps(W)
        if peer is not None:   # self-check outside the timed region: peer stores == NCCL all-gather
            torch.cuda.synchronize(dev)
            final_gather()
            torch.cuda.synchronize(dev)
            if (rank == 0 or not root_only) and not torch.equal(gathered, peer.buf):
                raise RuntimeError("peer-store gather differs from the NCCL gather")
        barrier()
        sampler = ClockSampler(local_rank) if rank == 0 else None
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if world > 1:
            # the host barrier above releases the ranks a few hundred microseconds apart; a device-side barrier on the
            # timed stream lines the GPUs up so that the start skew is not billed to the final exchange
            if peer is not None:
                peer.fence()
            else:
                dist.all_reduce(torch.zeros(1, device=dev))
        def line_up():
            if world > 1:
                if peer is not None:
                    peer.fence()
                else:
                    dist.all_reduce(torch.zeros(1, device=dev))

        def timed_region(n_steps):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            line_up()
            a.record(ext)
            run_steps(n_steps)
            b.record(ext)
            return a, b

        e0.record(ext)
        run_steps(K)
        e1.record(ext)
        barrier()
        first_ms = e0.elapsed_time(e1)
        if world > 1:   # every rank must run the same number of regions (each holds device-side barriers)
            fm = torch.tensor([first_ms], dtype=torch.float64, device=dev)
            dist.all_reduce(fm, op=dist.ReduceOp.MAX)
            first_ms = float(fm.item())
        # The K-step region lasts ~1.5 ms: it is repeated until >= 50 ms of device time have been timed, so that clock
        # samples land inside timed work; the reported figure is the MEDIAN region (each = K steps + the one exchange)
        reps = int(min(200, max(1, np.ceil(50.0 / max(first_ms, 1e-3)))))
        evs = [timed_region(K) for _ in range(reps)]
        barrier()
        region_ms = sorted([first_ms] + [a.elapsed_time(b) for a, b in evs])
        ms_total = region_ms[len(region_ms) // 2]
        # the same job with ONE step: 1M cases per GPU solved once + their gather (nothing amortised)
        evs1 = [timed_region(1) for _ in range(30)]
        barrier()
        single_ms = sorted(a.elapsed_time(b) for a, b in evs1)[15]
        # dominant kernel alone: CUDA-event pair around each fused-kernel launch (obgpu_get_timings)
        kern_ms = []
        for _ in range(K):
            ctx.solve_static(keep_Ff=False, compact_stress=compact)
            ctx.synchronize()
            kern_ms.append(ctx.timings()["solveStatic.2_solve"])
        kernel_name = ctx.kernel_info()
        launches_per_step = int(ctx.timings()["launches"])
        barrier()
        # end-to-end through the public host-buffer entry point: H2D + solve + D2H inside the timed region
        for _ in range(min(W, 3)):
            ctx.solve_cases_host(L, h_nl, h_q, h_eai, h_U, h_Fb, h_S, h_info, compact_stress=compact)
        barrier()
        t0 = time.perf_counter()
        for _ in range(K):
            ctx.solve_cases_host(L, h_nl, h_q, h_eai, h_U, h_Fb, h_S, h_info, compact_stress=compact)
        torch.cuda.synchronize(dev)
        e2e_s = time.perf_counter() - t0
        e2e_launches = ctx.timings()["launches"]
        barrier()
        clocks = sampler.stop() if sampler else None
    if h_info.any():
        raise SystemExit("bench.py: a system reported a non-positive pivot")
    t = torch.tensor([ms_total, e2e_s, float(np.mean(kern_ms)), single_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total, e2e_s, kern_avg_ms, single_ms = t.tolist()
    if rank == 0:
        peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(peaks_path):
            peak, peak_src = json.load(open(peaks_path))["hbm_gbs"], "MEASURED_PEAKS.json hbm_gbs (measured copy)"
        else:
            peak, peak_src = 6650.0, "fallback 6.65 TB/s (B200_PROFILING.md)"
        value = world * L * K / (ms_total * 1e-3)
        alg = ALG_BYTES_PER_CASE + (72 if args.per_case_section else 0)  # + the 3E per-case section values read
        achieved = alg * L / (kern_avg_ms * 1e-3) / 1e9
        jit = kernel_name.startswith("k_tiny_jit")
        traffic = None
        if jit and not args.per_case_section:
            traffic, traffic_src = (TRAFFIC_BYTES_PER_CASE, TRAFFIC_SOURCE) if args.full_stress else (TRAFFIC_BYTES_PER_CASE_COMPACT, TRAFFIC_SOURCE_COMPACT)
        in_bytes = (h_nl.nbytes + h_q.nbytes + (h_eai.nbytes if h_eai is not None else 0))
        out_bytes = h_U.nbytes + h_Fb.nbytes + h_S.nbytes + h_info.nbytes
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms_total / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic", "config": config_dict(args, world),
            "dof_per_s": value * N_DOF,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": None if traffic is None else traffic * L,
                         "traffic_source": traffic_src if traffic is not None else None,
                         "kernel": kernel_name,
                         "frac_counting_summary_record": (alg + 32) * L / (kern_avg_ms * 1e-3) / 1e9 / peak
                         if launches_per_step == 1 else None,  # the kernel also writes the 32-B per-case summary
                         "kernel_ms": kern_avg_ms, "algorithmic_bytes_per_launch": alg * L,
                         "peak_source": peak_src},
            "e2e": {"value": world * L * K / e2e_s, "unit": UNIT, "h2d_bytes_per_step": in_bytes,
                    "d2h_bytes_per_step": out_bytes,
                    "api": "obgpu_solve_cases_host (pinned host buffers in and out, chunked over 3 streams)"},
            "timed_regions": {"count": len(region_ms), "ms_min": region_ms[0], "ms_median": region_ms[len(region_ms) // 2],
                              "ms_max": region_ms[-1], "note": "value = world*cases*steps / median region"},
            "single_pass_job": {"value": world * L / (single_ms * 1e-3), "unit": UNIT, "ms": single_ms,
                                "what": "ONE step (every GPU solves its cases once) + the gather of all its summaries, "
                                        "median of 30; the K-step figure above exchanges the summaries of its last step "
                                        "only, i.e. 1/%d summary per counted solve" % K},
            "gather": ("none (1 GPU)" if world == 1 else
                       ("gather to rank 0" if root_only else "all-gather") +
                       (": peer stores of the last step's kernel over NVLink (symmetric memory)" if peer is not None else ": NCCL")),
            "gpu_launches": int(K * launches_per_step), "e2e_launches_per_step": int(e2e_launches),
            "clocks": clocks,
        }
        if not args.no_cpu_baseline:
            cores = os.cpu_count() or 1
            kind = cpu_kind()[0]
            # the reference's own code is ~3x slower per case than the port: keep the leg within its 10-30 s budget
            sample = min(L, 400_000 if kind == "reference" else 1_000_000)
            sample1 = min(sample, 100_000 if kind == "reference" else 200_000)
            dt_all = cpu_run(args, sample, cores)
            dt_1 = cpu_run(args, sample1, 1)
            line["cpu_baseline"] = {
                "value": sample / dt_all, "unit": UNIT, "cores": cores, "kind": kind, "what": CPU_NOTE[kind],
                "value_1core": sample1 / dt_1,
                "sample": "%d cases of the same workload on %d host threads (and %d cases on 1 thread): "
                          "solveStatic+postProcCalcStress, one call per case" % (sample, cores, sample1)}
        emit(line)
    ctx.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


# dram__bytes_read.sum + dram__bytes_write.sum of one k_tiny_jit launch (summaries bound, F_f skipped), per case, from the
# committed `ncu --set full` capture of the current kernel at 1M cases: 56.03 MB read + 361.29 MB written (the kernel
# hands 420 MB of records to L2; what is still dirty in L2 when the kernel ends is not in the DRAM counter)
TRAFFIC_BYTES_PER_CASE = 417.3
TRAFFIC_SOURCE = "profiles/r01o_ncu_full_k_tiny_jit_summary.txt (1M-case launch, scaled per case)"
# compact (N, Vy, Mz) records: dram__bytes_read 56.03 MB + dram__bytes_write 217.43 MB per 1M-case launch; below the 296 MB
# algorithmic figure (+ 32 MB of summaries) because the tail of a launch's stores is still dirty in the 126 MB L2
TRAFFIC_BYTES_PER_CASE_COMPACT = 273.5
TRAFFIC_SOURCE_COMPACT = "profiles/r02e_tiny_jit_compact_ncu_full_summary.txt (1M-case launch, scaled per case)"


def peak_info():
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        return json.load(open(peaks_path))["hbm_gbs"], "MEASURED_PEAKS.json hbm_gbs (measured copy)"
    return 6650.0, "fallback 6.65 TB/s (B200_PROFILING.md)"


_JSON_FD = None


def emit(line):
    """The one JSON line, on the process's ORIGINAL stdout."""
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def main():
    global _JSON_FD
    args = parse_args()
    # stdout carries exactly one JSON line: native libraries (the NCCL banner, NVRTC) write to fd 1 behind Python's back,
    # so fd 1 is pointed at stderr for the life of the process and the line goes to a saved copy of the real stdout
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.config == "c3":
        import bench_frames
        if args.impl == "reference":
            return bench_frames.reference(args, rank, world, emit)
        return bench_frames.ours(args, rank, local_rank, world, emit, ClockSampler, peak_info)
    if args.config in ("c4", "c5"):
        import bench_lattice
        if args.impl == "reference":
            return bench_lattice.reference(args, rank, world, emit)
        return bench_lattice.ours(args, rank, local_rank, world, emit, ClockSampler, peak_info)
    if args.impl == "reference":
        return reference_arm(args, rank, world)
    return ours_arm(args, rank, local_rank, world)


if __name__ == "__main__":
    sys.exit(main())
