// capi_sparse.cu — global K assembly (K1+K4), BC gather (K5) and the PCG route (K7-K9, K10) for large structures.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>

#include "ctx.h"

namespace obgpu {

int ensure_device_batch(obgpu_ctx* c);  // capi_general.cu

static int ensure_plan(obgpu_ctx* c) {
  if (c->dev_plan) return OBGPU_OK;
  int rc = ensure_device_batch(c);
  if (rc != OBGPU_OK) return rc;
  std::string err;
  if (!c->sym.has_pattern) {  // pattern was skipped for a very large batch: build it now
    rc = run_symbolic(c->hb, c->sym, err, true);
    if (rc != OBGPU_OK) return fail(c, rc, err);
  }
  rc = build_sparse_plan(c->hb, c->sym, c->plan, err);
  if (rc != OBGPU_OK) return fail(c, rc, err);
  const HostBatch& hb = c->hb;
  const SparsePlan& P = c->plan;
  cudaStream_t st = c->stream;
  std::vector<int32_t> node0(hb.Etot()), f0(hb.Etot()), eg(hb.ELtot());
  for (int64_t s = 0; s < hb.S; s++) {
    for (int64_t e = hb.elem_off[s]; e < hb.elem_off[s + 1]; e++) {
      node0[e] = int32_t(hb.node_off[s]);
      f0[e] = int32_t(c->sym.free_off[s]);
    }
    for (int64_t l = hb.eload_off[s]; l < hb.eload_off[s + 1]; l++) eg[l] = int32_t(hb.elem_off[s] + hb.eload_elem[l]);
  }
  OB_CUDA(c, c->d_elem_node0.upload(node0, st));
  OB_CUDA(c, c->d_elem_f0.upload(f0, st));
  OB_CUDA(c, c->d_eload_elem_g.upload(eg, st));
  OB_CUDA(c, c->d_seg_ptr.upload(P.seg_ptr, st));
  OB_CUDA(c, c->d_seg_src.upload(P.seg_src, st));
  if (P.asm_max_loc > 0) {
    OB_CUDA(c, c->d_blk_elem_ptr.upload(P.blk_elem_ptr, st));
    OB_CUDA(c, c->d_blk_elems.upload(P.blk_elems, st));
    OB_CUDA(c, c->d_seg_loc.upload(P.seg_loc, st));
  }
  OB_CUDA(c, c->d_sl_ptr.upload(P.sl_ptr, st));
  OB_CUDA(c, c->d_ff_col.upload(P.ff_col, st));
  OB_CUDA(c, c->d_ff_src.upload(P.ff_src, st));
  OB_CUDA(c, c->d_ff_diag.upload(P.ff_diag, st));
  OB_CUDA(c, c->d_bf_rowptr.upload(P.bf_rowptr, st));
  OB_CUDA(c, c->d_bf_col.upload(P.bf_col, st));
  OB_CUDA(c, c->d_bf_src.upload(P.bf_src, st));
  OB_CUDA(c, c->d_fseg_ptr.upload(P.fseg_ptr, st));
  OB_CUDA(c, c->d_fseg_src.upload(P.fseg_src, st));
  OB_CUDA(c, c->d_eload_g.reserve(std::max<size_t>(size_t(hb.ELtot()) * 6, 1)));
  OB_CUDA(c, c->d_Eq.reserve(std::max<size_t>(c->sym.nb_tot(), 1)));
  OB_CUDA(c, cudaStreamSynchronize(st));  // node0/f0/eg are stack vectors
  c->dev_plan = true;
  return OBGPU_OK;
}

static SparseDev sparse_dev(obgpu_ctx* c, int64_t k /* K copy */) {
  SparseDev d;
  d.Etot = c->hb.Etot(); d.ELtot = c->hb.ELtot(); d.nnz = c->sym.nnz_tot();
  d.nf = c->sym.nf_tot(); d.nb = c->sym.nb_tot();
  d.nnz_ff = int64_t(c->plan.ff_col.size()); d.nnz_bf = int64_t(c->plan.bf_col.size());
  d.elem_nodes = reinterpret_cast<const uint2*>(c->d_elem_nodes.p);
  d.elem_type = c->d_elem_type.p;
  d.node_xy = reinterpret_cast<const double2*>(c->d_node_xy.p);
  d.elem_node0 = c->d_elem_node0.p; d.elem_f0 = c->d_elem_f0.p;
  d.node_slot = c->d_node_slot.p; d.el_ptr = c->d_el_ptr.p; d.el_order = c->d_el_order.p;
  d.eload_elem_g = c->d_eload_elem_g.p; d.eload_dir = c->d_eload_dir.p;
  d.eload_type = c->hb.has_load_kinds ? c->d_eload_type.p : nullptr;
  d.eload_q2 = c->hb.has_load_kinds ? c->d_eload_q2.p : nullptr;
  d.seg_ptr = c->d_seg_ptr.p; d.seg_src = c->d_seg_src.p;
  static const bool fused_ok = !(getenv("OBGPU_ASM_FUSED") && atoi(getenv("OBGPU_ASM_FUSED")) == 0);
  d.asm_max_loc = fused_ok ? c->plan.asm_max_loc : 0;
  d.blk_elem_ptr = c->d_blk_elem_ptr.p; d.blk_elems = c->d_blk_elems.p; d.seg_loc = c->d_seg_loc.p;
  d.sl_ptr = c->d_sl_ptr.p; d.ff_col = c->d_ff_col.p; d.ff_src = c->d_ff_src.p; d.ff_diag = c->d_ff_diag.p;
  d.bf_rowptr = c->d_bf_rowptr.p; d.bf_col = c->d_bf_col.p; d.bf_src = c->d_bf_src.p;
  d.fseg_ptr = c->d_fseg_ptr.p; d.fseg_src = c->d_fseg_src.p;
  d.elem_k = c->d_elem_k.p + size_t(k) * d.Etot * ELEM_K;
  d.kval = c->d_kval.p + size_t(k) * d.nnz;
  d.ff_val = c->d_ff_val.p; d.bf_val = c->d_bf_val.p; d.eload_g = c->d_eload_g.p; d.Eq = c->d_Eq.p;
  return d;
}

// K1 + K4 for every K copy, and F_f for every case (what assembleProblem() leaves in BuildProblemInfo).
int assemble_general(obgpu_ctx* c) {
  int rc = ensure_plan(c);
  if (rc != OBGPU_OK) return rc;
  if (c->hb.has_rotz) return fail(c, OBGPU_EUNSUPPORTED, "nodal rotated frames are not implemented on the GPU path yet");
  const int64_t L = c->L, LK = c->has_case_eai ? L : 1;
  c->LK = LK;
  const int64_t Etot = c->hb.Etot(), nnz = c->sym.nnz_tot(), nf = c->sym.nf_tot();
  const bool fused = sparse_dev(c, 0).asm_max_loc > 0;
  if (!fused) OB_CUDA(c, c->d_elem_k.reserve(size_t(LK) * Etot * ELEM_K));
  c->elem_k_valid = !fused;
  OB_CUDA(c, c->d_kval.reserve(size_t(LK) * std::max<int64_t>(nnz, 1)));
  OB_CUDA(c, c->d_ff_val.reserve(std::max<size_t>(c->plan.ff_col.size(), 1)));
  OB_CUDA(c, c->d_bf_val.reserve(std::max<size_t>(c->plan.bf_col.size(), 1)));
  OB_CUDA(c, c->d_Ff.reserve(size_t(L) * std::max<int64_t>(nf, 1)));
  OB_CUDA(c, cudaEventRecord(c->ev[6], c->stream));
  double launches = 0;
  for (int64_t k = 0; k < LK; k++) {
    SparseDev d = sparse_dev(c, k);
    const double* eai = c->has_case_eai ? c->d_case_eai.p + size_t(k) * 3 * Etot : c->d_elem_EAI.p;
    if (fused) {  // K1 fused into K4: no element-block array in HBM
      OB_CUDA(c, launch_assemble_fused(d, eai, c->stream));
      launches += 1;
    } else {
      OB_CUDA(c, launch_elem_stiff(d, eai, c->stream));
      OB_CUDA(c, launch_assemble(d, c->stream));
      launches += 2;
    }
  }
  for (int64_t cs = 0; cs < L; cs++) {
    SparseDev d = sparse_dev(c, 0);
    const double* nl = c->has_case_nl ? c->d_case_nl.p + size_t(cs) * c->hb.NLtot() : c->d_nload_val.p;
    const double* q = c->has_case_q ? c->d_case_q.p + size_t(cs) * c->hb.ELtot() : c->d_eload_q.p;
    const double* eai_cs = c->has_case_eai ? c->d_case_eai.p + size_t(cs) * 3 * Etot : c->d_elem_EAI.p;
    OB_CUDA(c, launch_load_vector(d, nl, q, eai_cs, c->d_Ff.p + size_t(cs) * nf, c->stream));
    launches += d.ELtot ? 2 : 1;
  }
  OB_CUDA(c, cudaEventRecord(c->ev[7], c->stream));
  c->t_have[3] = true;
  c->launches = launches;
  c->assembled = true;
  return OBGPU_OK;
}

int solve_sparse(obgpu_ctx* c, const obgpu_solve_opts& o) {
  if (o.method != OBGPU_PCG_JACOBI)
    return fail(c, OBGPU_EUNSUPPORTED, "only OBGPU_PCG_JACOBI is implemented on the sparse route so far");
  int rc = assemble_general(c);
  if (rc != OBGPU_OK) return rc;
  const Symbolic& sy = c->sym;
  const int64_t L = c->L, S = c->hb.S, nf = sy.nf_tot(), nb = sy.nb_tot(), Etot = c->hb.Etot();
  double launches = c->launches;
  OB_CUDA(c, c->d_Uf.reserve(size_t(L) * std::max<int64_t>(nf, 1)));
  OB_CUDA(c, c->d_Fb.reserve(size_t(L) * std::max<int64_t>(nb, 1)));
  OB_CUDA(c, c->d_info.reserve(size_t(L) * S));
  OB_CUDA(c, c->d_fail.reserve(1));
  if (o.with_stress) OB_CUDA(c, c->d_stress.reserve(size_t(L) * Etot * 12));
  // PCG work vectors in one slab (p | ap | r | z | minv).  The matrix (360 MB at C4) streams through the 126 MB L2
  // once per iteration and would evict them; an access-policy window marks the slab persisting (everything else on
  // the stream is treated as streaming), so the three vector kernels and the gathers of the SpMV run out of L2.
  const size_t nfp = (size_t(std::max<int64_t>(nf, 1)) + 31) & ~size_t(31);
  OB_CUDA(c, c->d_pcgvec.reserve(5 * nfp));
  double* const v_p = c->d_pcgvec.p;
  double* const v_ap = v_p + nfp;
  double* const v_r = v_ap + nfp;
  double* const v_z = v_r + nfp;
  double* const v_minv = v_z + nfp;
  bool l2_window = false;
  cudaStreamAttrValue l2_av;
  std::memset(&l2_av, 0, sizeof(l2_av));
  {
    static const bool want = !(getenv("OBGPU_L2_PERSIST") && atoi(getenv("OBGPU_L2_PERSIST")) == 0);
    int max_persist = 0, max_window = 0;
    cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, c->device);
    cudaDeviceGetAttribute(&max_window, cudaDevAttrMaxAccessPolicyWindowSize, c->device);
    const size_t slab = 5 * nfp * sizeof(double);
    if (want && max_persist > 0 && max_window > 0 && slab > (8u << 20)) {
      const size_t win = std::min(slab, size_t(max_window));
      const size_t carve = std::min(win, size_t(max_persist));
      if (cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, carve) == cudaSuccess) {
        cudaStreamAttrValue av;
        std::memset(&av, 0, sizeof(av));
        av.accessPolicyWindow.base_ptr = v_p;
        av.accessPolicyWindow.num_bytes = win;
        av.accessPolicyWindow.hitRatio = float(double(carve) / double(win));
        av.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
        av.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
        l2_window = cudaStreamSetAttribute(c->stream, cudaStreamAttributeAccessPolicyWindow, &av) == cudaSuccess;
        l2_av = av;
      }
      cudaGetLastError();  // the window is an optimisation: never fail the solve over it
    }
  }
  int NC = 1;  // structures in flight
  if (S > 1) {
    NC = int(std::min<int64_t>(S, obgpu_ctx::NPCG));
    if (const char* ev = getenv("OBGPU_PCG_CONCURRENT")) NC = std::max(1, std::min(NC, atoi(ev)));
  }
  std::vector<cudaStream_t> pst(size_t(NC), c->stream);
  if (NC > 1) {
    if (!c->pcg_ev) OB_CUDA(c, cudaEventCreateWithFlags(&c->pcg_ev, cudaEventDisableTiming));
    for (int k = 0; k < NC; k++) {
      if (!c->pcg_stream[k]) OB_CUDA(c, cudaStreamCreateWithFlags(&c->pcg_stream[k], cudaStreamNonBlocking));
      pst[size_t(k)] = c->pcg_stream[k];
    }
  }
  if (l2_window)
    for (int k = 0; k < NC && NC > 1; k++) cudaStreamSetAttribute(pst[size_t(k)], cudaStreamAttributeAccessPolicyWindow, &l2_av);
  const size_t part_stride = size_t(2) * c->sm_count * 8;
  OB_CUDA(c, c->d_scal.reserve(size_t(16) * obgpu_ctx::NPCG));
  OB_CUDA(c, c->d_partial.reserve(part_stride * obgpu_ctx::NPCG));
  OB_CUDA(c, c->d_counter.reserve(obgpu_ctx::NPCG));
  if (!c->h_scal) OB_CUDA(c, cudaHostAlloc((void**)&c->h_scal, 8 * obgpu_ctx::NPCG * sizeof(double), cudaHostAllocDefault));
  OB_CUDA(c, cudaMemsetAsync(c->d_counter.p, 0, obgpu_ctx::NPCG * sizeof(unsigned int), c->stream));
  OB_CUDA(c, cudaMemsetAsync(c->d_fail.p, 0, sizeof(int32_t), c->stream));
  OB_CUDA(c, cudaMemsetAsync(c->d_info.p, 0, size_t(L) * S * sizeof(int32_t), c->stream));
  OB_CUDA(c, cudaEventRecord(c->ev[2], c->stream));
  const double rtol = o.rtol > 0 ? o.rtol : 1e-12;
  static const bool use_graph = !(getenv("OBGPU_PCG_GRAPH") && atoi(getenv("OBGPU_PCG_GRAPH")) == 0);
  const int chk = use_graph ? 24 : 25;  // iterations between convergence checks (even when replayed as a graph)
  double iters_total = 0, worst_rel = 0;
  bool all_converged = true;
  std::vector<int32_t> h_info(size_t(L) * S, 0);
  for (int64_t cs = 0; cs < L; cs++) {
    const int64_t k = c->has_case_eai ? cs : 0;
    SparseDev d = sparse_dev(c, k);
    if (cs == 0 || c->has_case_eai) {  // K5: BC elimination as gathers, Jacobi diagonal
      OB_CUDA(c, launch_bc_gather(d, c->stream));
      OB_CUDA(c, launch_inv_diag(nf, d.ff_diag, d.ff_val, v_minv, c->stream));
      launches += 3;
    }
    const double* nl = c->has_case_nl ? c->d_case_nl.p + size_t(cs) * c->hb.NLtot() : c->d_nload_val.p;
    const double* q = c->has_case_q ? c->d_case_q.p + size_t(cs) * c->hb.ELtot() : c->d_eload_q.p;
    const double* eai = c->has_case_eai ? c->d_case_eai.p + size_t(cs) * 3 * Etot : c->d_elem_EAI.p;
    if (L > 1) {  // Eq (equivalent loads on bounded DOFs) of THIS case; F_f was filled by assemble_general
      OB_CUDA(c, launch_load_vector(d, nl, q, eai, c->d_Ff.p + size_t(cs) * nf, c->stream));
      launches += 2;
    }
    // ---- one PCG per structure; up to NC of them in flight on their own streams, so that the launch gaps and the
    // 25-iteration convergence checks of one structure are covered by the kernels of the others (C5: 8 structures of
    // 0.5 M DOF per GPU were launch-latency bound one after the other) ----
    struct Job {
      int64_t s = -1; PcgArgs a; int it = 0, max_iter = 0; bool busy = false;
      cudaGraph_t graph = nullptr; cudaGraphExec_t exec = nullptr;  // `chk` iterations of THIS structure as one launch
    };
    auto drop_graph = [](Job& jb) {
      if (jb.exec) cudaGraphExecDestroy(jb.exec);
      if (jb.graph) cudaGraphDestroy(jb.graph);
      jb.exec = nullptr; jb.graph = nullptr;
    };
    std::vector<Job> job;
    job.resize(size_t(NC));
    int64_t next_s = 0;
    auto start_next = [&](int k) -> cudaError_t {
      Job& jb = job[size_t(k)];
      jb.busy = false;
      while (next_s < S && sy.free_off[next_s + 1] == sy.free_off[next_s]) next_s++;
      if (next_s >= S) return cudaSuccess;
      const int64_t st_i = next_s++;
      PcgArgs a;
      a.r0 = int32_t(sy.free_off[st_i]); a.r1 = int32_t(sy.free_off[st_i + 1]);
      a.slptr = d.sl_ptr; a.col = d.ff_col; a.val = d.ff_val; a.minv = v_minv;
      a.b = c->d_Ff.p + size_t(cs) * nf;
      a.x = c->d_Uf.p + size_t(cs) * nf;
      a.r = v_r; a.z = v_z; a.p = v_p; a.ap = v_ap;
      a.ctas_per_sm = NC > 1 ? 2 : 0;  // measured at C5's share (8 x 0.5 M DOF): 8 -> 1.84 s, 4 -> 1.67 s, 2 -> 1.62 s
      a.scal = c->d_scal.p + 16 * k; a.partial = c->d_partial.p + size_t(k) * part_stride; a.counter = c->d_counter.p + k;
      jb.s = st_i; jb.a = a; jb.it = 0; jb.busy = true;
      const int nrows = a.r1 - a.r0;
      jb.max_iter = o.max_iter > 0 ? o.max_iter : int(20.0 * std::sqrt(double(nrows))) + 1000;
      cudaError_t e = launch_pcg_init(a, c->sm_count, pst[size_t(k)]);
      launches += 1;
      if (e != cudaSuccess) return e;
      return cudaMemcpyAsync(c->h_scal + 8 * k, a.scal, 8 * sizeof(double), cudaMemcpyDeviceToHost, pst[size_t(k)]);
    };
    if (NC > 1) {
      OB_CUDA(c, cudaEventRecord(c->pcg_ev, c->stream));
      for (int k = 0; k < NC; k++) OB_CUDA(c, cudaStreamWaitEvent(pst[size_t(k)], c->pcg_ev, 0));
    }
    for (int k = 0; k < NC; k++) OB_CUDA(c, start_next(k));
    for (bool any = true; any;) {
      any = false;
      for (int k = 0; k < NC; k++) {
        Job& jb = job[size_t(k)];
        if (!jb.busy) continue;
        any = true;
        OB_CUDA(c, cudaStreamSynchronize(pst[size_t(k)]));
        double rr = c->h_scal[8 * k + 3];
        const double bb = c->h_scal[8 * k + 4];
        if (jb.it == 0) rr = bb;
        bool conv = false, stop = false;
        if (!(rr == rr)) stop = true;                              // NaN: breakdown (K_ff not positive definite)
        else if (rr <= rtol * rtol * bb) { conv = true; stop = true; }
        else if (jb.it >= jb.max_iter) stop = true;
        if (!stop) {
          const int n_it = std::min(chk, jb.max_iter - jb.it);
          if (use_graph && n_it == chk && (jb.it & 1) == 0) {
            // the 3*chk kernels of a check interval replayed as one graph launch: the host no longer has to feed
            // ~13 us per kernel launch to a loop whose kernels take 5-20 us each (chk is even: same rz ping-pong parity)
            if (!jb.exec) {
              OB_CUDA(c, cudaStreamBeginCapture(pst[size_t(k)], cudaStreamCaptureModeRelaxed));
              cudaError_t ce = cudaSuccess;
              for (int q2 = 0; q2 < chk && ce == cudaSuccess; q2++) ce = launch_pcg_iter(jb.a, q2, c->sm_count, pst[size_t(k)]);
              const cudaError_t ee = cudaStreamEndCapture(pst[size_t(k)], &jb.graph);
              OB_CUDA(c, ce);
              OB_CUDA(c, ee);
              OB_CUDA(c, cudaGraphInstantiate(&jb.exec, jb.graph, 0));
            }
            OB_CUDA(c, cudaGraphLaunch(jb.exec, pst[size_t(k)]));
            jb.it += chk;
            launches += 3.0 * n_it;
          } else {
            for (int q2 = 0; q2 < n_it; q2++, jb.it++) OB_CUDA(c, launch_pcg_iter(jb.a, jb.it, c->sm_count, pst[size_t(k)]));
            launches += 3.0 * n_it;
          }
          OB_CUDA(c, cudaMemcpyAsync(c->h_scal + 8 * k, jb.a.scal, 8 * sizeof(double), cudaMemcpyDeviceToHost, pst[size_t(k)]));
          continue;
        }
        iters_total += jb.it;
        const double rel = bb > 0 ? std::sqrt(rr / bb) : 0.0;
        worst_rel = std::max(worst_rel, rel == rel ? rel : 1.0);
        if (!conv) {
          all_converged = false;
          h_info[size_t(cs) * S + jb.s] = -1;
        }
        drop_graph(jb);
        OB_CUDA(c, start_next(k));
      }
    }
    if (NC > 1)
      for (int k = 0; k < NC; k++) {
        OB_CUDA(c, cudaEventRecord(c->pcg_ev, pst[size_t(k)]));
        OB_CUDA(c, cudaStreamWaitEvent(c->stream, c->pcg_ev, 0));
      }
    // reactions F_b = K_bf U_f - Eq, element end forces
    OB_CUDA(c, launch_reactions(d, c->d_Uf.p + size_t(cs) * nf, c->d_Fb.p + size_t(cs) * nb, c->stream));
    launches += 1;
    if (o.with_stress) {
      OB_CUDA(c, launch_stress(d, eai, q, c->d_Uf.p + size_t(cs) * nf, c->d_stress.p + size_t(cs) * Etot * 12, c->stream));
      launches += 1;
    }
  }
  if (!all_converged)
    OB_CUDA(c, cudaMemcpyAsync(c->d_info.p, h_info.data(), h_info.size() * sizeof(int32_t), cudaMemcpyHostToDevice, c->stream));
  OB_CUDA(c, cudaEventRecord(c->ev[3], c->stream));
  OB_CUDA(c, cudaStreamSynchronize(c->stream));
  if (l2_window) {
    cudaStreamAttrValue av;
    std::memset(&av, 0, sizeof(av));
    cudaStreamSetAttribute(c->stream, cudaStreamAttributeAccessPolicyWindow, &av);
    for (int k = 0; k < NC && NC > 1; k++) cudaStreamSetAttribute(pst[size_t(k)], cudaStreamAttributeAccessPolicyWindow, &av);
    cudaCtxResetPersistingL2Cache();
    cudaGetLastError();
  }
  c->t_have[1] = true;
  c->launches = launches;
  c->pcg_iters = iters_total;
  c->pcg_relres = worst_rel;
  c->solved = true;
  c->have_ff = true;
  c->kernel_info = "k_spmv_dot (SELL-32 SpMV of the Jacobi-PCG iteration)";
  c->have_stress = o.with_stress != 0;
  if (!all_converged) return fail(c, OBGPU_ENOTCONVERGED, "PCG did not converge (or broke down: K_ff not positive definite?)");
  return OBGPU_OK;
}

int download_general(obgpu_ctx* c, int which, void* out, bool& handled) {
  handled = false;
  if (which != OBGPU_RES_KVAL && which != OBGPU_RES_ELEM_K && !(which == OBGPU_RES_FF && c->assembled && !c->solved))
    return OBGPU_OK;
  handled = true;
  if (!c->assembled) return fail(c, OBGPU_ESEQUENCE, "obgpu_download(K) before obgpu_assemble");
  const void* src;
  size_t bytes;
  if (which == OBGPU_RES_KVAL) { src = c->d_kval.p; bytes = size_t(c->LK) * c->sym.nnz_tot() * 8; }
  else if (which == OBGPU_RES_ELEM_K) {
    if (!c->elem_k_valid) {  // the fused assembly never materialises the element blocks: produce them on demand
      OB_CUDA(c, c->d_elem_k.reserve(size_t(c->LK) * c->hb.Etot() * ELEM_K));
      for (int64_t k = 0; k < c->LK; k++) {
        SparseDev d = sparse_dev(c, k);
        const double* eai = c->has_case_eai ? c->d_case_eai.p + size_t(k) * 3 * c->hb.Etot() : c->d_elem_EAI.p;
        OB_CUDA(c, launch_elem_stiff(d, eai, c->stream));
      }
      c->elem_k_valid = true;
    }
    src = c->d_elem_k.p; bytes = size_t(c->LK) * c->hb.Etot() * ELEM_K * 8;
  }
  else { src = c->d_Ff.p; bytes = size_t(c->L) * c->sym.nf_tot() * 8; }
  if (bytes) OB_CUDA(c, cudaMemcpyAsync(out, src, bytes, cudaMemcpyDeviceToHost, c->stream));
  OB_CUDA(c, cudaStreamSynchronize(c->stream));
  return OBGPU_OK;
}

}  // namespace obgpu
