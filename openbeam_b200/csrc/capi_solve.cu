// capi_solve.cu — solveStatic / postProcCalcStress entry points, downloads, the host-buffer sweep.
#include <algorithm>
#include <cstring>

#include "ctx.h"

using namespace obgpu;

namespace obgpu {
int solve_general(obgpu_ctx* c, const obgpu_solve_opts& o);   // capi_general.cu
int assemble_general(obgpu_ctx* c);                            // capi_general.cu
int download_general(obgpu_ctx* c, int which, void* host_out, bool& handled);
int profile_kernel(obgpu_ctx* c, int which, int reps, double* ms_avg, double* alg_bytes);  // capi_sparse.cu
int result_count_general(obgpu_ctx* c, int which, int64_t* count);
int symbolic_general(obgpu_ctx* c, int which, int32_t* out, int64_t* count);
int materialize_host_batch(obgpu_ctx* c);
}

static int ensure_symbolic(obgpu_ctx* c) {
  if (!c->uploaded) return fail(c, OBGPU_ESEQUENCE, "no batch uploaded");
  if (!c->sym.valid) return obgpu_symbolic(c);
  return OBGPU_OK;
}

static int alloc_results(obgpu_ctx* c, int64_t L, bool stress) {
  const Symbolic& s = c->sym;
  OB_CUDA(c, c->d_Uf.reserve(size_t(L) * std::max<int64_t>(s.nf_tot(), 1)));
  OB_CUDA(c, c->d_Ff.reserve(size_t(L) * std::max<int64_t>(s.nf_tot(), 1)));
  OB_CUDA(c, c->d_Fb.reserve(size_t(L) * std::max<int64_t>(s.nb_tot(), 1)));
  OB_CUDA(c, c->d_info.reserve(size_t(L) * c->hb.S));
  OB_CUDA(c, c->d_fail.reserve(1));
  if (stress) OB_CUDA(c, c->d_stress.reserve(size_t(L) * c->hb.Etot() * 12));
  return OBGPU_OK;
}

SummaryArgs obgpu::summary_args(obgpu_ctx* c, double* d_out) {
  SummaryArgs a;
  a.L = c->L; a.S = c->hb.S; a.nf_tot = c->sym.nf_tot(); a.nb_tot = c->sym.nb_tot();
  a.free_off = c->d_free_off.p; a.bnd_off = c->d_bnd_off.p; a.bnd_dofid = c->d_bnd_dofid.p;
  a.U_f = c->d_Uf.p; a.F_b = c->d_Fb.p; a.info = c->d_info.p; a.out = d_out;
  return a;
}

static TinyArgs tiny_args(obgpu_ctx* c, int64_t L, const double* nl, const double* q, const double* eai, double* Uf,
                          double* Fb, double* stress, double* Ff, int32_t* info) {
  TinyArgs a;
  a.prog_i = c->prog.ints.data();
  a.prog_d = c->prog.dbls.data();
  a.n_ints = int(c->prog.ints.size());
  a.n_dbls = int(c->prog.dbls.size());
  a.n_cases = L;
  a.case_nl = nl; a.case_q = q; a.case_eai = eai;
  a.U_f = Uf; a.F_b = Fb; a.stress = stress; a.F_f = Ff; a.info = info;
  a.fail = c->d_fail.p;
  a.summ = nullptr; a.bnd_dofid = nullptr; a.summ_fallback = nullptr; a.n_summ_peer = 0;
  a.stress_compact = 0;
  for (double*& p : a.summ_peer) p = nullptr;
  return a;
}

extern "C" {

int obgpu_solve_static(obgpu_ctx* c, const obgpu_solve_opts* opts) {
  if (!c) return OBGPU_EINVAL;
  OB_NEED_DEVICE(c);
  int rc = ensure_symbolic(c);
  if (rc != OBGPU_OK) return rc;
  obgpu_solve_opts o;
  o.method = OBGPU_BATCHED_LDLT; o.with_stress = 1; o.rtol = 1e-12; o.max_iter = 0; o.flags = 0;
  if (opts) o = *opts;
  c->summ_written = nullptr;
  c->last_method = o.method; c->last_rtol = o.rtol; c->last_max_iter = o.max_iter;
  const bool want_compact = o.with_stress && (o.flags & OBGPU_SOLVE_COMPACT_STRESS);
  c->have_stress_c = false;
  if (o.method == OBGPU_BATCHED_LDLT && c->prog.valid && !c->have_xy_len) {
    rc = alloc_results(c, c->L, o.with_stress != 0 && !want_compact);
    if (rc != OBGPU_OK) return rc;
    if (want_compact) OB_CUDA(c, c->d_stress_c.reserve(size_t(c->L) * c->hb.Etot() * 6));
    OB_CUDA(c, cudaMemsetAsync(c->d_fail.p, 0, sizeof(int32_t), c->stream));
    TinyArgs a = tiny_args(c, c->L, c->has_case_nl ? c->d_case_nl.p : nullptr, c->has_case_q ? c->d_case_q.p : nullptr,
                           c->has_case_eai ? c->d_case_eai.p : nullptr, c->d_Uf.p, c->d_Fb.p,
                           !o.with_stress ? nullptr : want_compact ? c->d_stress_c.p : c->d_stress.p,
                           (o.flags & OBGPU_SOLVE_SKIP_FF) ? nullptr : c->d_Ff.p, c->d_info.p);
    a.stress_compact = want_compact ? 1 : 0;
    c->have_ff = !(o.flags & OBGPU_SOLVE_SKIP_FF);
    SummaryArgs sa;
    if (c->d_summ) {  // bound summaries: written by the solve kernel itself (or by k_summaries right behind it)
      sa = summary_args(c, c->d_summ);
      a.summ = c->d_summ; a.bnd_dofid = c->h_bnd_dofid.data(); a.summ_fallback = &sa;
      a.n_summ_peer = c->n_summ_peer;
      for (int p = 0; p < c->n_summ_peer; p++) a.summ_peer[p] = c->summ_peer[p];
    }
    OB_CUDA(c, cudaEventRecord(c->ev[2], c->stream));
    cudaError_t le = launch_tiny_fused(c->prog.nf, c->prog.n, a, c->sm_count, c->stream, &c->kernel_info);
    bool compact_pass = false;
    if (le == cudaErrorNotYetImplemented) {  // this kernel only writes the six-component record: compact it afterwards
      cudaGetLastError();
      OB_CUDA(c, c->d_stress.reserve(size_t(c->L) * c->hb.Etot() * 12));
      a.stress = c->d_stress.p;
      a.stress_compact = 0;
      le = launch_tiny_fused(c->prog.nf, c->prog.n, a, c->sm_count, c->stream, &c->kernel_info);
      compact_pass = le == cudaSuccess;
    }
    if (le == cudaErrorNotSupported && c->prog.has_load_kinds) {  // no NVRTC here and non-uniform loads: general route
      c->summ_written = nullptr;
      cudaGetLastError();
      goto general_route;   // (not a plain return: the compact-record pass below must follow this solve too)
    }
    OB_CUDA(c, le);
    if (compact_pass) OB_CUDA(c, launch_compact_stress(c->d_stress.p, c->d_stress_c.p, c->L * c->hb.Etot(), c->stream));
    OB_CUDA(c, cudaEventRecord(c->ev[3], c->stream));
    c->t_have[1] = true;
    c->summ_written = c->d_summ;
    c->launches = ((c->d_summ && std::strncmp(c->kernel_info, "k_tiny_jit", 10) != 0) ? 2 + c->n_summ_peer : 1) + (compact_pass ? 1 : 0);
    c->pcg_iters = 0;
    c->solved = true;
    c->have_stress = o.with_stress != 0 && (!want_compact || compact_pass);
    c->have_stress_c = want_compact;
    return OBGPU_OK;
  }
general_route:
  rc = solve_general(c, o);
  if ((rc == OBGPU_OK || rc == OBGPU_ESINGULAR || rc == OBGPU_ENOTCONVERGED) && want_compact && c->have_stress) {
    // the warp-per-structure and sparse routes write the six-component record: one more pass compacts it
    OB_CUDA(c, c->d_stress_c.reserve(size_t(c->L) * c->hb.Etot() * 6));
    OB_CUDA(c, launch_compact_stress(c->d_stress.p, c->d_stress_c.p, c->L * c->hb.Etot(), c->stream));
    c->have_stress_c = true;
    c->launches += 1;
  }
  return rc;
}

const char* obgpu_kernel_info(const obgpu_ctx* c) { return c ? c->kernel_info : ""; }

int obgpu_bind_summary_peers(obgpu_ctx* c, int n, double* const* bufs) {
  if (!c || n < 0 || n > 8 || (n > 0 && !bufs)) return OBGPU_EINVAL;
  for (int i = 0; i < n; i++)
    if (!bufs[i] || reinterpret_cast<uintptr_t>(bufs[i]) % 32) return fail(c, OBGPU_EINVAL, "peer summary buffers must be non-null and 32-byte aligned");
  if (n > 0 && !c->d_summ) return fail(c, OBGPU_ESEQUENCE, "obgpu_bind_summary_peers needs a bound local buffer (obgpu_bind_summaries)");
  c->n_summ_peer = n;
  for (int i = 0; i < 8; i++) c->summ_peer[i] = i < n ? bufs[i] : nullptr;
  return OBGPU_OK;
}

int obgpu_bind_summaries(obgpu_ctx* c, double* d_out) {
  if (!c) return OBGPU_EINVAL;
  if (reinterpret_cast<uintptr_t>(d_out) % 32) return fail(c, OBGPU_EINVAL, "summary buffer must be 32-byte aligned");
  c->d_summ = d_out;
  c->summ_written = nullptr;
  if (!d_out) c->n_summ_peer = 0;
  return OBGPU_OK;
}

int obgpu_assemble(obgpu_ctx* c) {
  if (!c) return OBGPU_EINVAL;
  OB_NEED_DEVICE(c);
  int rc = ensure_symbolic(c);
  if (rc != OBGPU_OK) return rc;
  return assemble_general(c);
}

int obgpu_stress(obgpu_ctx* c) {
  if (!c) return OBGPU_EINVAL;
  if (!c->solved) return fail(c, OBGPU_ESEQUENCE, "obgpu_stress before obgpu_solve_static");
  if (c->have_stress || c->have_stress_c) return OBGPU_OK;
  obgpu_solve_opts o;
  // the same route as the solve it follows (a PCG solve must not be replaced by an LDL^T one, nor refused because the
  // structure is too large for the batched route); the fused kernels recompute, the sparse route re-solves
  o.method = c->last_method; o.with_stress = 1; o.rtol = c->last_rtol; o.max_iter = c->last_max_iter; o.flags = 0;
  return obgpu_solve_static(c, &o);
}

int obgpu_device_ptr(obgpu_ctx* c, int which, void** p) {
  if (!c || !p) return OBGPU_EINVAL;
  OB_NEED_DEVICE(c);
  switch (which) {
    case OBGPU_RES_UF: *p = c->d_Uf.p; break;
    case OBGPU_RES_FB: *p = c->d_Fb.p; break;
    case OBGPU_RES_STRESS: *p = c->d_stress.p; break;
    case OBGPU_RES_STRESS_COMPACT: *p = c->d_stress_c.p; break;
    case OBGPU_RES_INFO: *p = c->d_info.p; break;
    case OBGPU_RES_FF:
      if (!c->have_ff) return fail(c, OBGPU_ESEQUENCE, "F_f was not kept (OBGPU_SOLVE_SKIP_FF)");
      *p = c->d_Ff.p; break;
    case OBGPU_RES_UB: *p = c->d_Ub.p; break;
    default: return OBGPU_EINVAL;
  }
  return OBGPU_OK;
}

int obgpu_download(obgpu_ctx* c, int which, void* out) {
  if (!c || !out) return OBGPU_EINVAL;
  OB_NEED_DEVICE(c);
  bool handled = false;
  int rc = download_general(c, which, out, handled);
  if (handled) return rc;
  if (which == OBGPU_RES_UB && c->sym.valid) {   // BuildProblemInfo::U_b is a product of the symbolic phase
    const size_t bytes = size_t(c->sym.nb_tot()) * 8;
    if (bytes) OB_CUDA(c, cudaMemcpyAsync(out, c->d_Ub.p, bytes, cudaMemcpyDeviceToHost, c->stream));
    OB_CUDA(c, cudaStreamSynchronize(c->stream));
    return OBGPU_OK;
  }
  if (!c->solved) return fail(c, OBGPU_ESEQUENCE, "obgpu_download before obgpu_solve_static");
  const Symbolic& s = c->sym;
  const void* src = nullptr;
  size_t bytes = 0;
  switch (which) {
    case OBGPU_RES_UF: src = c->d_Uf.p; bytes = size_t(c->L) * s.nf_tot() * 8; break;
    case OBGPU_RES_FF:
      if (!c->have_ff) return fail(c, OBGPU_ESEQUENCE, "F_f was not kept (OBGPU_SOLVE_SKIP_FF)");
      src = c->d_Ff.p; bytes = size_t(c->L) * s.nf_tot() * 8; break;
    case OBGPU_RES_FB: src = c->d_Fb.p; bytes = size_t(c->L) * s.nb_tot() * 8; break;
    case OBGPU_RES_UB: src = c->d_Ub.p; bytes = size_t(s.nb_tot()) * 8; break;
    case OBGPU_RES_STRESS:
      if (!c->have_stress) return fail(c, OBGPU_ESEQUENCE, "six-component stress was not computed (with_stress = 0, or OBGPU_SOLVE_COMPACT_STRESS: read OBGPU_RES_STRESS_COMPACT)");
      src = c->d_stress.p; bytes = size_t(c->L) * c->hb.Etot() * 12 * 8; break;
    case OBGPU_RES_STRESS_COMPACT:
      if (!c->have_stress_c) return fail(c, OBGPU_ESEQUENCE, "no compact stress records (solve with OBGPU_SOLVE_COMPACT_STRESS)");
      src = c->d_stress_c.p; bytes = size_t(c->L) * c->hb.Etot() * 6 * 8; break;
    case OBGPU_RES_INFO: src = c->d_info.p; bytes = size_t(c->L) * c->hb.S * 4; break;
    default: return fail(c, OBGPU_EINVAL, "unknown result id");
  }
  OB_CUDA(c, cudaEventRecord(c->ev[4], c->stream));
  if (bytes) OB_CUDA(c, cudaMemcpyAsync(out, src, bytes, cudaMemcpyDeviceToHost, c->stream));
  OB_CUDA(c, cudaEventRecord(c->ev[5], c->stream));
  c->t_have[2] = true;
  OB_CUDA(c, cudaStreamSynchronize(c->stream));
  if (which == OBGPU_RES_INFO) {
    const int32_t* info = static_cast<const int32_t*>(out);
    for (int64_t i = 0; i < c->L * c->hb.S; i++)
      if (info[i] != 0) return fail(c, OBGPU_ESINGULAR, obgpu_status_string(OBGPU_ESINGULAR));
  }
  return OBGPU_OK;
}

int obgpu_get_timings(const obgpu_ctx* cc, double out[8]) {
  if (!cc || !out) return OBGPU_EINVAL;
  obgpu_ctx* c = const_cast<obgpu_ctx*>(cc);
  for (int i = 0; i < 8; i++) out[i] = 0;
  out[0] = c->t_symbolic_ms;
  float ms = 0;
  if (c->t_have[0] && cudaEventSynchronize(c->ev[1]) == cudaSuccess && cudaEventElapsedTime(&ms, c->ev[0], c->ev[1]) == cudaSuccess) out[4] = ms;
  if (c->t_have[1] && cudaEventSynchronize(c->ev[3]) == cudaSuccess && cudaEventElapsedTime(&ms, c->ev[2], c->ev[3]) == cudaSuccess) out[2] = ms;
  if (c->t_have[2] && cudaEventSynchronize(c->ev[5]) == cudaSuccess && cudaEventElapsedTime(&ms, c->ev[4], c->ev[5]) == cudaSuccess) out[5] = ms;
  if (c->t_have[3] && cudaEventSynchronize(c->ev[7]) == cudaSuccess && cudaEventElapsedTime(&ms, c->ev[6], c->ev[7]) == cudaSuccess) out[1] = ms;
  out[6] = c->launches;
  out[7] = c->pcg_iters;
  cudaGetLastError();
  return OBGPU_OK;
}

int obgpu_profile_kernel(obgpu_ctx* c, int which, int reps, double* ms_avg, double* alg_bytes) {
  if (!c) return OBGPU_EINVAL;
  OB_NEED_DEVICE(c);
  return profile_kernel(c, which, reps, ms_avg, alg_bytes);
}

int obgpu_set_stress_length_poses(obgpu_ctx* c, const double* node_xy) {
  if (!c) return OBGPU_EINVAL;
  OB_NEED_DEVICE(c);
  if (!c->uploaded) return fail(c, OBGPU_ESEQUENCE, "obgpu_set_stress_length_poses before obgpu_batch_upload");
  c->have_xy_len = false;
  if (!node_xy) return OBGPU_OK;
  OB_CUDA(c, c->d_node_xy_len.upload(node_xy, size_t(2 * c->hb.Ntot()), c->stream));
  OB_CUDA(c, cudaStreamSynchronize(c->stream));
  c->have_xy_len = true;
  return OBGPU_OK;
}

int obgpu_result_count(obgpu_ctx* c, int which, int64_t* count) {
  if (!c || !count || !c->sym.valid) return OBGPU_ESEQUENCE;
  const Symbolic& s = c->sym;
  const int64_t L = c->L, LK = c->has_case_eai ? c->L : 1, E = c->hb.Etot();
  switch (which) {
    case OBGPU_RES_UF: case OBGPU_RES_FF: *count = L * s.nf_tot(); return OBGPU_OK;
    case OBGPU_RES_FB: *count = L * s.nb_tot(); return OBGPU_OK;
    case OBGPU_RES_UB: *count = s.nb_tot(); return OBGPU_OK;
    case OBGPU_RES_STRESS: *count = L * E * 12; return OBGPU_OK;
    case OBGPU_RES_STRESS_COMPACT: *count = L * E * 6; return OBGPU_OK;
    case OBGPU_RES_INFO: *count = L * c->hb.S; return OBGPU_OK;
    case OBGPU_RES_KVAL: *count = LK * s.nnz_tot(); return OBGPU_OK;
    case OBGPU_RES_ELEM_K: *count = LK * E * ELEM_K; return OBGPU_OK;
  }
  if (c->device_batch_mode && !c->host_batch_valid) {
    int rc = materialize_host_batch(c);
    if (rc == OBGPU_OK) rc = obgpu_symbolic(c);
    if (rc != OBGPU_OK) return rc;
  }
  return result_count_general(c, which, count);
}

int obgpu_symbolic_count(obgpu_ctx* c, int which, int64_t* count) {
  if (!c || !count || !c->sym.valid) return OBGPU_ESEQUENCE;
  const Symbolic& s = c->sym;
  switch (which) {
    case OBGPU_SYM_DOF_NODE: case OBGPU_SYM_DOF_ID: *count = s.n_tot(); return OBGPU_OK;
    case OBGPU_SYM_FREE: *count = s.nf_tot(); return OBGPU_OK;
    case OBGPU_SYM_BOUNDED: *count = s.nb_tot(); return OBGPU_OK;
    case OBGPU_SYM_DOF_INDEX: *count = 6 * c->hb.Ntot(); return OBGPU_OK;
    case OBGPU_SYM_K_ROWPTR: *count = s.n_tot() + c->hb.S; return OBGPU_OK;
    case OBGPU_SYM_K_COL: *count = s.nnz_tot(); return OBGPU_OK;
    case OBGPU_SYM_TRI_NNZ: *count = s.T_tot(); return OBGPU_OK;
    case OBGPU_SYM_KFF_ROWPTR: case OBGPU_SYM_KFF_COL: return symbolic_general(c, which, nullptr, count);
  }
  return OBGPU_EINVAL;
}

int obgpu_pcg_info(const obgpu_ctx* c, double out[4]) {
  if (!c || !out) return OBGPU_EINVAL;
  out[0] = c->pcg_iters;
  out[1] = c->pcg_relres;
  out[2] = c->pcg_restarts;
  out[3] = c->pcg_levels;
  return OBGPU_OK;
}

// Host-buffer sweep: chunks of cases stream H2D -> fused kernel -> D2H over NSLOT streams so the PCIe
// copies in both directions overlap each other and the compute.
int obgpu_solve_cases_host(obgpu_ctx* c, int64_t L, const double* nl, const double* q, const double* eai, double* U_f,
                           double* F_b, double* stress, int32_t* info) {
  return obgpu_solve_cases_host_flags(c, L, nl, q, eai, U_f, F_b, stress, info, 0);
}

int obgpu_solve_cases_host_flags(obgpu_ctx* c, int64_t L, const double* nl, const double* q, const double* eai, double* U_f,
                            double* F_b, double* stress, int32_t* info, int flags) {
  if (!c) return OBGPU_EINVAL;
  const bool compact = stress && (flags & OBGPU_SOLVE_COMPACT_STRESS);
  const int SR = compact ? 6 : 12;
  OB_NEED_DEVICE(c);
  int rc = ensure_symbolic(c);
  if (rc != OBGPU_OK) return rc;
  if (!c->prog.valid || c->hb.S != 1)
    return fail(c, OBGPU_EUNSUPPORTED, "obgpu_solve_cases_host needs one shared small topology (use update_cases + solve_static)");
  if (L < 1) return fail(c, OBGPU_EINVAL, "n_cases must be >= 1");
  const int64_t NL = c->hb.NLtot(), EL = c->hb.ELtot(), E = c->hb.Etot(), nf = c->sym.nf_tot(), nb = c->sym.nb_tot();
  if (NL == 0) nl = nullptr;
  if (EL == 0) q = nullptr;
  const int64_t CH = std::min<int64_t>(L, 1 << 17);
  OB_CUDA(c, c->d_fail.reserve(1));
  OB_CUDA(c, cudaMemsetAsync(c->d_fail.p, 0, sizeof(int32_t), c->stream));
  OB_CUDA(c, cudaEventRecord(c->ev[0], c->stream));
  for (int i = 0; i < obgpu_ctx::NSLOT; i++) {
    auto& sl = c->slot[i];
    if (nl) OB_CUDA(c, sl.nl.reserve(size_t(CH) * NL));
    if (q) OB_CUDA(c, sl.q.reserve(size_t(CH) * EL));
    if (eai) OB_CUDA(c, sl.eai.reserve(size_t(CH) * 3 * E));
    OB_CUDA(c, sl.Uf.reserve(size_t(CH) * std::max<int64_t>(nf, 1)));
    OB_CUDA(c, sl.Fb.reserve(size_t(CH) * std::max<int64_t>(nb, 1)));
    if (stress) OB_CUDA(c, sl.stress.reserve(size_t(CH) * E * SR));
    OB_CUDA(c, sl.info.reserve(size_t(CH)));
    OB_CUDA(c, cudaStreamWaitEvent(c->slot_stream[i], c->ev[0], 0));
  }
  int64_t launches = 0;
  for (int64_t c0 = 0, k = 0; c0 < L; c0 += CH, k++) {
    const int64_t n = std::min(CH, L - c0);
    const int si = int(k % obgpu_ctx::NSLOT);
    auto& sl = c->slot[si];
    cudaStream_t st = c->slot_stream[si];
    if (nl) OB_CUDA(c, cudaMemcpyAsync(sl.nl.p, nl + c0 * NL, size_t(n) * NL * 8, cudaMemcpyHostToDevice, st));
    if (q) OB_CUDA(c, cudaMemcpyAsync(sl.q.p, q + c0 * EL, size_t(n) * EL * 8, cudaMemcpyHostToDevice, st));
    if (eai) OB_CUDA(c, cudaMemcpyAsync(sl.eai.p, eai + c0 * 3 * E, size_t(n) * 3 * E * 8, cudaMemcpyHostToDevice, st));
    TinyArgs a = tiny_args(c, n, nl ? sl.nl.p : nullptr, q ? sl.q.p : nullptr, eai ? sl.eai.p : nullptr, sl.Uf.p, sl.Fb.p,
                           stress ? sl.stress.p : nullptr, nullptr, sl.info.p);
    a.stress_compact = compact ? 1 : 0;
    cudaError_t le = launch_tiny_fused(c->prog.nf, c->prog.n, a, c->sm_count, st, &c->kernel_info);
    if (le == cudaErrorNotYetImplemented) {  // six-component kernel + compaction pass
      cudaGetLastError();
      OB_CUDA(c, sl.stress_full.reserve(size_t(CH) * E * 12));
      a.stress = sl.stress_full.p;
      a.stress_compact = 0;
      le = launch_tiny_fused(c->prog.nf, c->prog.n, a, c->sm_count, st, &c->kernel_info);
      if (le == cudaSuccess) {
        le = launch_compact_stress(sl.stress_full.p, sl.stress.p, n * E, st);
        launches++;
      }
    }
    if (le == cudaErrorNotSupported)
      return fail(c, OBGPU_EUNSUPPORTED, "the host-buffer sweep with non-uniform element loads needs the run-time specialised kernel (NVRTC)");
    OB_CUDA(c, le);
    launches++;
    if (U_f && nf) OB_CUDA(c, cudaMemcpyAsync(U_f + c0 * nf, sl.Uf.p, size_t(n) * nf * 8, cudaMemcpyDeviceToHost, st));
    if (F_b && nb) OB_CUDA(c, cudaMemcpyAsync(F_b + c0 * nb, sl.Fb.p, size_t(n) * nb * 8, cudaMemcpyDeviceToHost, st));
    if (stress) OB_CUDA(c, cudaMemcpyAsync(stress + c0 * E * SR, sl.stress.p, size_t(n) * E * SR * 8, cudaMemcpyDeviceToHost, st));
    if (info) OB_CUDA(c, cudaMemcpyAsync(info + c0, sl.info.p, size_t(n) * 4, cudaMemcpyDeviceToHost, st));
  }
  int32_t failed = 0;
  for (int i = 0; i < obgpu_ctx::NSLOT; i++) OB_CUDA(c, cudaStreamSynchronize(c->slot_stream[i]));
  OB_CUDA(c, cudaMemcpyAsync(&failed, c->d_fail.p, 4, cudaMemcpyDeviceToHost, c->stream));
  OB_CUDA(c, cudaStreamSynchronize(c->stream));
  c->launches = double(launches);
  if (failed) return fail(c, OBGPU_ESINGULAR, obgpu_status_string(OBGPU_ESINGULAR));
  return OBGPU_OK;
}

}  // extern "C"
