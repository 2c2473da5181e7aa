// capi_sparse.cu — global K assembly (K1+K4), BC gather (K5) and the PCG route (K7-K9, K10) for large structures.
#include <algorithm>
#include <cmath>
#include <cstdio>
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
  std::vector<int32_t> node0(hb.Etot()), f0(hb.Etot()), b0(hb.Etot()), nf0(hb.Ntot()), eg(hb.ELtot());
  for (int64_t s = 0; s < hb.S; s++) {
    for (int64_t e = hb.elem_off[s]; e < hb.elem_off[s + 1]; e++) {
      node0[e] = int32_t(hb.node_off[s]);
      f0[e] = int32_t(c->sym.free_off[s]);
      b0[e] = int32_t(c->sym.bnd_off[s]);
    }
    for (int64_t n = hb.node_off[s]; n < hb.node_off[s + 1]; n++) nf0[n] = int32_t(c->sym.free_off[s]);
    for (int64_t l = hb.eload_off[s]; l < hb.eload_off[s + 1]; l++) eg[l] = int32_t(hb.elem_off[s] + hb.eload_elem[l]);
  }
  OB_CUDA(c, c->d_elem_node0.upload(node0, st));
  OB_CUDA(c, c->d_elem_f0.upload(f0, st));
  OB_CUDA(c, c->d_elem_b0.upload(b0, st));
  OB_CUDA(c, c->d_node_f0.upload(nf0, st));
  OB_CUDA(c, c->d_bb_rowptr.upload(P.bb_rowptr, st));
  OB_CUDA(c, c->d_bb_col.upload(P.bb_col, st));
  OB_CUDA(c, c->d_bb_src.upload(P.bb_src, st));
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
  if (P.ff_blk.valid) {
    OB_CUDA(c, c->d_ff_bslptr.upload(P.ff_blk.sl_ptr, st));
    OB_CUDA(c, c->d_ff_bcol.upload(P.ff_blk.col, st));
    OB_CUDA(c, c->d_ff_bmap.upload(P.ff_blk.map, st));
    OB_CUDA(c, c->d_ff_bval.reserve(P.ff_blk.map.size()));
  }
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

// multigrid plan (host, once per topology) and its device arrays
static int ensure_mg(obgpu_ctx* c) {
  if (c->dev_mg) return OBGPU_OK;
  std::string err;
  int rc = build_mg_plan(c->hb, c->sym, c->plan, c->mg, err);
  if (rc != OBGPU_OK) return fail(c, rc, err);
  const size_t nl = c->mg.lev.size();
  if (nl > size_t(MG_MAX_LEVELS)) return fail(c, OBGPU_EUNSUPPORTED, "multigrid hierarchy deeper than MG_MAX_LEVELS");
  c->mg_lev.clear();
  c->mg_lev.resize(nl);
  cudaStream_t st = c->stream;
  for (size_t l = 1; l < nl; l++) {
    const MgLevel& L = c->mg.lev[l];
    obgpu_ctx::MgLevelBufs& B = c->mg_lev[l];
    OB_CUDA(c, B.sl_ptr.upload(L.sl_ptr, st));
    OB_CUDA(c, B.col.upload(L.col, st));
    OB_CUDA(c, B.f_agg.upload(L.f_agg, st));
    OB_CUDA(c, B.f_kind.upload(L.f_kind, st));
    OB_CUDA(c, B.f_w.upload(L.f_w, st));
    OB_CUDA(c, B.mem_ptr.upload(L.mem_ptr, st));
    OB_CUDA(c, B.mem.upload(L.mem, st));
    OB_CUDA(c, B.val.reserve(std::max<size_t>(L.col.size(), 1)));
    B.nblk = 0;
    if (L.blk.valid) {
      OB_CUDA(c, B.bsl_ptr.upload(L.blk.sl_ptr, st));
      OB_CUDA(c, B.bcol.upload(L.blk.col, st));
      OB_CUDA(c, B.bmap.upload(L.blk.map, st));
      OB_CUDA(c, B.bval.reserve(L.blk.map.size()));
      B.nblk = int64_t(L.blk.col.size());
    }
    OB_CUDA(c, B.minv.reserve(size_t(L.n)));
    OB_CUDA(c, B.vec.reserve(size_t(5) * L.n));
    OB_CUDA(c, cudaMemsetAsync(B.vec.p, 0, size_t(5) * L.n * sizeof(double), st));
    OB_CUDA(c, cudaMemsetAsync(B.minv.p, 0, size_t(L.n) * sizeof(double), st));
  }
  OB_CUDA(c, c->d_mg_lam.reserve(MG_MAX_LEVELS));
  OB_CUDA(c, c->d_mg_cinv.reserve(size_t(obgpu_ctx::NPCG) * MgPlan::DENSE_MAX * MgPlan::DENSE_MAX));
  OB_CUDA(c, c->d_mg_prow.reserve(size_t(obgpu_ctx::NPCG) * 2 * (MgPlan::DENSE_MAX + 2)));
  OB_CUDA(c, cudaStreamSynchronize(st));
  c->dev_mg = true;
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
  d.elem_node0 = c->d_elem_node0.p; d.elem_f0 = c->d_elem_f0.p; d.elem_b0 = c->d_elem_b0.p; d.node_f0 = c->d_node_f0.p;
  d.Ntot = c->hb.Ntot();
  d.node_rot = c->hb.has_rotz ? c->d_node_rot.p : nullptr;
  d.U_b = c->any_ub ? c->d_Ub.p : nullptr;
  d.bb_rowptr = c->d_bb_rowptr.p; d.bb_col = c->d_bb_col.p; d.bb_src = c->d_bb_src.p; d.bb_val = c->d_bb_val.p;
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
  const int64_t L = c->L, LK = c->has_case_eai ? L : 1;
  c->LK = LK;
  const int64_t Etot = c->hb.Etot(), nnz = c->sym.nnz_tot(), nf = c->sym.nf_tot();
  const bool fused = sparse_dev(c, 0).asm_max_loc > 0;
  if (!fused) OB_CUDA(c, c->d_elem_k.reserve(size_t(LK) * Etot * ELEM_K));
  c->elem_k_valid = !fused;
  OB_CUDA(c, c->d_kval.reserve(size_t(LK) * std::max<int64_t>(nnz, 1)));
  OB_CUDA(c, c->d_ff_val.reserve(std::max<size_t>(c->plan.ff_col.size(), 1)));
  OB_CUDA(c, c->d_bf_val.reserve(std::max<size_t>(c->plan.bf_col.size(), 1)));
  OB_CUDA(c, c->d_bb_val.reserve(std::max<size_t>(c->plan.bb_col.size(), 1)));
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
  if (o.method != OBGPU_PCG_JACOBI && o.method != OBGPU_PCG_MG)
    return fail(c, OBGPU_EINVAL, "obgpu_solve_static: unknown method (OBGPU_BATCHED_LDLT, OBGPU_PCG_JACOBI, OBGPU_PCG_MG)");
  const bool use_mg = o.method == OBGPU_PCG_MG;
  int rc = assemble_general(c);
  if (rc != OBGPU_OK) return rc;
  if (use_mg) {
    rc = ensure_mg(c);
    if (rc != OBGPU_OK) return rc;
  }
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
  const size_t nvec = use_mg ? 8 : 5;  // + the V-cycle's fine-level residual and two Chebyshev directions
  OB_CUDA(c, c->d_pcgvec.reserve(nvec * nfp));
  double* const v_p = c->d_pcgvec.p;
  double* const v_ap = v_p + nfp;
  double* const v_r = v_ap + nfp;
  double* const v_z = v_r + nfp;
  double* const v_minv = v_z + nfp;
  double* const v_mr = v_minv + nfp;
  double* const v_md0 = v_mr + nfp;
  double* const v_md1 = v_md0 + nfp;
  bool l2_window = false;
  cudaStreamAttrValue l2_av;
  std::memset(&l2_av, 0, sizeof(l2_av));
  {
    static const bool want = !(getenv("OBGPU_L2_PERSIST") && atoi(getenv("OBGPU_L2_PERSIST")) == 0);
    int max_persist = 0, max_window = 0;
    cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, c->device);
    cudaDeviceGetAttribute(&max_window, cudaDevAttrMaxAccessPolicyWindowSize, c->device);
    const size_t slab = nvec * nfp * sizeof(double);
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
  c->prof_valid = false;
  static const bool use_graph = !(getenv("OBGPU_PCG_GRAPH") && atoi(getenv("OBGPU_PCG_GRAPH")) == 0);
  // iterations between convergence checks (even when replayed as a graph); a multigrid iteration is ~40 kernels and
  // the whole solve ~100 iterations, so it is checked every other iteration
  // The multigrid V-cycle reads FP32 copies of its block-SELL matrices (vectors, products and sums stay FP64): a
  // preconditioner needs no more, and the matrices are most of the bytes its kernels move.  OBGPU_MG_FP32=0: FP64 copies.
  const bool mg_fp32 = use_mg && !(getenv("OBGPU_MG_FP32") && atoi(getenv("OBGPU_MG_FP32")) == 0);
  int chk = use_mg ? 2 : use_graph ? 24 : 25;   // iterations between convergence checks (even: the rz ping-pong parity)
  if (const char* ev = getenv("OBGPU_PCG_CHECK")) chk = std::max(2, atoi(ev) & ~1);
  double restarts_total = 0, worst_floor = 0;
  std::vector<double> mg_lam(MG_MAX_LEVELS, 0.0);
  double iters_total = 0, worst_rel = 0;
  bool all_converged = true;
  std::vector<int32_t> h_info(size_t(L) * S, 0);
  for (int64_t cs = 0; cs < L; cs++) {
    const int64_t k = c->has_case_eai ? cs : 0;
    SparseDev d = sparse_dev(c, k);
    if (cs == 0 || c->has_case_eai) {  // K5: BC elimination as gathers, Jacobi diagonal
      OB_CUDA(c, launch_bc_gather(d, c->stream));
      launches += 2;
      if (c->plan.ff_blk.valid) {  // the 3x3 block-SELL copy the SpMV-class kernels stream (76 instead of 108 bytes per block)
        OB_CUDA(c, launch_gather(int64_t(c->plan.ff_blk.map.size()), c->d_ff_bmap.p, d.ff_val, c->d_ff_bval.p, c->stream));
        launches += 1;
      }
      if (!use_mg) {
        OB_CUDA(c, launch_inv_diag(nf, d.ff_diag, d.ff_val, v_minv, c->stream));
        launches += 1;
      } else {
        if (c->plan.ff_blk.valid && mg_fp32) {  // the V-cycle streams FP32 copies of the block matrices (40 B per block)
          const int64_t nblk0 = int64_t(c->plan.ff_blk.map.size() / 9);
          OB_CUDA(c, c->d_ff_bval32.reserve(size_t(10 * nblk0)));
          OB_CUDA(c, launch_gather_blk_f32(nblk0, c->d_ff_bmap.p, d.ff_val, c->d_ff_bval32.p, c->stream));
          launches += 1;
        }
        // numeric setup of the hierarchy from the current K_ff: Galerkin products level by level, Jacobi diagonals and
        // Gershgorin bounds of D^-1 A (the Chebyshev intervals); the coarsest inverses are built per structure below
        const size_t nl = c->mg.lev.size();
        OB_CUDA(c, cudaMemsetAsync(c->d_mg_lam.p, 0, MG_MAX_LEVELS * sizeof(double), c->stream));
        OB_CUDA(c, launch_mg_diag(0, nf, d.sl_ptr, d.ff_col, d.ff_val, v_minv, c->d_mg_lam.p, c->stream));
        launches += 1;
        for (size_t l = 1; l < nl; l++) {
          const MgLevel& ML = c->mg.lev[l];
          obgpu_ctx::MgLevelBufs& B = c->mg_lev[l];
          MgGalerkinArgs g;
          g.n_coarse = ML.n;
          g.f_slptr = l == 1 ? d.sl_ptr : c->mg_lev[l - 1].sl_ptr.p;
          g.f_col = l == 1 ? d.ff_col : c->mg_lev[l - 1].col.p;
          g.f_val = l == 1 ? d.ff_val : c->mg_lev[l - 1].val.p;
          g.f_agg = B.f_agg.p; g.f_kind = B.f_kind.p; g.f_w = B.f_w.p;
          g.mem_ptr = B.mem_ptr.p; g.mem = B.mem.p;
          g.c_slptr = B.sl_ptr.p; g.c_col = B.col.p; g.c_val = B.val.p;
          OB_CUDA(c, launch_mg_galerkin(g, c->stream));
          if (B.nblk) {
            if (mg_fp32) {
              OB_CUDA(c, B.bval32.reserve(size_t(10 * B.nblk)));
              OB_CUDA(c, launch_gather_blk_f32(B.nblk, B.bmap.p, B.val.p, B.bval32.p, c->stream));
            } else {
              OB_CUDA(c, launch_gather(9 * B.nblk, B.bmap.p, B.val.p, B.bval.p, c->stream));
            }
            launches += 1;
          }
          OB_CUDA(c, launch_mg_diag(0, ML.n, B.sl_ptr.p, B.col.p, B.val.p, B.minv.p, c->d_mg_lam.p + l, c->stream));
          launches += 2;
        }
        OB_CUDA(c, cudaMemcpyAsync(mg_lam.data(), c->d_mg_lam.p, MG_MAX_LEVELS * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
        OB_CUDA(c, cudaStreamSynchronize(c->stream));
      }
    }
    const double* nl = c->has_case_nl ? c->d_case_nl.p + size_t(cs) * c->hb.NLtot() : c->d_nload_val.p;
    const double* q = c->has_case_q ? c->d_case_q.p + size_t(cs) * c->hb.ELtot() : c->d_eload_q.p;
    const double* eai = c->has_case_eai ? c->d_case_eai.p + size_t(cs) * 3 * Etot : c->d_elem_EAI.p;
    if (c->any_ub) {  // U_f = K_ff^-1 (F_f - K_bf^T U_b), solver_static.cpp:37-43
      OB_CUDA(c, c->d_rhs.reserve(size_t(std::max<int64_t>(nf, 1))));
    }
    if (L > 1) {  // Eq (equivalent loads on bounded DOFs) of THIS case; F_f was filled by assemble_general
      OB_CUDA(c, launch_load_vector(d, nl, q, eai, c->d_Ff.p + size_t(cs) * nf, c->stream));
      launches += 2;
    }
    if (c->any_ub) {
      OB_CUDA(c, launch_rhs_prescribed(sparse_dev(c, k), c->d_Ff.p + size_t(cs) * nf, c->d_rhs.p, c->stream));
      launches += 1;
    }
    // ---- one PCG per structure; up to NC of them in flight on their own streams, so that the launch gaps and the
    // 25-iteration convergence checks of one structure are covered by the kernels of the others (C5: 8 structures of
    // 0.5 M DOF per GPU were launch-latency bound one after the other) ----
    struct Job {
      int64_t s = -1; PcgArgs a; int it = 0, max_iter = 0; bool busy = false;
      bool verifying = false;   // the recurrence says converged: the true residual b - A x is being computed
      int restarts = 0; double prev_true = -1;
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
      const bool blk0 = c->plan.ff_blk.valid;
      a.bslptr = blk0 ? c->d_ff_bslptr.p : nullptr; a.bcol = blk0 ? c->d_ff_bcol.p : nullptr; a.bval = blk0 ? c->d_ff_bval.p : nullptr;
      a.b = c->any_ub ? c->d_rhs.p : c->d_Ff.p + size_t(cs) * nf;
      a.x = c->d_Uf.p + size_t(cs) * nf;
      a.r = v_r; a.z = v_z; a.p = v_p; a.ap = v_ap;
      a.ctas_per_sm = NC > 1 ? 2 : 0;  // measured at C5's share (8 x 0.5 M DOF): 8 -> 1.84 s, 4 -> 1.67 s, 2 -> 1.62 s
      a.scal = c->d_scal.p + 16 * k; a.partial = c->d_partial.p + size_t(k) * part_stride; a.counter = c->d_counter.p + k;
      a.mg = nullptr;
      if (use_mg) {
        MgJob& J = c->mg_job[k];
        J = MgJob();
        J.nlev = c->mg.nlev[st_i];
        J.dense = c->mg.dense[st_i] != 0;
        J.cinv = c->d_mg_cinv.p + size_t(k) * MgPlan::DENSE_MAX * MgPlan::DENSE_MAX;
        if (const char* ev = getenv("OBGPU_MG_OMEGA")) J.omega = atof(ev);
        for (int l = 0; l < J.nlev; l++) {
          MgLevelDev& V = J.lv[l];
          if (l == 0) {
            V.r0 = a.r0; V.r1 = a.r1;
            V.slptr = d.sl_ptr; V.col = d.ff_col; V.val = d.ff_val; V.minv = v_minv;
            static const bool fine_blk = !(getenv("OBGPU_MG_FINE_BLOCK") && atoi(getenv("OBGPU_MG_FINE_BLOCK")) == 0);
            V.bslptr = fine_blk ? a.bslptr : nullptr; V.bcol = fine_blk ? a.bcol : nullptr; V.bval = fine_blk ? a.bval : nullptr;
            V.bval32 = (fine_blk && a.bval && mg_fp32) ? c->d_ff_bval32.p : nullptr;
            V.x = v_z; V.b = v_r; V.r = v_mr; V.d0 = v_md0; V.d1 = v_md1;
          } else {
            const MgLevel& ML = c->mg.lev[size_t(l)];
            obgpu_ctx::MgLevelBufs& B = c->mg_lev[size_t(l)];
            V.r0 = ML.row_off[st_i]; V.r1 = V.r0 + ML.nrows[st_i];
            V.slptr = B.sl_ptr.p; V.col = B.col.p; V.val = B.val.p; V.minv = B.minv.p;
            V.bslptr = B.nblk ? B.bsl_ptr.p : nullptr; V.bcol = B.nblk ? B.bcol.p : nullptr; V.bval = B.nblk ? B.bval.p : nullptr;
            V.bval32 = (B.nblk && mg_fp32) ? B.bval32.p : nullptr;
            V.x = B.vec.p; V.b = V.x + ML.n; V.r = V.b + ML.n; V.d0 = V.r + ML.n; V.d1 = V.d0 + ML.n;
          }
          if (l + 1 < J.nlev) {  // transfer to the next coarser level (arrays indexed by this level's rows)
            obgpu_ctx::MgLevelBufs& BC = c->mg_lev[size_t(l) + 1];
            V.f_agg = BC.f_agg.p; V.f_kind = BC.f_kind.p; V.f_w = BC.f_w.p; V.mem_ptr = BC.mem_ptr.p; V.mem = BC.mem.p;
          } else {
            V.f_agg = nullptr; V.f_kind = nullptr; V.f_w = nullptr; V.mem_ptr = nullptr; V.mem = nullptr;
          }
          // Chebyshev(2) on [lam/8, lam]
          const double lmax = mg_lam[size_t(l)] > 0 ? mg_lam[size_t(l)] : 1.0, lmin = lmax / 8.0;
          const double theta = 0.5 * (lmax + lmin), delta = 0.5 * (lmax - lmin), sigma = theta / delta;
          const double rho0 = 1.0 / sigma, rho1 = 1.0 / (2.0 * sigma - rho0);
          V.c0 = 1.0 / theta; V.c1 = rho1 * rho0; V.c2 = 2.0 * rho1 / delta;
        }
        if (J.dense) {
          const MgLevelDev& C = J.lv[J.nlev - 1];
          cudaError_t e = launch_mg_dense_inverse(C.r0, C.r1 - C.r0, C.slptr, C.col, C.val, const_cast<double*>(J.cinv),
                                                  c->d_mg_prow.p + size_t(k) * 2 * (MgPlan::DENSE_MAX + 2), pst[size_t(k)]);
          launches += 1;
          if (e != cudaSuccess) return e;
        }
        a.mg = &J;
      }
      if (!c->prof_valid) {  // kept for obgpu_profile_kernel: the first structure's iteration arguments
        c->prof_args = a;
        c->prof_args.mg = nullptr;
        c->prof_mg = use_mg;
        if (use_mg) c->prof_lv0 = c->mg_job[k].lv[0];
        c->prof_valid = true;
      }
      jb.s = st_i; jb.a = a; jb.it = 0; jb.busy = true; jb.verifying = false; jb.restarts = 0; jb.prev_true = -1;
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
        if (jb.it == 0 && !jb.verifying) rr = bb;
        bool conv = false, stop = false;
        if (!(rr == rr)) stop = true;                              // NaN: breakdown (K_ff not positive definite)
        else if (jb.verifying) {
          // rr is now ||F_f - K_ff U_f||^2 computed from U_f itself.  Below the tolerance: done.  Otherwise the
          // recurrence drifted from the true residual: restart CG from it (at most 3 times, and only while a restart
          // still gains something — a tolerance below what FP64 can attain for this K_ff is reported, not looped on)
          jb.verifying = false;
          // FP64 floor of this residual: 16 eps || |K_ff| |U_f| ||_2 (a stiff structure under a soft load pattern cannot
          // be told from converged below it); a tolerance under the floor is met at the floor
          const double floor2 = 256.0 * 4.930380657631324e-32 * c->h_scal[8 * k + 5];
          worst_floor = std::max(worst_floor, bb > 0 ? std::sqrt(floor2 / bb) : 0.0);
          if (rr <= rtol * rtol * bb) { conv = true; stop = true; }
          else if (jb.restarts >= 3 || jb.it >= jb.max_iter || (jb.prev_true >= 0 && rr > 0.25 * jb.prev_true)) {
            stop = true;
            conv = rr <= floor2;   // restarts no longer gain anything: at the floor the solve is as converged as FP64 allows
          } else {
            jb.prev_true = rr;
            jb.restarts++;
            restarts_total += 1;
            OB_CUDA(c, launch_pcg_restart(jb.a, jb.it, c->sm_count, pst[size_t(k)]));
            launches += use_mg ? 40 : 1;
          }
        } else if (rr <= rtol * rtol * bb || jb.it >= jb.max_iter) {
          jb.verifying = true;
          OB_CUDA(c, launch_pcg_true_residual(jb.a, c->sm_count, pst[size_t(k)]));
          launches += 1;
          OB_CUDA(c, cudaMemcpyAsync(c->h_scal + 8 * k, jb.a.scal, 8 * sizeof(double), cudaMemcpyDeviceToHost, pst[size_t(k)]));
          continue;
        }
        if (!stop) {
          const int n_it = std::min(chk, std::max(1, jb.max_iter - jb.it));
          const double per_iter = jb.a.mg ? 3.0 + 1.0 + 6.0 * (jb.a.mg->nlev - 1) + 1.0 : 3.0;
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
            launches += per_iter * n_it;
          } else {
            for (int q2 = 0; q2 < n_it; q2++, jb.it++) OB_CUDA(c, launch_pcg_iter(jb.a, jb.it, c->sm_count, pst[size_t(k)]));
            launches += per_iter * n_it;
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
  c->pcg_restarts = restarts_total;
  c->pcg_levels = use_mg ? double(c->mg.lev.size()) : 1.0;
  c->solved = true;
  c->have_ff = true;
  const bool blk_info = c->plan.ff_blk.valid;
  c->kernel_info = use_mg ? (blk_info ? "k_mg_spmv_blk (3x3 block-SELL SpMV + fused Chebyshev step of the multigrid-PCG V-cycle)"
                                      : "k_mg_spmv (SELL-32 SpMV + fused Chebyshev step of the multigrid-PCG V-cycle)")
                          : (blk_info ? "k_spmv_dot_blk (3x3 block-SELL SpMV of the Jacobi-PCG iteration)"
                                      : "k_spmv_dot (SELL-32 SpMV of the Jacobi-PCG iteration)");
  c->have_stress = o.with_stress != 0;
  if (!all_converged) {
    char msg[384];
    snprintf(msg, sizeof msg, "PCG did not converge: worst true relative residual ||F_f - K_ff U_f||/||F_f|| = %.3e against "
             "rtol %.1e, FP64 floor %.1e (max_iter reached or K_ff not positive definite)", worst_rel, rtol, worst_floor);
    return fail(c, OBGPU_ENOTCONVERGED, msg);
  }
  return OBGPU_OK;
}

// Times `reps` back-to-back launches of one iteration kernel of the last PCG solve on the first structure's rows with
// CUDA events on the ctx stream (bench.py's live roofline figure).  The vectors it touches are scratch of the finished
// solve (p, ap, z and the V-cycle's work vectors): U_f, F_b and the stresses are not disturbed.
int profile_kernel(obgpu_ctx* c, int which, int reps, double* ms_avg, double* alg_bytes) {
  if (!c->prof_valid || !c->solved) return fail(c, OBGPU_ESEQUENCE, "obgpu_profile_kernel: no PCG solve to profile");
  if (which == OBGPU_PROF_MG_SPMV && !c->prof_mg) return fail(c, OBGPU_ESEQUENCE, "obgpu_profile_kernel: the last solve was not OBGPU_PCG_MG");
  if (which != OBGPU_PROF_SPMV && which != OBGPU_PROF_MG_SPMV) return fail(c, OBGPU_EINVAL, "obgpu_profile_kernel: unknown kernel id");
  reps = std::max(1, reps);
  const PcgArgs& a = c->prof_args;
  // entries of the structure's slices (padding included: it is read like the rest)
  const int64_t s0 = a.r0 >> 5, s1 = (int64_t(a.r1) + 31) >> 5;
  const int64_t entries = int64_t(c->plan.sl_ptr[s1]) - int64_t(c->plan.sl_ptr[s0]), n = a.r1 - a.r0;
  cudaEvent_t e0, e1;
  OB_CUDA(c, cudaEventCreate(&e0));
  OB_CUDA(c, cudaEventCreate(&e1));
  for (int pass = 0; pass < 2; pass++) {  // pass 0 = warm-up
    OB_CUDA(c, cudaEventRecord(e0, c->stream));
    for (int i = 0; i < reps; i++) {
      if (which == OBGPU_PROF_SPMV) OB_CUDA(c, launch_spmv_only(a, c->sm_count, c->stream));
      else OB_CUDA(c, launch_mg_cheb_only(c->prof_lv0, c->sm_count, c->stream));
    }
    OB_CUDA(c, cudaEventRecord(e1, c->stream));
    OB_CUDA(c, cudaEventSynchronize(e1));
  }
  float ms = 0;
  OB_CUDA(c, cudaEventElapsedTime(&ms, e0, e1));
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  if (ms_avg) *ms_avg = double(ms) / reps;
  // SpMV = 12 B per stored entry + 20 B per row (x gathered once, y written, p read for the dot);
  // the fused Chebyshev step additionally reads r, D^-1, x and writes r, d, x: 12 B per entry + 56 B per row
  if (alg_bytes) {
    if (a.bval) {  // 3x3 block-SELL: 72 B of values + 4 B of column per stored block
      const int64_t bs0 = (a.r0 / 3) >> 5, bs1 = (int64_t(a.r1) / 3 + 31) >> 5;
      const double blocks = double(c->plan.ff_blk.sl_ptr[size_t(bs1)]) - double(c->plan.ff_blk.sl_ptr[size_t(bs0)]);
      const bool f32 = which != OBGPU_PROF_SPMV && c->mg_job[0].lv[0].bval32 != nullptr;   // the V-cycle's FP32 matrix copy
      *alg_bytes = (f32 ? 44.0 : 76.0) * blocks + (which == OBGPU_PROF_SPMV ? 20.0 : 56.0) * double(n);
    } else {
      *alg_bytes = 12.0 * double(entries) + (which == OBGPU_PROF_SPMV ? 20.0 : 56.0) * double(n);
    }
  }
  return OBGPU_OK;
}

static int ensure_pattern(obgpu_ctx* c) {
  if (c->sym.has_pattern) return OBGPU_OK;
  std::string err;
  const int rc = run_symbolic(c->hb, c->sym, err, true);
  if (rc != OBGPU_OK) return fail(c, rc, err);
  return OBGPU_OK;
}

// extents of the K_ff / K_bf / K_bb views per K copy
static void subk_counts(const Symbolic& sy, int64_t S, int64_t& ff, int64_t& bf, int64_t& bb) {
  ff = bf = bb = 0;
  for (int64_t s = 0; s < S; s++) {
    const int64_t nf = sy.free_off[s + 1] - sy.free_off[s], nb = sy.bnd_off[s + 1] - sy.bnd_off[s];
    ff += nf * nf; bf += nb * nf; bb += nb * nb;
  }
}

// true-sparse K_ff (lower CSR over free indices) pattern, from the pattern of K
static void kff_csr_pattern(const obgpu_ctx* c, std::vector<int32_t>* rowptr, std::vector<int32_t>* col, std::vector<int32_t>* src) {
  const Symbolic& sy = c->sym;
  const int64_t S = c->hb.S;
  for (int64_t s = 0; s < S; s++) {
    const int64_t k0 = sy.dof_off[s], n = sy.dof_off[s + 1] - k0, z0 = sy.nnz_off[s];
    const int32_t* rp = &sy.k_rowptr[size_t(k0 + s)];
    int32_t run = 0;
    if (rowptr) rowptr->push_back(0);
    for (int64_t hi = 0; hi < n; hi++) {
      if (sy.dof_type[size_t(k0 + hi)] < 0) continue;
      for (int32_t z = rp[hi]; z < rp[hi + 1]; z++) {
        const int32_t tlo = sy.dof_type[size_t(k0 + sy.k_col[size_t(z0 + z)])];
        if (tlo < 0) continue;
        if (col) col->push_back(tlo);
        if (src) src->push_back(int32_t(z0 + z));
        run++;
      }
      if (rowptr) rowptr->push_back(run);
    }
  }
}

int symbolic_general(obgpu_ctx* c, int which, int32_t* out, int64_t* count) {  // OBGPU_SYM_KFF_*
  int rc = ensure_pattern(c);
  if (rc != OBGPU_OK) return rc;
  std::vector<int32_t> v;
  if (which == OBGPU_SYM_KFF_ROWPTR) kff_csr_pattern(c, &v, nullptr, nullptr);
  else kff_csr_pattern(c, nullptr, &v, nullptr);
  if (count) *count = int64_t(v.size());
  if (out && !v.empty()) std::memcpy(out, v.data(), v.size() * sizeof(int32_t));
  return OBGPU_OK;
}

int result_count_general(obgpu_ctx* c, int which, int64_t* count) {
  int64_t ff, bf, bb;
  subk_counts(c->sym, c->hb.S, ff, bf, bb);
  const int64_t LK = c->has_case_eai ? c->L : 1;
  switch (which) {
    case OBGPU_RES_KFF_DENSE: *count = LK * ff; return OBGPU_OK;
    case OBGPU_RES_KBF_DENSE: *count = LK * bf; return OBGPU_OK;
    case OBGPU_RES_KBB_DENSE: *count = LK * bb; return OBGPU_OK;
    case OBGPU_RES_KFF_CSR_VAL: {
      int rc = ensure_pattern(c);
      if (rc != OBGPU_OK) return rc;
      std::vector<int32_t> col;
      kff_csr_pattern(c, nullptr, &col, nullptr);
      *count = LK * int64_t(col.size());
      return OBGPU_OK;
    }
  }
  return OBGPU_EINVAL;
}

// BuildProblemInfo::K_ff / K_bf / K_bb: the values of K (assembled on the device) regrouped on the host into the
// reference's dense-structured sub-matrices — data movement only, no arithmetic.
static int download_subk(obgpu_ctx* c, int which, double* out) {
  if (!c->assembled) return fail(c, OBGPU_ESEQUENCE, "obgpu_download(K_ff/K_bf/K_bb) before obgpu_assemble");
  int rc = ensure_pattern(c);
  if (rc != OBGPU_OK) return rc;
  const Symbolic& sy = c->sym;
  const int64_t S = c->hb.S, nnz = sy.nnz_tot(), LK = c->LK;
  std::vector<double> kval(size_t(LK * nnz));
  if (!kval.empty()) OB_CUDA(c, cudaMemcpyAsync(kval.data(), c->d_kval.p, kval.size() * 8, cudaMemcpyDeviceToHost, c->stream));
  OB_CUDA(c, cudaStreamSynchronize(c->stream));
  if (which == OBGPU_RES_KFF_CSR_VAL) {
    std::vector<int32_t> src;
    kff_csr_pattern(c, nullptr, nullptr, &src);
    for (int64_t k = 0; k < LK; k++)
      for (size_t i = 0; i < src.size(); i++) out[size_t(k) * src.size() + i] = kval[size_t(k * nnz + src[i])];
    return OBGPU_OK;
  }
  int64_t ff, bf, bb;
  subk_counts(sy, S, ff, bf, bb);
  const int64_t per = which == OBGPU_RES_KFF_DENSE ? ff : which == OBGPU_RES_KBF_DENSE ? bf : bb;
  if (per > (int64_t(1) << 28)) return fail(c, OBGPU_EUNSUPPORTED, "dense-structured K_ff/K_bf/K_bb view too large: read OBGPU_RES_KFF_CSR_VAL");
  std::memset(out, 0, size_t(LK * per) * 8);
  for (int64_t k = 0; k < LK; k++) {
    double* o = out + k * per;
    int64_t base = 0;
    for (int64_t s = 0; s < S; s++) {
      const int64_t k0 = sy.dof_off[s], n = sy.dof_off[s + 1] - k0, z0 = sy.nnz_off[s];
      const int64_t nf = sy.free_off[s + 1] - sy.free_off[s], nb = sy.bnd_off[s + 1] - sy.bnd_off[s];
      const int32_t* rp = &sy.k_rowptr[size_t(k0 + s)];
      for (int64_t hi = 0; hi < n; hi++) {
        const int32_t thi = sy.dof_type[size_t(k0 + hi)];
        for (int32_t z = rp[hi]; z < rp[hi + 1]; z++) {
          const int32_t tlo = sy.dof_type[size_t(k0 + sy.k_col[size_t(z0 + z)])];
          const double v = kval[size_t(k * nnz + z0 + z)];   // K(lo, hi), lo <= hi
          if (which == OBGPU_RES_KFF_DENSE) {
            if (thi >= 0 && tlo >= 0) o[base + int64_t(tlo) * nf + thi] = v;              // K_ff(c = f(hi), r = f(lo)): lower
          } else if (which == OBGPU_RES_KBB_DENSE) {
            if (thi < 0 && tlo < 0) o[base + int64_t(-thi - 1) * nb + (-tlo - 1)] = v;    // K_bb(b(lo), b(hi)): upper
          } else {
            if (thi < 0 && tlo >= 0) o[base + int64_t(tlo) * nb + (-thi - 1)] = v;        // K_bf(b, f)
            else if (thi >= 0 && tlo < 0) o[base + int64_t(thi) * nb + (-tlo - 1)] = v;
          }
        }
      }
      base += which == OBGPU_RES_KFF_DENSE ? nf * nf : which == OBGPU_RES_KBF_DENSE ? nb * nf : nb * nb;
    }
  }
  return OBGPU_OK;
}

int download_general(obgpu_ctx* c, int which, void* out, bool& handled) {
  handled = false;
  if (which >= OBGPU_RES_KFF_DENSE && which <= OBGPU_RES_KFF_CSR_VAL) {
    handled = true;
    return download_subk(c, which, static_cast<double*>(out));
  }
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
