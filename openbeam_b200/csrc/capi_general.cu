// capi_general.cu — general (any-topology) path: device copies of the batch, the warp-per-structure fused
// kernel for small structures, summaries.  Assembly of the global K and the PCG route live in capi_sparse.cu.
#include <cstdlib>
#include <algorithm>

#include "ctx.h"

namespace obgpu {

int solve_sparse(obgpu_ctx* c, const obgpu_solve_opts& o);  // capi_sparse.cu
int materialize_host_batch(obgpu_ctx* c);                     // capi_devsym.cu
int build_frame_classes(obgpu_ctx* c, const std::vector<int32_t>& nf, const std::vector<int32_t>& nb, const std::vector<int32_t>& hbw);

// Size classes of the warp-per-structure kernel: per-warp shared memory is sized for the largest structure a launch
// serves, so structures are bucketed by their need (band + vectors, in doubles) and every class gets its own launch and
// occupancy.  Inputs per structure: n_f, n_b and the half bandwidth the kernel will use.
int build_frame_classes(obgpu_ctx* c, const std::vector<int32_t>& nf, const std::vector<int32_t>& nb, const std::vector<int32_t>& hbw) {
  const int64_t S = int64_t(nf.size());
  constexpr int NCLS = FRAME_NCLS;
  auto cls_of = [&](int64_t s) { return frame_class_of(hbw[size_t(s)], nf[size_t(s)], nb[size_t(s)]); };
  c->max_band = c->max_nf = c->max_nb = 0;
  std::vector<uint8_t> cls;
  cls.resize(size_t(S));
  std::vector<int64_t> cnt(NCLS + 1, 0);
  for (int64_t s = 0; s < S; s++) {
    cls[size_t(s)] = uint8_t(cls_of(s));
    cnt[cls[size_t(s)] + 1]++;
    c->max_band = int(std::max<int64_t>(c->max_band, int64_t(frame_ld(hbw[size_t(s)])) * nf[size_t(s)]));
    c->max_nf = std::max(c->max_nf, nf[size_t(s)]);
    c->max_nb = std::max(c->max_nb, nb[size_t(s)]);
  }
  for (int k = 0; k < NCLS; k++) cnt[k + 1] += cnt[k];
  std::vector<int32_t> order;
  order.resize(size_t(S));
  c->frame_classes.clear();
  std::vector<int64_t> pos(cnt.begin(), cnt.end() - 1);
  for (int64_t s = 0; s < S; s++) order[size_t(pos[cls[size_t(s)]]++)] = int32_t(s);
  for (int k = 0; k < NCLS; k++) {
    if (cnt[k + 1] == cnt[k]) continue;
    obgpu_ctx::FrameClass fc{cnt[k], cnt[k + 1], 0, 0, 0, 0};
    for (int64_t i = cnt[k]; i < cnt[k + 1]; i++) {
      const size_t s = size_t(order[size_t(i)]);
      fc.max_band = int(std::max<int64_t>(fc.max_band, int64_t(frame_ld(hbw[s])) * nf[s]));
      fc.max_nf = std::max(fc.max_nf, nf[s]);
      fc.max_nb = std::max(fc.max_nb, nb[s]);
      fc.max_hb = std::max(fc.max_hb, hbw[s]);
    }
    c->frame_classes.push_back(fc);
  }
  OB_CUDA(c, c->d_frame_order.upload(order, c->stream));
  OB_CUDA(c, cudaStreamSynchronize(c->stream));  // order is a stack vector
  return OBGPU_OK;
}

// Upload the batch arrays and the per-structure records once per (upload, symbolic).
int ensure_device_batch(obgpu_ctx* c) {
  if (c->dev_batch) return OBGPU_OK;
  const HostBatch& hb = c->hb;
  const Symbolic& sy = c->sym;
  cudaStream_t st = c->stream;
  OB_CUDA(c, c->d_node_xy.upload(hb.node_xy, st));
  if (hb.has_rotz) OB_CUDA(c, c->d_node_rot.upload(hb.node_rotz, st));
  if (hb.has_load_kinds) {
    OB_CUDA(c, c->d_eload_type.upload(hb.eload_type, st));
    OB_CUDA(c, c->d_eload_q2.upload(hb.eload_q2, st));
  }
  OB_CUDA(c, c->d_elem_nodes.upload(hb.elem_nodes, st));
  OB_CUDA(c, c->d_elem_type.upload(hb.elem_type, st));
  OB_CUDA(c, c->d_elem_EAI.upload(hb.elem_EAI, st));
  OB_CUDA(c, c->d_nload_val.upload(hb.nload_val, st));
  OB_CUDA(c, c->d_eload_q.upload(hb.eload_q, st));
  OB_CUDA(c, c->d_eload_dir.upload(hb.eload_dir, st));
  // Band reduction for the batched-LDLt route.  The reference numbers DOFs by node id, and mesh() appends the
  // intermediate nodes after all original ones, so a meshed beam's K_ff has a band as wide as the structure.  Where a
  // reverse Cuthill-McKee order of the nodes narrows the band, the route factors K_ff in that order: the slots handed to
  // the kernel are renumbered and U_f / F_f go back to the reference's numbering on the way out (free_inv).  Structures
  // whose band does not shrink keep their numbering (and their bit-for-bit results).
  std::vector<int32_t> fslot(sy.node_slot), fnl(sy.nload_slot), finv(size_t(sy.nf_tot()));
  std::vector<int32_t> half_bw(sy.half_bw);
  c->frame_renumbered = false;
  for (int64_t s = 0; s < hb.S; s++) {
    const int64_t n0 = hb.node_off[s], N = hb.node_off[s + 1] - n0, e0 = hb.elem_off[s], E = hb.elem_off[s + 1] - e0;
    const int64_t f0 = sy.free_off[s], nf = sy.free_off[s + 1] - f0;
    for (int64_t i = 0; i < nf; i++) finv[size_t(f0 + i)] = int32_t(i);
    if (sy.half_bw[s] <= 12 || getenv("OBGPU_FRAME_NO_RCM")) continue;
    std::vector<std::vector<int32_t>> adj;
    adj.resize(size_t(N));
    for (int64_t e = 0; e < E; e++) {
      const int32_t a = int32_t(hb.elem_nodes[2 * (e0 + e)]), b = int32_t(hb.elem_nodes[2 * (e0 + e) + 1]);
      if (a != b) { adj[size_t(a)].push_back(b); adj[size_t(b)].push_back(a); }
    }
    auto bfs = [&](int32_t root, std::vector<int32_t>& out, std::vector<char>& seen) {
      size_t head = out.size();
      out.push_back(root); seen[size_t(root)] = 1;
      while (head < out.size()) {
        const int32_t u = out[head++];
        std::vector<int32_t> nb;
        for (int32_t v : adj[size_t(u)]) if (!seen[size_t(v)]) { seen[size_t(v)] = 1; nb.push_back(v); }
        std::sort(nb.begin(), nb.end(), [&](int32_t x, int32_t y) {
          return adj[size_t(x)].size() != adj[size_t(y)].size() ? adj[size_t(x)].size() < adj[size_t(y)].size() : x < y; });
        nb.erase(std::unique(nb.begin(), nb.end()), nb.end());
        for (int32_t v : nb) out.push_back(v);
      }
    };
    std::vector<int32_t> ord;
    std::vector<char> seen(size_t(N), 0);
    for (int64_t start = 0; start < N; start++) {
      if (seen[size_t(start)]) continue;
      // pseudo-peripheral root of this component: the last node of a BFS from its lowest-degree node
      std::vector<int32_t> comp; std::vector<char> tmp(seen);
      bfs(int32_t(start), comp, tmp);
      int32_t root = comp[0];
      for (int32_t v : comp) if (adj[size_t(v)].size() < adj[size_t(root)].size()) root = v;
      std::vector<int32_t> far; std::vector<char> tmp2(seen);
      bfs(root, far, tmp2);
      bfs(far.back(), ord, seen);
    }
    std::reverse(ord.begin(), ord.end());
    std::vector<int32_t> perm(size_t(nf), -1);  // reference free index -> internal slot
    int32_t next = 0;
    for (int32_t nd : ord)
      for (int a = 0; a < 3; a++) {
        const int32_t t = sy.node_slot[size_t(3 * (n0 + nd) + a)];
        if (t >= 0) perm[size_t(t)] = next++;
      }
    if (next != nf) continue;  // free DOFs outside the planar slots: leave this structure alone
    int32_t bw = 0;
    for (int64_t e = 0; e < E; e++) {
      int32_t lo = INT32_MAX, hi = -1;
      for (int f = 0; f < 2; f++)
        for (int a = 0; a < 3; a++) {
          const int32_t t = sy.node_slot[size_t(3 * (n0 + hb.elem_nodes[2 * (e0 + e) + f]) + a)];
          if (t >= 0) { lo = std::min(lo, perm[size_t(t)]); hi = std::max(hi, perm[size_t(t)]); }
        }
      if (hi >= 0) bw = std::max(bw, hi - lo);
    }
    if (bw >= sy.half_bw[s]) continue;
    half_bw[size_t(s)] = bw;
    c->frame_renumbered = true;
    for (int64_t i = 0; i < 3 * N; i++) {
      int32_t& t = fslot[size_t(3 * n0 + i)];
      if (t >= 0) t = perm[size_t(t)];
    }
    for (int64_t l = hb.nload_off[s]; l < hb.nload_off[s + 1]; l++)
      if (fnl[size_t(l)] >= 0) fnl[size_t(l)] = perm[size_t(fnl[size_t(l)])];
    for (int64_t i = 0; i < nf; i++) finv[size_t(f0 + perm[size_t(i)])] = int32_t(i);
  }
  OB_CUDA(c, c->d_node_slot.upload(sy.node_slot, st));    // reference numbering: the sparse route's kernels
  OB_CUDA(c, c->d_nload_slot.upload(sy.nload_slot, st));
  if (c->frame_renumbered) {
    OB_CUDA(c, c->d_frame_slot.upload(fslot, st));
    OB_CUDA(c, c->d_frame_nlslot.upload(fnl, st));
    OB_CUDA(c, c->d_frame_inv.upload(finv, st));
  }
  OB_CUDA(c, c->d_el_ptr.upload(sy.el_ptr, st));
  OB_CUDA(c, c->d_el_order.upload(sy.el_order, st));
  std::vector<FrameMeta> meta(hb.S);
  c->max_band = c->max_nf = c->max_nb = 0;
  for (int64_t s = 0; s < hb.S; s++) {
    FrameMeta& m = meta[s];
    m.n0 = hb.node_off[s]; m.e0 = hb.elem_off[s]; m.nl0 = hb.nload_off[s];
    m.f0 = sy.free_off[s]; m.b0 = sy.bnd_off[s];
    m.N = int32_t(hb.node_off[s + 1] - hb.node_off[s]);
    m.E = int32_t(hb.elem_off[s + 1] - hb.elem_off[s]);
    m.NL = int32_t(hb.nload_off[s + 1] - hb.nload_off[s]);
    m.n = int32_t(sy.dof_off[s + 1] - sy.dof_off[s]);
    m.nf = int32_t(sy.free_off[s + 1] - sy.free_off[s]);
    m.nb = int32_t(sy.bnd_off[s + 1] - sy.bnd_off[s]);
    m.hb = std::min(half_bw[size_t(s)], std::max(m.nf - 1, 0));
    m.pad = 0;
    c->max_band = std::max<int64_t>(c->max_band, int64_t(frame_ld(m.hb)) * m.nf);
    c->max_nf = std::max(c->max_nf, m.nf);
    c->max_nb = std::max(c->max_nb, m.nb);
  }
  OB_CUDA(c, c->d_meta.upload(meta, st));
  {
    std::vector<int32_t> v_nf, v_nb, v_hb;
    v_nf.resize(size_t(hb.S)); v_nb.resize(size_t(hb.S)); v_hb.resize(size_t(hb.S));
    for (int64_t s = 0; s < hb.S; s++) { v_nf[size_t(s)] = meta[s].nf; v_nb[size_t(s)] = meta[s].nb; v_hb[size_t(s)] = meta[s].hb; }
    const int rc = build_frame_classes(c, v_nf, v_nb, v_hb);
    if (rc != OBGPU_OK) return rc;
  }
  OB_CUDA(c, cudaStreamSynchronize(st));  // meta, order are stack vectors
  c->dev_batch = true;
  return OBGPU_OK;
}

static bool frame_kernel_fits(const obgpu_ctx* c) {
  return frame_smem_per_warp(c->max_band, c->max_nf, c->max_nb) * FRAME_WARPS <= size_t(227) * 1024;
}

int solve_general(obgpu_ctx* c, const obgpu_solve_opts& o) {
  int rc = ensure_device_batch(c);
  if (rc != OBGPU_OK) return rc;
  const bool any_ub = c->any_ub;
  if (o.method != OBGPU_BATCHED_LDLT || !frame_kernel_fits(c)) {
    if (c->device_batch_mode && !c->host_batch_valid) {  // the sparse route plans on the host: bring the batch there once
      rc = materialize_host_batch(c);
      if (rc == OBGPU_OK) rc = obgpu_symbolic(c);
      if (rc == OBGPU_OK) rc = ensure_device_batch(c);
      if (rc != OBGPU_OK) return rc;
    }
    if (o.method == OBGPU_BATCHED_LDLT)
      return fail(c, OBGPU_EUNSUPPORTED, "structure too large for the batched LDLt kernel: use a PCG method");
    return solve_sparse(c, o);
  }
  const Symbolic& sy = c->sym;
  const int64_t L = c->L;
  OB_CUDA(c, c->d_Uf.reserve(size_t(L) * std::max<int64_t>(sy.nf_tot(), 1)));
  OB_CUDA(c, c->d_Ff.reserve(size_t(L) * std::max<int64_t>(sy.nf_tot(), 1)));
  OB_CUDA(c, c->d_Fb.reserve(size_t(L) * std::max<int64_t>(sy.nb_tot(), 1)));
  OB_CUDA(c, c->d_info.reserve(size_t(L) * c->hb.S));
  OB_CUDA(c, c->d_fail.reserve(1));
  if (o.with_stress) OB_CUDA(c, c->d_stress.reserve(size_t(L) * c->hb.Etot() * 12));
  OB_CUDA(c, cudaMemsetAsync(c->d_fail.p, 0, sizeof(int32_t), c->stream));
  FrameArgs a;
  a.S = c->hb.S; a.L = L;
  a.Ntot = c->hb.Ntot(); a.Etot = c->hb.Etot(); a.NLtot = c->hb.NLtot(); a.ELtot = c->hb.ELtot();
  a.nf_tot = sy.nf_tot(); a.nb_tot = sy.nb_tot();
  a.meta = c->d_meta.p;
  a.node_xy = reinterpret_cast<const double2*>(c->d_node_xy.p);
  a.elem_nodes = reinterpret_cast<const uint2*>(c->d_elem_nodes.p);
  a.elem_type = c->d_elem_type.p; a.elem_EAI = c->d_elem_EAI.p;
  a.node_slot = c->frame_renumbered ? c->d_frame_slot.p : c->d_node_slot.p;
  a.nload_slot = c->frame_renumbered ? c->d_frame_nlslot.p : c->d_nload_slot.p;
  a.nload_val = c->d_nload_val.p;
  a.el_ptr = c->d_el_ptr.p; a.el_order = c->d_el_order.p; a.eload_q = c->d_eload_q.p; a.eload_dir = c->d_eload_dir.p;
  a.eload_type = c->hb.has_load_kinds ? c->d_eload_type.p : nullptr;
  a.eload_q2 = c->hb.has_load_kinds ? c->d_eload_q2.p : nullptr;
  a.U_b = any_ub ? c->d_Ub.p : nullptr;
  a.node_rot = c->hb.has_rotz ? c->d_node_rot.p : nullptr;
  a.node_xy_len = c->have_xy_len ? reinterpret_cast<const double2*>(c->d_node_xy_len.p) : nullptr;
  a.case_nl = c->has_case_nl ? c->d_case_nl.p : nullptr;
  a.case_q = c->has_case_q ? c->d_case_q.p : nullptr;
  a.case_eai = c->has_case_eai ? c->d_case_eai.p : nullptr;
  a.U_f = c->d_Uf.p; a.F_b = c->d_Fb.p; a.F_f = c->d_Ff.p;
  a.stress = o.with_stress ? c->d_stress.p : nullptr;
  a.info = c->d_info.p; a.fail = c->d_fail.p;
  a.order = c->d_frame_order.p;
  a.free_inv = c->frame_renumbered ? c->d_frame_inv.p : nullptr;
  // One load case, identity node frames, U_b = 0 and at least a few thousand structures: the three-kernel path (element
  // arithmetic of stages A and C in kernels of their own, kernels_frame.cu).  Everything else: the fused kernel.
  // (OBGPU_FRAME_SPLIT=0 switches it off, OBGPU_FRAME_SPLIT_MIN=n moves the structure-count threshold: tests, A/B runs)
  const char* env_split = std::getenv("OBGPU_FRAME_SPLIT");
  const char* env_min = std::getenv("OBGPU_FRAME_SPLIT_MIN");
  const int64_t split_min = env_min ? std::atoll(env_min) : 4096;
  const bool split = !(env_split && env_split[0] == '0') && L == 1 && !a.U_b && !a.node_rot && !a.node_xy_len && !a.case_nl && !a.case_q && !a.case_eai &&
                     c->hb.S >= split_min && c->max_nb <= 1024;
  a.erec = nullptr;
  if (split) {
    OB_CUDA(c, c->d_erec.reserve(size_t(c->hb.Etot()) * FRAME_EREC));
    a.erec = c->d_erec.p;
  }
  OB_CUDA(c, cudaEventRecord(c->ev[2], c->stream));
  if (split) OB_CUDA(c, launch_frame_pre(a, c->sm_count, c->stream));
  for (auto it = c->frame_classes.rbegin(); it != c->frame_classes.rend(); ++it) {  // largest structures first
    const obgpu_ctx::FrameClass& fc = *it;
    a.o0 = fc.o0; a.o1 = fc.o1;
    a.max_band = fc.max_band; a.max_nf = fc.max_nf; a.max_nb = fc.max_nb; a.max_hb = fc.max_hb;
    OB_CUDA(c, launch_frame_fused(a, c->sm_count, c->stream));
  }
  if (split) OB_CUDA(c, launch_frame_post(a, c->max_nb, c->sm_count, c->stream));
  OB_CUDA(c, cudaEventRecord(c->ev[3], c->stream));
  c->t_have[1] = true;
  c->launches = double(c->frame_classes.size()) + (split ? 2 : 0);
  c->pcg_iters = 0;
  c->solved = true;
  c->have_ff = true;
  c->kernel_info = split ? "k_frame_pre + k_frame_fused<split> + k_frame_post (warp per structure, banded LDLt in shared memory)"
                         : "k_frame_fused (warp per structure, banded LDLt in shared memory)";
  c->have_stress = o.with_stress != 0;
  return OBGPU_OK;
}

}  // namespace obgpu

extern "C" int obgpu_summaries_device(obgpu_ctx* c, double* d_out) {
  if (!c || !d_out) return OBGPU_EINVAL;
  OB_NEED_DEVICE(c);
  if (!c->solved) return obgpu::fail(c, OBGPU_ESEQUENCE, "obgpu_summaries_device before obgpu_solve_static");
  if (reinterpret_cast<uintptr_t>(d_out) % 32) return obgpu::fail(c, OBGPU_EINVAL, "summary buffer must be 32-byte aligned");
  if (c->summ_written == d_out) return OBGPU_OK;  // bound buffer: the solve's own kernel already wrote them
  const obgpu::SummaryArgs a = obgpu::summary_args(c, d_out);
  OB_CUDA(c, obgpu::launch_summaries(a, c->sm_count, c->stream));
  c->launches += 1;
  return OBGPU_OK;
}
