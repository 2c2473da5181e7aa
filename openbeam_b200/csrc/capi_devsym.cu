// capi_devsym.cu — "device batch" mode for batches of MANY small structures (BASELINE config 3): obgpu_batch_upload_device
// sends the caller's arrays straight to HBM (the host keeps only the five offset arrays) and obgpu_symbolic runs K3 as
// device kernels (kernels_symbolic.cu) instead of the per-structure host loop.  For 524 288 different frames this turns
// 3.5 s of host copying + 1.5 s of host symbolic (round 1) into two PCIe transfers and three kernels.
//
// The batched-LDLt route, summaries and all result downloads work directly on a device batch.  Anything that needs the
// host-side description — obgpu_get_symbolic's integer artefacts, the K pattern, the sparse/PCG route, the thread-per-case
// kernel — first materialises it (one download of the batch + the host symbolic phase); results are identical either way.
#include <algorithm>
#include <chrono>
#include <cstring>

#include "ctx.h"

namespace obgpu {

static const char* sym_error_text(int code) {
  switch (code) {
    case SYM_E_ELEM_NODE: return "element node index out of range (ASSERT_(conected_nodes_ids[k] < nNodes))";
    case SYM_E_UNINIT: return "Uninitialized parameter E/A/Iz";
    case SYM_E_ELEM_TYPE: return "unknown element type";
    case SYM_E_ZERO_LEN: return "zero-length element (ASSERT_(L > 0))";
    case SYM_E_NONPLANAR: return "element with out-of-plane DOFs";
    case SYM_E_CONSTRAINT: return "constraint node/dof out of range";
    case SYM_E_NLOAD: return "node load node/dof out of range";
    case SYM_E_ELOAD: return "element load element index / type out of range";
    case SYM_E_LOAD_DIR: return "element load direction is not a unit vector";
    case SYM_E_LOAD_ELEM: return "Unsupported element type for this element load (or load point outside the beam)";
    case SYM_E_NO_DOFS: return "structure without DOFs (ASSERT_(nDOFs > 0))";
  }
  return "malformed structure";
}

int upload_cases_public(obgpu_ctx* c, int64_t L, const double* nl, const double* q, const double* eai);  // capi.cu

int batch_upload_device(obgpu_ctx* c, const obgpu_batch_desc* d) {
  if (d->n_structures < 1) return fail(c, OBGPU_EINVAL, "n_structures must be >= 1");
  if (!d->node_off || !d->elem_off || !d->con_off || !d->nload_off || !d->eload_off)
    return fail(c, OBGPU_EINVAL, "offset arrays are mandatory");
  HostBatch& hb = c->hb;
  {  // reset, keeping the capacity of the five offset vectors (a job that uploads batch after batch reuses the pages)
    std::vector<int64_t> keep[5] = {std::move(hb.node_off), std::move(hb.elem_off), std::move(hb.con_off), std::move(hb.nload_off),
                                    std::move(hb.eload_off)};
    hb = HostBatch();
    hb.node_off = std::move(keep[0]); hb.elem_off = std::move(keep[1]); hb.con_off = std::move(keep[2]);
    hb.nload_off = std::move(keep[3]); hb.eload_off = std::move(keep[4]);
    for (auto* v : {&hb.node_off, &hb.elem_off, &hb.con_off, &hb.nload_off, &hb.eload_off}) v->clear();
  }
  const int64_t S = hb.S = d->n_structures;
  // totals straight from the caller's arrays: the copies below are issued first, the offset arrays are checked on the host
  // while the DMA runs
  const int64_t N = d->node_off[S], E = d->elem_off[S], C = d->con_off[S], NL = d->nload_off[S], EL = d->eload_off[S];
  if (N < 1 || E < 1) return fail(c, OBGPU_EINVAL, "empty structure (ASSERT_(nDOFs > 0))");
  if (C < 0 || NL < 0 || EL < 0) return fail(c, OBGPU_EINVAL, "offset arrays must be non-decreasing");
  if (N >= INT32_MAX / 3 || E >= INT32_MAX / 21) return fail(c, OBGPU_EUNSUPPORTED, "batch too large for 32-bit slots: split it");
  if (!d->node_xy || !d->elem_nodes || !d->elem_type || !d->elem_EAI) return fail(c, OBGPU_EINVAL, "null geometry array");
  if ((C && (!d->con_node || !d->con_dof || !d->con_val)) || (NL && (!d->nload_node || !d->nload_dof || !d->nload_val)) ||
      (EL && (!d->eload_elem || !d->eload_q || !d->eload_dir)))
    return fail(c, OBGPU_EINVAL, "null constraint/load array");
  hb.has_rotz = d->node_rotz != nullptr;          // all-zero angles cost the general kernel instantiation, nothing else
  hb.has_load_kinds = d->eload_type != nullptr && EL > 0;
  cudaStream_t st = c->stream;
  OB_CUDA(c, cudaEventRecord(c->ev[0], st));
  OB_CUDA(c, c->d_off_node.upload(d->node_off, size_t(S + 1), st));
  OB_CUDA(c, c->d_off_elem.upload(d->elem_off, size_t(S + 1), st));
  OB_CUDA(c, c->d_off_con.upload(d->con_off, size_t(S + 1), st));
  OB_CUDA(c, c->d_off_nload.upload(d->nload_off, size_t(S + 1), st));
  OB_CUDA(c, c->d_off_eload.upload(d->eload_off, size_t(S + 1), st));
  OB_CUDA(c, c->d_node_xy.upload(d->node_xy, size_t(2 * N), st));
  if (hb.has_rotz) OB_CUDA(c, c->d_node_rot.upload(d->node_rotz, size_t(N), st));
  OB_CUDA(c, c->d_elem_nodes.upload(d->elem_nodes, size_t(2 * E), st));
  OB_CUDA(c, c->d_elem_type.upload(d->elem_type, size_t(E), st));
  OB_CUDA(c, c->d_elem_EAI.upload(d->elem_EAI, size_t(3 * E), st));
  OB_CUDA(c, c->d_con_node.upload(d->con_node, size_t(C), st));
  OB_CUDA(c, c->d_con_dof.upload(d->con_dof, size_t(C), st));
  OB_CUDA(c, c->d_con_val.upload(d->con_val, size_t(C), st));
  OB_CUDA(c, c->d_nload_node.upload(d->nload_node, size_t(NL), st));
  OB_CUDA(c, c->d_nload_dof.upload(d->nload_dof, size_t(NL), st));
  OB_CUDA(c, c->d_nload_val.upload(d->nload_val, size_t(NL), st));
  OB_CUDA(c, c->d_eload_elem.upload(d->eload_elem, size_t(EL), st));
  OB_CUDA(c, c->d_eload_q.upload(d->eload_q, size_t(EL), st));
  OB_CUDA(c, c->d_eload_dir.upload(d->eload_dir, size_t(3 * EL), st));
  if (hb.has_load_kinds) {
    OB_CUDA(c, c->d_eload_type.upload(d->eload_type, size_t(EL), st));
    if (d->eload_q2) OB_CUDA(c, c->d_eload_q2.upload(d->eload_q2, size_t(EL), st));
    else {
      OB_CUDA(c, c->d_eload_q2.reserve(size_t(EL)));
      OB_CUDA(c, cudaMemsetAsync(c->d_eload_q2.p, 0, size_t(EL) * 8, st));
    }
  }
  OB_CUDA(c, cudaEventRecord(c->ev[1], st));
  c->t_have[0] = true;
  // host copies of the offset arrays (totals, on-demand per-structure offsets) + their checks, while the copies are in flight
  hb.node_off.assign(d->node_off, d->node_off + S + 1);
  hb.elem_off.assign(d->elem_off, d->elem_off + S + 1);
  hb.con_off.assign(d->con_off, d->con_off + S + 1);
  hb.nload_off.assign(d->nload_off, d->nload_off + S + 1);
  hb.eload_off.assign(d->eload_off, d->eload_off + S + 1);
  for (auto* v : {&hb.node_off, &hb.elem_off, &hb.con_off, &hb.nload_off, &hb.eload_off}) {
    bool ok = (*v)[0] == 0;
    for (int64_t s = 0; s < S && ok; s++) ok = (*v)[s + 1] >= (*v)[s];
    if (!ok) {
      cudaStreamSynchronize(st);
      c->uploaded = false;
      return fail(c, OBGPU_EINVAL, (*v)[0] != 0 ? "offset arrays must start at 0" : "offset arrays must be non-decreasing");
    }
  }
  hb.L = d->n_cases < 1 ? 1 : d->n_cases;
  c->uploaded = true;
  c->device_batch_mode = true;
  c->host_batch_valid = false;
  c->sym = Symbolic();
  c->dev_batch = false;
  c->dev_plan = false;
  c->dev_mg = false;
  c->assembled = false;
  c->prog = TinyProgram();
  c->have_xy_len = false;
  const int rc = upload_cases_public(c, hb.L, d->case_nload_val, d->case_eload_q, d->case_elem_EAI);
  if (rc != OBGPU_OK) return rc;
  OB_CUDA(c, cudaStreamSynchronize(st));  // the caller's buffers may be reused as soon as this returns
  return OBGPU_OK;
}

// Size classes of the warp-per-structure kernel (the same table as ensure_device_batch, capi_general.cu)
int build_frame_classes(obgpu_ctx* c, const std::vector<int32_t>& nf, const std::vector<int32_t>& nb, const std::vector<int32_t>& hbw);

int symbolic_device(obgpu_ctx* c) {
  const auto t0 = std::chrono::steady_clock::now();
  const HostBatch& hb = c->hb;
  const int64_t S = hb.S, N = hb.Ntot(), E = hb.Etot(), NL = hb.NLtot(), EL = hb.ELtot(), C = hb.Ctot();
  cudaStream_t st = c->stream;
  OB_CUDA(c, c->d_sym_mask.reserve(size_t(N)));
  OB_CUDA(c, c->d_sym_cval.reserve(size_t(3 * N)));
  OB_CUDA(c, c->d_sym_counts.reserve(size_t(4 * S)));
  OB_CUDA(c, c->d_dof_off.reserve(size_t(S + 1)));
  OB_CUDA(c, c->d_free_off.reserve(size_t(S + 1)));
  OB_CUDA(c, c->d_bnd_off.reserve(size_t(S + 1)));
  OB_CUDA(c, c->d_node_slot.reserve(size_t(3 * N)));
  OB_CUDA(c, c->d_nload_slot.reserve(std::max<size_t>(size_t(NL), 1)));
  OB_CUDA(c, c->d_el_ptr.reserve(size_t(E + 1)));
  OB_CUDA(c, c->d_el_order.reserve(std::max<size_t>(size_t(EL), 1)));
  OB_CUDA(c, c->d_sym_cursor.reserve(size_t(E)));
  OB_CUDA(c, c->d_meta.reserve(size_t(S)));
  // a bounded DOF needs a constraint entry: n_b <= C bounds U_b and the DOF ids without knowing n_b yet
  OB_CUDA(c, c->d_Ub.reserve(std::max<size_t>(size_t(C), 1)));
  OB_CUDA(c, c->d_bnd_dofid.reserve(std::max<size_t>(size_t(C), 1)));
  OB_CUDA(c, c->d_sym_error.reserve(4));
  const int64_t nchunk = (S + 1023) / 1024;
  OB_CUDA(c, c->d_sym_partial.reserve(size_t(3 * (nchunk + 1))));
  OB_CUDA(c, c->d_sym_cls.reserve(size_t(S)));
  OB_CUDA(c, c->d_sym_cls_stats.reserve(size_t(FRAME_NCLS) * 5 + FRAME_NCLS));
  OB_CUDA(c, c->d_frame_order.reserve(size_t(S)));
  OB_CUDA(c, cudaMemsetAsync(c->d_sym_error.p, 0, 4 * sizeof(int32_t), st));
  OB_CUDA(c, cudaMemsetAsync(c->d_sym_cls_stats.p, 0, (size_t(FRAME_NCLS) * 5 + FRAME_NCLS) * sizeof(int32_t), st));
  SymDev d;
  d.S = S;
  d.node_off = c->d_off_node.p; d.elem_off = c->d_off_elem.p; d.con_off = c->d_off_con.p;
  d.nload_off = c->d_off_nload.p; d.eload_off = c->d_off_eload.p;
  d.node_xy = reinterpret_cast<const double2*>(c->d_node_xy.p);
  d.node_rotz = hb.has_rotz ? c->d_node_rot.p : nullptr;
  d.elem_nodes = reinterpret_cast<const uint2*>(c->d_elem_nodes.p);
  d.elem_type = c->d_elem_type.p; d.elem_EAI = c->d_elem_EAI.p;
  d.con_node = c->d_con_node.p; d.con_dof = c->d_con_dof.p; d.con_val = c->d_con_val.p;
  d.nload_node = c->d_nload_node.p; d.nload_dof = c->d_nload_dof.p;
  d.eload_elem = c->d_eload_elem.p; d.eload_type = hb.has_load_kinds ? c->d_eload_type.p : nullptr;
  d.eload_dir = c->d_eload_dir.p; d.eload_q2 = hb.has_load_kinds ? c->d_eload_q2.p : nullptr;
  d.node_mask = c->d_sym_mask.p; d.node_cval = c->d_sym_cval.p; d.counts = c->d_sym_counts.p;
  d.dof_off = c->d_dof_off.p; d.free_off = c->d_free_off.p; d.bnd_off = c->d_bnd_off.p;
  d.node_slot = c->d_node_slot.p; d.nload_slot = c->d_nload_slot.p; d.el_ptr = c->d_el_ptr.p; d.el_order = c->d_el_order.p;
  d.el_cursor = c->d_sym_cursor.p; d.meta = c->d_meta.p; d.U_b = c->d_Ub.p; d.bnd_dofid = c->d_bnd_dofid.p;
  d.error = c->d_sym_error.p;
  d.scan_partial = c->d_sym_partial.p; d.cls = c->d_sym_cls.p; d.cls_stats = c->d_sym_cls_stats.p;
  OB_CUDA(c, launch_symbolic_device(d, st));
  // what comes back to the host: the error word, the three totals and 18 x 5 class statistics — a few hundred bytes
  int32_t err[4] = {0, 0, 0, 0};
  int32_t stats[FRAME_NCLS * 5];
  int64_t totals[3] = {0, 0, 0};
  OB_CUDA(c, cudaMemcpyAsync(err, c->d_sym_error.p, sizeof(err), cudaMemcpyDeviceToHost, st));
  OB_CUDA(c, cudaMemcpyAsync(stats, c->d_sym_cls_stats.p, sizeof(stats), cudaMemcpyDeviceToHost, st));
  OB_CUDA(c, cudaMemcpyAsync(totals, c->d_sym_partial.p + 3 * nchunk, sizeof(totals), cudaMemcpyDeviceToHost, st));
  OB_CUDA(c, cudaStreamSynchronize(st));
  if (err[0] != 0) {
    char msg[256];
    snprintf(msg, sizeof msg, "%s (structure %d of the batch)", sym_error_text(err[0]), err[1]);
    return fail(c, OBGPU_EINVAL, msg);
  }
  Symbolic& sy = c->sym;
  const bool keep = sy.valid && !sy.has_pattern && int64_t(sy.dof_off.size()) == S + 1 && sy.dof_node.empty();
  if (!keep) {
    sy = Symbolic();
    sy.dof_off.assign(size_t(S + 1), 0); sy.free_off.assign(size_t(S + 1), 0); sy.bnd_off.assign(size_t(S + 1), 0);
    sy.nnz_off.assign(size_t(S + 1), 0); sy.tri_off.assign(size_t(S + 1), 0);
  }
  // per-structure offsets stay on the device (obgpu_get_offsets fetches them on demand); the totals size every result
  sy.dof_off.back() = totals[0]; sy.free_off.back() = totals[1]; sy.bnd_off.back() = totals[2];
  c->host_offsets_valid = false;
  sy.has_pattern = false;
  sy.valid = true;
  c->any_ub = err[2] != 0;
  c->frame_renumbered = false;
  // size classes from the device statistics; the order array is filled by one more kernel from the class starts
  c->frame_classes.clear();
  c->max_band = c->max_nf = c->max_nb = 0;
  int32_t starts[FRAME_NCLS];
  int64_t run = 0;
  for (int k = 0; k < FRAME_NCLS; k++) {
    const int32_t* sk = stats + 5 * k;
    starts[k] = int32_t(run);
    if (sk[0] > 0) {
      c->frame_classes.push_back(obgpu_ctx::FrameClass{run, run + sk[0], sk[1], sk[2], sk[3], sk[4]});
      c->max_band = std::max(c->max_band, sk[1]);
      c->max_nf = std::max(c->max_nf, sk[2]);
      c->max_nb = std::max(c->max_nb, sk[3]);
    }
    run += sk[0];
  }
  int32_t* d_cursor = c->d_sym_cls_stats.p + FRAME_NCLS * 5;
  OB_CUDA(c, cudaMemcpyAsync(d_cursor, starts, sizeof(starts), cudaMemcpyHostToDevice, st));
  OB_CUDA(c, launch_symbolic_order(S, c->d_sym_cls.p, d_cursor, c->d_frame_order.p, st));
  OB_CUDA(c, cudaStreamSynchronize(st));  // starts is a stack array
  c->dev_batch = true;
  c->t_symbolic_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
  return OBGPU_OK;
}

// per-structure DOF / free / bounded offsets of a device batch, fetched when a caller asks for them
int fetch_host_offsets(obgpu_ctx* c) {
  if (c->host_offsets_valid) return OBGPU_OK;
  const size_t n = size_t(c->hb.S + 1);
  OB_CUDA(c, cudaMemcpyAsync(c->sym.dof_off.data(), c->d_dof_off.p, n * 8, cudaMemcpyDeviceToHost, c->stream));
  OB_CUDA(c, cudaMemcpyAsync(c->sym.free_off.data(), c->d_free_off.p, n * 8, cudaMemcpyDeviceToHost, c->stream));
  OB_CUDA(c, cudaMemcpyAsync(c->sym.bnd_off.data(), c->d_bnd_off.p, n * 8, cudaMemcpyDeviceToHost, c->stream));
  OB_CUDA(c, cudaStreamSynchronize(c->stream));
  c->host_offsets_valid = true;
  return OBGPU_OK;
}

// Host copy of a device batch + the host symbolic phase, for the callers that need the integer artefacts on the host.
int materialize_host_batch(obgpu_ctx* c) {
  if (!c->device_batch_mode || c->host_batch_valid) return OBGPU_OK;
  HostBatch& hb = c->hb;
  const int64_t N = hb.Ntot(), E = hb.Etot(), C = hb.Ctot(), NL = hb.NLtot(), EL = hb.ELtot();
  cudaStream_t st = c->stream;
  auto down = [&](auto& vec, const auto& buf, size_t n) -> cudaError_t {
    vec.resize(n);
    if (n == 0) return cudaSuccess;
    return cudaMemcpyAsync(vec.data(), buf.p, n * sizeof(vec[0]), cudaMemcpyDeviceToHost, st);
  };
  OB_CUDA(c, down(hb.node_xy, c->d_node_xy, size_t(2 * N)));
  if (hb.has_rotz) OB_CUDA(c, down(hb.node_rotz, c->d_node_rot, size_t(N)));
  OB_CUDA(c, down(hb.elem_nodes, c->d_elem_nodes, size_t(2 * E)));
  OB_CUDA(c, down(hb.elem_type, c->d_elem_type, size_t(E)));
  OB_CUDA(c, down(hb.elem_EAI, c->d_elem_EAI, size_t(3 * E)));
  OB_CUDA(c, down(hb.con_node, c->d_con_node, size_t(C)));
  OB_CUDA(c, down(hb.con_dof, c->d_con_dof, size_t(C)));
  OB_CUDA(c, down(hb.con_val, c->d_con_val, size_t(C)));
  OB_CUDA(c, down(hb.nload_node, c->d_nload_node, size_t(NL)));
  OB_CUDA(c, down(hb.nload_dof, c->d_nload_dof, size_t(NL)));
  OB_CUDA(c, down(hb.nload_val, c->d_nload_val, size_t(NL)));
  OB_CUDA(c, down(hb.eload_elem, c->d_eload_elem, size_t(EL)));
  OB_CUDA(c, down(hb.eload_q, c->d_eload_q, size_t(EL)));
  OB_CUDA(c, down(hb.eload_dir, c->d_eload_dir, size_t(3 * EL)));
  if (hb.has_load_kinds) {
    OB_CUDA(c, down(hb.eload_type, c->d_eload_type, size_t(EL)));
    OB_CUDA(c, down(hb.eload_q2, c->d_eload_q2, size_t(EL)));
  }
  OB_CUDA(c, cudaStreamSynchronize(st));
  if (!hb.has_load_kinds) {
    hb.eload_type.assign(size_t(EL), OBGPU_LOAD_UNIFORM);
    hb.eload_q2.assign(size_t(EL), 0.0);
  }
  if (hb.has_rotz) {
    bool any = false;
    for (double r : hb.node_rotz) any |= r != 0.0;
    if (!any) { hb.has_rotz = false; hb.node_rotz.clear(); }
  }
  c->host_batch_valid = true;
  return OBGPU_OK;
}

}  // namespace obgpu
