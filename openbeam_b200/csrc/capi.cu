// capi.cu — extern "C" surface of libobgpu (include/obgpu.h): context, upload, symbolic, solve, download.
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>

#include "ctx.h"

namespace obgpu {
int fail(obgpu_ctx* c, int code, const std::string& msg) {
  if (c) c->err = msg;
  return code;
}
int cuda_fail(obgpu_ctx* c, cudaError_t e, const char* where) {
  if (c) c->err = std::string("CUDA error: ") + cudaGetErrorString(e) + " at " + where;
  cudaGetLastError();
  return OBGPU_ECUDA;
}
template <class T>
static void copy_vec(std::vector<T>& dst, const T* src, int64_t n) {
  dst.assign(src, src + n);
}
}  // namespace obgpu
namespace obgpu {
int batch_upload_device(obgpu_ctx* c, const obgpu_batch_desc* d);   // capi_devsym.cu
int symbolic_device(obgpu_ctx* c);
int materialize_host_batch(obgpu_ctx* c);
int symbolic_general(obgpu_ctx* c, int which, int32_t* out, int64_t* count);   // capi_sparse.cu
int fetch_host_offsets(obgpu_ctx* c);                                           // capi_devsym.cu
}
using namespace obgpu;

extern "C" {

const char* obgpu_status_string(int s) {
  switch (s) {
    case OBGPU_OK: return "ok";
    case OBGPU_EINVAL: return "invalid structure description";
    case OBGPU_ECUDA: return "CUDA error (no usable sm_100 device?)";
    case OBGPU_ESINGULAR:  // the reference's exception text, solver_static.cpp:102-104
      return "[CFiniteElementProblem::solveStatic] Ill-conditioned matrix. Too few restrictions?";
    case OBGPU_ENOTCONVERGED: return "PCG did not converge";
    case OBGPU_EUNSUPPORTED: return "input outside what the GPU path implements";
    case OBGPU_ESEQUENCE: return "call sequence violated";
  }
  return "unknown status";
}

int obgpu_create(int device, obgpu_ctx** out) {
  if (!out) return OBGPU_EINVAL;
  *out = nullptr;
  if (device == OBGPU_DEVICE_NONE) {  // symbolic-only ctx: integer host phase works, every compute call fails
    auto* c = new obgpu_ctx;
    c->device = OBGPU_DEVICE_NONE;
    *out = c;
    return OBGPU_OK;
  }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) {
    cudaGetLastError();
    return OBGPU_ECUDA;  // no CPU fallback: the product path fails loudly without a GPU
  }
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return OBGPU_ECUDA;
  if (prop.major < 10) return OBGPU_ECUDA;  // kernels are built for sm_100a only
  if (cudaSetDevice(device) != cudaSuccess) return OBGPU_ECUDA;
  auto* c = new obgpu_ctx;
  c->device = device;
  c->sm_count = prop.multiProcessorCount;
  bool ok = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) == cudaSuccess;
  for (int i = 0; i < obgpu_ctx::NSLOT && ok; i++)
    ok = cudaStreamCreateWithFlags(&c->slot_stream[i], cudaStreamNonBlocking) == cudaSuccess;
  for (int i = 0; i < 8 && ok; i++) ok = cudaEventCreate(&c->ev[i]) == cudaSuccess;
  if (!ok) {
    delete c;
    return OBGPU_ECUDA;
  }
  *out = c;
  return OBGPU_OK;
}

void obgpu_destroy(obgpu_ctx* c) {
  if (!c) return;
  if (c->device == OBGPU_DEVICE_NONE) {
    delete c;
    return;
  }
  cudaSetDevice(c->device);
  cudaDeviceSynchronize();
  if (c->h_scal) cudaFreeHost(c->h_scal);
  for (auto& s : c->pcg_stream)
    if (s) cudaStreamDestroy(s);
  if (c->pcg_ev) cudaEventDestroy(c->pcg_ev);
  for (auto& e : c->ev)
    if (e) cudaEventDestroy(e);
  for (auto& s : c->slot_stream)
    if (s) cudaStreamDestroy(s);
  if (c->stream) cudaStreamDestroy(c->stream);
  delete c;
}

const char* obgpu_last_error(const obgpu_ctx* c) { return c ? c->err.c_str() : "null ctx"; }

int obgpu_host_alloc(void** p, size_t bytes) {
  return cudaHostAlloc(p, bytes, cudaHostAllocDefault) == cudaSuccess ? OBGPU_OK : OBGPU_ECUDA;
}
int obgpu_host_free(void* p) { return cudaFreeHost(p) == cudaSuccess ? OBGPU_OK : OBGPU_ECUDA; }

static int upload_cases(obgpu_ctx* c, int64_t L, const double* nl, const double* q, const double* eai);
extern "C++" {
namespace obgpu {
int upload_cases_public(obgpu_ctx* c, int64_t L, const double* nl, const double* q, const double* eai) {
  return upload_cases(c, L, nl, q, eai);
}
}
}
static int upload_cases(obgpu_ctx* c, int64_t L, const double* nl, const double* q, const double* eai) {
  if (L < 1) return fail(c, OBGPU_EINVAL, "n_cases must be >= 1");
  c->L = L;
  c->has_case_nl = nl != nullptr && c->hb.NLtot() > 0;
  c->has_case_q = q != nullptr && c->hb.ELtot() > 0;
  c->has_case_eai = eai != nullptr;
  c->solved = false;
  c->assembled = false;
  if (c->device == OBGPU_DEVICE_NONE) return OBGPU_OK;
  OB_CUDA(c, cudaEventRecord(c->ev[0], c->stream));
  if (c->has_case_nl) OB_CUDA(c, c->d_case_nl.upload(nl, size_t(L) * c->hb.NLtot(), c->stream));
  if (c->has_case_q) OB_CUDA(c, c->d_case_q.upload(q, size_t(L) * c->hb.ELtot(), c->stream));
  if (c->has_case_eai) OB_CUDA(c, c->d_case_eai.upload(eai, size_t(L) * 3 * c->hb.Etot(), c->stream));
  OB_CUDA(c, cudaEventRecord(c->ev[1], c->stream));
  c->t_have[0] = true;
  c->solved = false;
  return OBGPU_OK;
}

int obgpu_batch_upload_device(obgpu_ctx* c, const obgpu_batch_desc* d) {
  if (!c || !d) return OBGPU_EINVAL;
  OB_NEED_DEVICE(c);
  return batch_upload_device(c, d);
}

int obgpu_batch_upload(obgpu_ctx* c, const obgpu_batch_desc* d) {
  if (!c || !d) return OBGPU_EINVAL;
  OB_DEVICE(c);
  c->device_batch_mode = false;
  c->host_batch_valid = true;
  if (d->n_structures < 1) return fail(c, OBGPU_EINVAL, "n_structures must be >= 1");
  if (!d->node_off || !d->elem_off || !d->con_off || !d->nload_off || !d->eload_off)
    return fail(c, OBGPU_EINVAL, "offset arrays are mandatory");
  HostBatch& hb = c->hb;
  hb = HostBatch();
  const int64_t S = hb.S = d->n_structures;
  copy_vec(hb.node_off, d->node_off, S + 1);
  copy_vec(hb.elem_off, d->elem_off, S + 1);
  copy_vec(hb.con_off, d->con_off, S + 1);
  copy_vec(hb.nload_off, d->nload_off, S + 1);
  copy_vec(hb.eload_off, d->eload_off, S + 1);
  for (auto* v : {&hb.node_off, &hb.elem_off, &hb.con_off, &hb.nload_off, &hb.eload_off}) {
    if ((*v)[0] != 0) return fail(c, OBGPU_EINVAL, "offset arrays must start at 0");
    for (int64_t s = 0; s < S; s++)
      if ((*v)[s + 1] < (*v)[s]) return fail(c, OBGPU_EINVAL, "offset arrays must be non-decreasing");
  }
  const int64_t N = hb.Ntot(), E = hb.Etot(), C = hb.Ctot(), NL = hb.NLtot(), EL = hb.ELtot();
  if (N < 1 || E < 1) return fail(c, OBGPU_EINVAL, "empty structure (ASSERT_(nDOFs > 0))");
  if (!d->node_xy || !d->elem_nodes || !d->elem_type || !d->elem_EAI) return fail(c, OBGPU_EINVAL, "null geometry array");
  if ((C && (!d->con_node || !d->con_dof || !d->con_val)) || (NL && (!d->nload_node || !d->nload_dof || !d->nload_val)) ||
      (EL && (!d->eload_elem || !d->eload_q || !d->eload_dir)))
    return fail(c, OBGPU_EINVAL, "null constraint/load array");
  copy_vec(hb.node_xy, d->node_xy, 2 * N);
  if (d->node_rotz) {
    copy_vec(hb.node_rotz, d->node_rotz, N);
    for (double r : hb.node_rotz)
      if (r != 0.0) hb.has_rotz = true;
  }
  copy_vec(hb.elem_nodes, d->elem_nodes, 2 * E);
  copy_vec(hb.elem_type, d->elem_type, E);
  copy_vec(hb.elem_EAI, d->elem_EAI, 3 * E);
  if (C) { copy_vec(hb.con_node, d->con_node, C); copy_vec(hb.con_dof, d->con_dof, C); copy_vec(hb.con_val, d->con_val, C); }
  if (NL) { copy_vec(hb.nload_node, d->nload_node, NL); copy_vec(hb.nload_dof, d->nload_dof, NL); copy_vec(hb.nload_val, d->nload_val, NL); }
  if (EL) {
    copy_vec(hb.eload_elem, d->eload_elem, EL); copy_vec(hb.eload_q, d->eload_q, EL); copy_vec(hb.eload_dir, d->eload_dir, 3 * EL);
    hb.eload_type.assign(EL, OBGPU_LOAD_UNIFORM);
    hb.eload_q2.assign(EL, 0.0);
    if (d->eload_type) copy_vec(hb.eload_type, d->eload_type, EL);
    if (d->eload_q2) copy_vec(hb.eload_q2, d->eload_q2, EL);
    for (int64_t i = 0; i < EL; i++) {
      const int lt = hb.eload_type[i];
      if (lt > OBGPU_LOAD_TEMPERATURE) return fail(c, OBGPU_EINVAL, "unknown element load type");
      if (lt != OBGPU_LOAD_UNIFORM) hb.has_load_kinds = true;
      if (lt == OBGPU_LOAD_TEMPERATURE) continue;  // no direction (CLoadConstTemperature.cpp)
      // ASSERT_NEAR_(|dir|^2, 1, 1e-6), CLoadDistributedUniform.cpp:37-38
      const double* v = &hb.eload_dir[3 * i];
      const double n2 = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
      if (!(n2 > 1 - 1e-6 && n2 < 1 + 1e-6)) return fail(c, OBGPU_EINVAL, "element load direction is not a unit vector");
    }
  }
  for (int64_t e = 0; e < E; e++) {  // UNINITIALIZED_VALUE checks, CElementBeam_2D_RR.cpp:64-91
    for (int k = 0; k < 3; k++)
      if (hb.elem_EAI[3 * e + k] == 1.7976931348623157e308) return fail(c, OBGPU_EINVAL, "Uninitialized parameter E/A/Iz");
    if (hb.elem_type[e] > OBGPU_SPRING_TORSION) return fail(c, OBGPU_EINVAL, "unknown element type");
  }
  // (load, element) pairs the reference throws on (CLoadDistributedTriangular.cpp:93-125,
  // CLoadConcentratedForce.cpp:77-84,114-116): refused here, at upload
  if (EL)
    for (int64_t s = 0; s < S; s++)
      for (int64_t l = hb.eload_off[s]; l < hb.eload_off[s + 1]; l++) {
        const int lt = hb.eload_type[l];
        const int64_t e = hb.elem_off[s] + hb.eload_elem[l];
        if (hb.eload_elem[l] >= uint32_t(hb.elem_off[s + 1] - hb.elem_off[s])) continue;  // reported by the symbolic phase
        if (lt == OBGPU_LOAD_UNIFORM && hb.elem_type[e] > OBGPU_BEAM2D_AA)   // CLoadDistributedUniform.cpp:85-125
          return fail(c, OBGPU_EINVAL, "Unsupported element type for load 'CLoadDistributedUniform'");
        if (lt == OBGPU_LOAD_TEMPERATURE && hb.elem_type[e] > OBGPU_BEAM2D_RD)  // CBaseElementBeam only (:40-49)
          return fail(c, OBGPU_EINVAL, "Unsupported element type for load 'CLoadConstTemperature'");
        if (lt != OBGPU_LOAD_TRIANGULAR && lt != OBGPU_LOAD_CONCENTRATED) continue;
        if (hb.elem_type[e] != OBGPU_BEAM2D_RR)
          return fail(c, OBGPU_EINVAL, lt == OBGPU_LOAD_TRIANGULAR
                                           ? "Unsupported element type for load 'CLoadDistributedTriangular'"
                                           : "Unsupported element type for load 'CLoadConcentratedForce'");
        if (lt == OBGPU_LOAD_CONCENTRATED) {
          const int64_t n0 = hb.node_off[s], Ns = hb.node_off[s + 1] - n0;
          if (hb.elem_nodes[2 * e] >= Ns || hb.elem_nodes[2 * e + 1] >= Ns) continue;  // reported by the symbolic phase
          const double* p0 = &hb.node_xy[2 * (n0 + hb.elem_nodes[2 * e])];
          const double* p1 = &hb.node_xy[2 * (n0 + hb.elem_nodes[2 * e + 1])];
          const double ax = p1[0] - p0[0], ay = p1[1] - p0[1], len = std::sqrt(ax * ax + ay * ay);
          const double plong = (ax * hb.eload_dir[3 * l] + ay * hb.eload_dir[3 * l + 1]) / len;
          if (std::fabs(plong) > 1e-5) return fail(c, OBGPU_EINVAL, "Only perpendicular forces supported by now!");
          if (hb.eload_q2[l] < 0 || len - hb.eload_q2[l] < 0)
            return fail(c, OBGPU_EINVAL, "Error: Point of application is out of the beam!");
        }
      }
  hb.L = d->n_cases < 1 ? 1 : d->n_cases;
  c->uploaded = true;
  c->sym = Symbolic();
  c->dev_batch = false;
  c->dev_plan = false;
  c->dev_mg = false;
  c->assembled = false;
  c->prog = TinyProgram();
  c->have_xy_len = false;
  return upload_cases(c, hb.L, d->case_nload_val, d->case_eload_q, d->case_elem_EAI);
}

int obgpu_update_cases(obgpu_ctx* c, int64_t n_cases, const double* nl, const double* q, const double* eai) {
  if (!c) return OBGPU_EINVAL;
  if (!c->uploaded) return fail(c, OBGPU_ESEQUENCE, "obgpu_update_cases before obgpu_batch_upload");
  OB_DEVICE(c);
  return upload_cases(c, n_cases, nl, q, eai);
}

int obgpu_symbolic(obgpu_ctx* c) {
  if (!c) return OBGPU_EINVAL;
  if (!c->uploaded) return fail(c, OBGPU_ESEQUENCE, "obgpu_symbolic before obgpu_batch_upload");
  OB_DEVICE(c);
  if (c->device_batch_mode && !c->host_batch_valid) return symbolic_device(c);
  const auto t0 = std::chrono::steady_clock::now();
  std::string err;
  c->dev_batch = false;
  c->dev_plan = false;
  c->dev_mg = false;
  c->assembled = false;
  // the K pattern costs O(T log T) host work: built eagerly for moderate batches, on demand otherwise
  int rc = run_symbolic(c->hb, c->sym, err, c->hb.Etot() <= (int64_t(1) << 22));
  if (rc != OBGPU_OK) return fail(c, rc, err);
  // ASSERT_(L > 0), CElementBeam_2D_RR.cpp:63 (beams; springs may join coincident nodes: identity orientation)
  for (int64_t s = 0; s < c->hb.S; s++)
    for (int64_t e = c->hb.elem_off[s]; e < c->hb.elem_off[s + 1]; e++) {
      if (c->hb.elem_type[e] > OBGPU_BEAM2D_RD) continue;
      const double* p0 = &c->hb.node_xy[2 * (c->hb.node_off[s] + c->hb.elem_nodes[2 * e])];
      const double* p1 = &c->hb.node_xy[2 * (c->hb.node_off[s] + c->hb.elem_nodes[2 * e + 1])];
      if (!((p1[0] - p0[0]) * (p1[0] - p0[0]) + (p1[1] - p0[1]) * (p1[1] - p0[1]) > 0))
        return fail(c, OBGPU_EINVAL, "zero-length element (ASSERT_(L > 0))");
    }
  c->prog = TinyProgram();
  const bool dev = c->device != OBGPU_DEVICE_NONE;
  if (c->hb.S == 1) {
    std::string perr;
    if (build_tiny_program(c->hb, c->sym, c->prog, perr) == OBGPU_OK && dev) {
      OB_CUDA(c, c->d_prog_i.upload(c->prog.ints, c->stream));
      OB_CUDA(c, c->d_prog_d.upload(c->prog.dbls, c->stream));
    }
  }
  if (dev) {
    std::vector<uint8_t> bdof(c->sym.bnd_idx.size());
    for (int64_t st = 0; st < c->hb.S; st++)
      for (int64_t i = c->sym.bnd_off[st]; i < c->sym.bnd_off[st + 1]; i++)
        bdof[i] = uint8_t(c->sym.dof_id[c->sym.dof_off[st] + c->sym.bnd_idx[i]]);
    OB_CUDA(c, c->d_bnd_dofid.upload(bdof, c->stream));
    c->h_bnd_dofid = bdof;
    OB_CUDA(c, c->d_free_off.upload(c->sym.free_off, c->stream));
    OB_CUDA(c, c->d_bnd_off.upload(c->sym.bnd_off, c->stream));
    OB_CUDA(c, c->d_Ub.upload(c->sym.U_b, c->stream));
    OB_CUDA(c, cudaStreamSynchronize(c->stream));
  }
  c->any_ub = false;
  for (double v : c->sym.U_b) c->any_ub |= v != 0.0;
  c->host_offsets_valid = true;
  c->t_symbolic_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
  return OBGPU_OK;
}

int obgpu_get_sizes(const obgpu_ctx* c, int64_t out[OBGPU_NSIZES]) {
  if (!c || !out || !c->uploaded) return OBGPU_EINVAL;
  const int64_t v[OBGPU_NSIZES] = {c->hb.S, c->L, c->hb.Ntot(), c->hb.Etot(), c->sym.n_tot(), c->sym.nf_tot(),
                                   c->sym.nb_tot(), c->sym.T_tot(), c->sym.nnz_tot(), c->hb.Ctot(), c->hb.NLtot(),
                                   c->hb.ELtot()};
  std::memcpy(out, v, sizeof(v));
  return OBGPU_OK;
}

int obgpu_get_offsets(const obgpu_ctx* c, int which, int64_t* out) {
  if (!c || !out || !c->sym.valid) return OBGPU_ESEQUENCE;
  if (!c->host_offsets_valid) {  // a device batch keeps its per-structure offsets in HBM until somebody asks
    const int rc = fetch_host_offsets(const_cast<obgpu_ctx*>(c));
    if (rc != OBGPU_OK) return rc;
  }
  const std::vector<int64_t>* v = nullptr;
  switch (which) {
    case OBGPU_OFF_DOF: v = &c->sym.dof_off; break;
    case OBGPU_OFF_FREE: v = &c->sym.free_off; break;
    case OBGPU_OFF_BOUNDED: v = &c->sym.bnd_off; break;
    case OBGPU_OFF_NNZ: v = &c->sym.nnz_off; break;
    case OBGPU_OFF_TRIPLET: v = &c->sym.tri_off; break;
    default: return OBGPU_EINVAL;
  }
  std::memcpy(out, v->data(), v->size() * sizeof(int64_t));
  return OBGPU_OK;
}

int obgpu_get_symbolic(const obgpu_ctx* cc, int which, int32_t* out) {
  if (!cc || !out || !cc->sym.valid) return OBGPU_ESEQUENCE;
  if (cc->device_batch_mode && !cc->host_batch_valid) {  // a device batch: bring it to the host once, run the host phase
    obgpu_ctx* m = const_cast<obgpu_ctx*>(cc);
    int rc = materialize_host_batch(m);
    if (rc == OBGPU_OK) rc = obgpu_symbolic(m);
    if (rc != OBGPU_OK) return rc;
  }
  const obgpu_ctx* c = cc;
  if (which == OBGPU_SYM_KFF_ROWPTR || which == OBGPU_SYM_KFF_COL)
    return symbolic_general(const_cast<obgpu_ctx*>(cc), which, out, nullptr);
  const std::vector<int32_t>* v = nullptr;
  switch (which) {
    case OBGPU_SYM_DOF_NODE: v = &c->sym.dof_node; break;
    case OBGPU_SYM_DOF_ID: v = &c->sym.dof_id; break;
    case OBGPU_SYM_FREE: v = &c->sym.free_idx; break;
    case OBGPU_SYM_BOUNDED: v = &c->sym.bnd_idx; break;
    case OBGPU_SYM_DOF_INDEX: v = &c->sym.dof_index; break;
    case OBGPU_SYM_K_ROWPTR: v = &c->sym.k_rowptr; break;
    case OBGPU_SYM_K_COL: v = &c->sym.k_col; break;
    case OBGPU_SYM_TRI_NNZ: v = &c->sym.tri_nnz; break;
    default: return OBGPU_EINVAL;
  }
  if (!v->empty()) std::memcpy(out, v->data(), v->size() * sizeof(int32_t));
  return OBGPU_OK;
}

int obgpu_get_stream(obgpu_ctx* c, void** s) {
  if (!c || !s) return OBGPU_EINVAL;
  OB_NEED_DEVICE(c);
  *s = (void*)c->stream;
  return OBGPU_OK;
}

int obgpu_synchronize(obgpu_ctx* c) {
  if (!c) return OBGPU_EINVAL;
  OB_NEED_DEVICE(c);
  OB_CUDA(c, cudaStreamSynchronize(c->stream));
  for (auto s : c->slot_stream) OB_CUDA(c, cudaStreamSynchronize(s));
  return OBGPU_OK;
}

}  // extern "C"
