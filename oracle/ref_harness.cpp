// ref_harness.cpp — the oracle's C ABI (ob_oracle.h) implemented ON TOP OF THE REFERENCE'S OWN, UNMODIFIED SOURCES.
//
// oracle/Makefile.ref compiles /root/reference/libopenbeam/src/{CFiniteElementProblem,…_assemble,…_solver_static,
// …_postproc_stress,CStructureProblem,CStructureProblem_mesh,CElement*,CBaseElementBeam,CLoad*,loads,internals,globals,
// types}.cpp where they lie, against the header stand-ins in oracle/ref_shim/ (Eigen and MRPT are not in this image:
// Eigen/ob_eigen_shim.h and mrpt/* restate exactly the subset those files use), and links them with this file into
// oracle/_ref/libob_ref.so.  Nothing of the reference is copied into the repo.  This file only drives the reference's
// public API the way apps/ob-solve and examples-code/example1.cpp do:
//   setNumberOfNodes / setNodePose, insertElement, updateAll, getDOFIndex + insertConstraint (parse_file.cpp:653-662),
//   addLoadAtDOF (parse_file.cpp:747-756), addLoadAtBeam, assembleProblem / solveStatic / postProcCalcStress,
// and copies the public result structs out.  TEST INFRASTRUCTURE: it pins oracle/ (the restatement) to the reference.
#include <openbeam/CStructureProblem.h>
#include <openbeam/elements.h>
#include <openbeam/loads.h>

#include <atomic>
#include <chrono>
#include <cstring>
#include <thread>

#include "ob_oracle.h"

using namespace openbeam;

namespace {

struct Spec {  // the structure as handed to obo_create
  int64_t N = 0, E = 0;
  std::vector<double> xy, rotz, eai, con_val, nl_val, el_q, el_dir, el_q2;
  bool has_rot = false;
  std::vector<uint32_t> en, con_node, nl_node, el_elem;
  std::vector<uint8_t> et, con_dof, nl_dof, el_kind;
};

CElement::Ptr make_element(int type, size_t a, size_t b, const double* p) {
  switch (type) {
    case OBO_RR: { auto e = std::make_shared<CElementBeam_2D_RR>(a, b); e->E = p[0]; e->A = p[1]; e->Iz = p[2]; return e; }
    case OBO_RA: { auto e = std::make_shared<CElementBeam_2D_RA>(a, b); e->E = p[0]; e->A = p[1]; e->Iz = p[2]; return e; }
    case OBO_AR: { auto e = std::make_shared<CElementBeam_2D_AR>(a, b); e->E = p[0]; e->A = p[1]; e->Iz = p[2]; return e; }
    case OBO_AA: { auto e = std::make_shared<CElementBeam_2D_AA>(a, b); e->E = p[0]; e->A = p[1]; e->Iz = p[2]; return e; }
    case OBO_RD: { auto e = std::make_shared<CElementBeam_2D_RD>(a, b); e->E = p[0]; e->A = p[1]; e->Iz = p[2]; return e; }
    case OBO_SPRING_1D: return std::make_shared<CElementSpring>(a, b, p[0]);
    case OBO_SPRING_XY: return std::make_shared<CElementSpringXY>(a, b, p[0], p[1]);
    case OBO_SPRING_DXDYRZ: return std::make_shared<CElementSpringDXDYRZ>(a, b, p[0], p[1], p[2]);
    case OBO_SPRING_TORSION: return std::make_shared<CElementTorsionSpring>(a, b, p[0]);
  }
  return nullptr;
}

void set_element_params(CElement* el, int type, const double* p) {
  if (type <= OBO_RD) {
    auto* b = dynamic_cast<CBaseElementBeam*>(el);
    b->E = p[0]; b->A = p[1]; b->Iz = p[2];
  } else if (type == OBO_SPRING_1D) dynamic_cast<CElementSpring*>(el)->K = p[0];
  else if (type == OBO_SPRING_XY) { auto* s = dynamic_cast<CElementSpringXY*>(el); s->Kx = p[0]; s->Ky = p[1]; }
  else if (type == OBO_SPRING_DXDYRZ) { auto* s = dynamic_cast<CElementSpringDXDYRZ*>(el); s->Kx = p[0]; s->Ky = p[1]; s->KRz = p[2]; }
  else if (type == OBO_SPRING_TORSION) dynamic_cast<CElementTorsionSpring*>(el)->K = p[0];
}

struct Built {  // a live reference problem + handles to the mutable load objects
  CStructureProblem P;
  std::vector<CLoadOnBeam::Ptr> eloads;
  std::vector<size_t> nl_dof_index;  // npos when the DOF does not exist
  StaticSolveProblemInfo sinfo;
  BuildProblemInfo binfo;            // of a bare assembleProblem() call
  StressInfo stress;
  bool assembled = false, solved = false, stressed = false;
};

std::unique_ptr<Built> build(const Spec& s) {
  auto B = std::make_unique<Built>();
  CStructureProblem& P = B->P;
  P.setNumberOfNodes(size_t(s.N));
  for (int64_t i = 0; i < s.N; i++)  // parse_file.cpp:457-477: TRotationTrans3D(x, y, z, rot_x, rot_y, rot_z)
    P.setNodePose(size_t(i), TRotationTrans3D(s.xy[2 * i], s.xy[2 * i + 1], 0, 0, 0, s.has_rot ? s.rotz[i] : 0.0));
  for (int64_t e = 0; e < s.E; e++) {
    auto el = make_element(s.et[e], s.en[2 * e], s.en[2 * e + 1], &s.eai[3 * e]);
    if (!el) return nullptr;
    P.insertElement(el);
  }
  P.updateAll();
  for (size_t i = 0; i < s.con_node.size(); i++) {  // parse_file.cpp:653-662: constraints on absent DOFs are skipped
    const size_t k = P.getDOFIndex(s.con_node[i], DoF_index(s.con_dof[i]));
    if (k != std::string::npos) P.insertConstraint(k, s.con_val[i]);
  }
  B->nl_dof_index.resize(s.nl_node.size());
  for (size_t i = 0; i < s.nl_node.size(); i++) {   // parse_file.cpp:747-756
    const size_t k = P.getDOFIndex(s.nl_node[i], DoF_index(s.nl_dof[i]));
    B->nl_dof_index[i] = k;
    if (k != std::string::npos) P.addLoadAtDOF(k, s.nl_val[i]);
  }
  for (size_t i = 0; i < s.el_elem.size(); i++) {
    const double* d = &s.el_dir[3 * i];
    CLoadOnBeam::Ptr l;
    const int kind = s.el_kind.empty() ? 0 : s.el_kind[i];
    const double q2 = s.el_q2.empty() ? 0.0 : s.el_q2[i];
    if (kind == 0) l = std::make_shared<CLoadDistributedUniform>(s.el_q[i], d[0], d[1], d[2]);
    else if (kind == 1) l = std::make_shared<CLoadDistributedTriangular>(s.el_q[i], q2, d[0], d[1], d[2]);
    else if (kind == 2) l = std::make_shared<CLoadConcentratedForce>(s.el_q[i], q2, d[0], d[1], d[2]);
    else if (kind == 3) l = std::make_shared<CLoadConstTemperature>(s.el_q[i]);
    else return nullptr;
    B->eloads.push_back(l);
    P.addLoadAtBeam(s.el_elem[i], l);
  }
  return B;
}

void apply_values(const Spec& s, Built& B) {  // load values and sections of the spec -> the live problem
  // nodal loads: several entries may hit one DOF (addLoadAtDOF accumulates): reset, then add in input order
  for (size_t i = 0; i < s.nl_node.size(); i++)
    if (B.nl_dof_index[i] != std::string::npos) B.P.setLoadAtDOF(B.nl_dof_index[i], 0);
  for (size_t i = 0; i < s.nl_node.size(); i++)
    if (B.nl_dof_index[i] != std::string::npos) B.P.addLoadAtDOF(B.nl_dof_index[i], s.nl_val[i]);
  for (size_t i = 0; i < B.eloads.size(); i++) {
    const int kind = s.el_kind.empty() ? 0 : s.el_kind[i];
    if (kind == 0) static_cast<CLoadDistributedUniform*>(B.eloads[i].get())->q = s.el_q[i];
    else if (kind == 1) static_cast<CLoadDistributedTriangular*>(B.eloads[i].get())->q_ini = s.el_q[i];
    else if (kind == 2) static_cast<CLoadConcentratedForce*>(B.eloads[i].get())->P = s.el_q[i];
    else static_cast<CLoadConstTemperature*>(B.eloads[i].get())->m_incr_temp = s.el_q[i];
  }
  for (int64_t e = 0; e < s.E; e++) set_element_params(B.P.getElement(size_t(e)).get(), s.et[e], &s.eai[3 * e]);
}

}  // namespace

// ExpressionEvaluator.cpp (the parser's exprtk front end, out of the path) is not part of this build: the parameter loaders
// that reference it only ever see numeric literals here.
num_t openbeam::EvaluationContext::evaluate(const std::string& sVarVal) const { return std::strtod(sVarVal.c_str(), nullptr); }

struct obo_problem {
  Spec spec;
  std::unique_ptr<Built> B;
  std::string err;
  bool dirty = true;
};

static Built* live(obo_problem* h) {
  if (!h->B) {
    h->B = build(h->spec);
    h->dirty = false;
  } else if (h->dirty) {
    apply_values(h->spec, *h->B);
    h->dirty = false;
  }
  return h->B.get();
}

template <class T, class F>
static int64_t emit(int64_t n, T* out, int64_t cap, F f) {
  if (out)
    for (int64_t i = 0; i < n && i < cap; i++) out[i] = f(i);
  return n;
}

extern "C" {

obo_problem* obo_create(int64_t n_nodes, const double* node_xy, const double* node_rotz, int64_t n_elems,
                        const uint32_t* elem_nodes, const uint8_t* elem_type, const double* elem_EAI, int64_t n_cons,
                        const uint32_t* con_node, const uint8_t* con_dof, const double* con_val, int64_t n_nloads,
                        const uint32_t* nl_node, const uint8_t* nl_dof, const double* nl_val, int64_t n_eloads,
                        const uint32_t* el_elem, const double* el_q, const double* el_dir) {
  openbeam::setVerbosityLevel(0);   // the library's OB_MESSAGE chatter (std::cout) has no place in a test harness
  auto* h = new obo_problem;
  Spec& s = h->spec;
  s.N = n_nodes; s.E = n_elems;
  s.xy.assign(node_xy, node_xy + 2 * n_nodes);
  s.has_rot = node_rotz != nullptr;
  if (node_rotz) s.rotz.assign(node_rotz, node_rotz + n_nodes);
  s.en.assign(elem_nodes, elem_nodes + 2 * n_elems);
  s.et.assign(elem_type, elem_type + n_elems);
  s.eai.assign(elem_EAI, elem_EAI + 3 * n_elems);
  if (n_cons) { s.con_node.assign(con_node, con_node + n_cons); s.con_dof.assign(con_dof, con_dof + n_cons); s.con_val.assign(con_val, con_val + n_cons); }
  if (n_nloads) { s.nl_node.assign(nl_node, nl_node + n_nloads); s.nl_dof.assign(nl_dof, nl_dof + n_nloads); s.nl_val.assign(nl_val, nl_val + n_nloads); }
  if (n_eloads) { s.el_elem.assign(el_elem, el_elem + n_eloads); s.el_q.assign(el_q, el_q + n_eloads); s.el_dir.assign(el_dir, el_dir + 3 * n_eloads); }
  for (int64_t e = 0; e < n_elems; e++)
    if (elem_type[e] > OBO_SPRING_TORSION) { delete h; return nullptr; }
  return h;
}

void obo_destroy(obo_problem* h) { delete h; }

int obo_set_eload_kinds(obo_problem* h, const uint8_t* kind, const double* q2) {
  const size_t n = h->spec.el_elem.size();
  for (size_t i = 0; i < n; i++)
    if (kind[i] > 3) return -1;
  h->spec.el_kind.assign(kind, kind + n);
  h->spec.el_q2.assign(q2, q2 + n);
  h->B.reset();
  return 0;
}

int obo_set_load_values(obo_problem* h, const double* nl_val, const double* el_q) {
  if (nl_val) h->spec.nl_val.assign(nl_val, nl_val + h->spec.nl_node.size());
  if (el_q) h->spec.el_q.assign(el_q, el_q + h->spec.el_elem.size());
  h->dirty = true;
  return 0;
}

int obo_set_elem_EAI(obo_problem* h, const double* eai) {
  h->spec.eai.assign(eai, eai + 3 * h->spec.E);
  h->dirty = true;
  return 0;
}

int obo_assemble(obo_problem* h, int /*skip_subK: the reference always extracts the sub-matrices*/) {
  try {
    Built* B = live(h);
    if (!B) { h->err = "bad element / load type"; return -1; }
    B->P.assembleProblem(B->binfo);
    B->assembled = true;
    B->solved = false;
    return 0;
  } catch (const std::exception& e) {
    h->err = e.what();
    return -1;
  }
}

int obo_solve_static(obo_problem* h, int algorithm) {
  try {
    Built* B = live(h);
    if (!B) { h->err = "bad element / load type"; return -1; }
    StaticSolverOptions opts;     // defaults: Dense_LLT, linear (CFiniteElementProblem.h:165-171)
    opts.nonLinearIterative = (algorithm & 2) != 0;   // ob_oracle.h: algorithm bit 1
    B->P.solveStatic(B->sinfo, opts);
    B->binfo = B->sinfo.build_info;
    B->assembled = B->solved = true;
    B->stressed = false;
    return 0;
  } catch (const std::exception& e) {
    h->err = e.what();
    return -1;
  }
}

int obo_stress(obo_problem* h) {
  try {
    Built* B = live(h);
    if (!B || !B->solved) { h->err = "stress before solve"; return -1; }
    B->P.postProcCalcStress(B->stress, B->sinfo);
    B->stressed = true;
    return 0;
  } catch (const std::exception& e) {
    h->err = e.what();
    return -1;
  }
}

const char* obo_last_error(const obo_problem* h) { return h->err.c_str(); }

void obo_get_sizes(const obo_problem* hc, int64_t out[8]) {
  auto* h = const_cast<obo_problem*>(hc);
  for (int i = 0; i < 8; i++) out[i] = 0;
  Built* B = live(h);
  if (!B) return;
  out[0] = h->spec.N; out[1] = h->spec.E;
  out[2] = int64_t(B->P.getProblemDoFs().size());
  if (B->assembled) { out[3] = int64_t(B->binfo.free_dof_indices.size()); out[4] = int64_t(B->binfo.bounded_dof_indices.size()); }
  out[5] = -1; out[6] = -1;  // triplets and the full K are locals of assembleProblem(): not observable through the reference's API
}

int64_t obo_get_i64(const obo_problem* hc, int which, int64_t* out, int64_t cap) {
  auto* h = const_cast<obo_problem*>(hc);
  Built* B = live(h);
  if (!B) return -1;
  const auto& dofs = B->P.getProblemDoFs();
  switch (which) {
    case OBO_I_DOF_NODE: return emit(int64_t(dofs.size()), out, cap, [&](int64_t i) { return int64_t(dofs[size_t(i)].nodeId); });
    case OBO_I_DOF_ID: return emit(int64_t(dofs.size()), out, cap, [&](int64_t i) { return int64_t(dofs[size_t(i)].dofAsInt()); });
    case OBO_I_FREE: return emit(int64_t(B->binfo.free_dof_indices.size()), out, cap, [&](int64_t i) { return int64_t(B->binfo.free_dof_indices[size_t(i)]); });
    case OBO_I_BOUNDED: return emit(int64_t(B->binfo.bounded_dof_indices.size()), out, cap, [&](int64_t i) { return int64_t(B->binfo.bounded_dof_indices[size_t(i)]); });
    case OBO_I_DOF_INDEX:
      return emit(h->spec.N * 6, out, cap, [&](int64_t i) {
        const size_t k = B->P.getDOFIndex(size_t(i / 6), DoF_index(i % 6));
        return k == std::string::npos ? int64_t(-1) : int64_t(k);
      });
    default: return -1;  // triplets / full-K pattern: locals of assembleProblem()
  }
}

int64_t obo_get_f64(const obo_problem* hc, int which, double* out, int64_t cap) {
  auto* h = const_cast<obo_problem*>(hc);
  Built* B = live(h);
  if (!B) return -1;
  const BuildProblemInfo& bi = B->binfo;
  auto dense = [&](const Eigen::SparseMatrix<num_t>& M) {
    const int64_t r = M.rows(), c = M.cols();
    return emit(r * c, out, cap, [&](int64_t i) { return M.coeff(int(i % r), int(i / r)); });  // col-major
  };
  switch (which) {
    case OBO_F_KFF_DENSE: return dense(bi.K_ff);
    case OBO_F_KBF_DENSE: return dense(bi.K_bf);
    case OBO_F_KBB_DENSE: return dense(bi.K_bb);
    case OBO_F_UB: return emit(int64_t(bi.U_b.size()), out, cap, [&](int64_t i) { return bi.U_b[i]; });
    case OBO_F_FF: return emit(int64_t(bi.F_f.size()), out, cap, [&](int64_t i) { return bi.F_f[i]; });
    case OBO_F_UF: return emit(int64_t(B->sinfo.U_f.size()), out, cap, [&](int64_t i) { return B->sinfo.U_f[i]; });
    case OBO_F_FB: return emit(int64_t(B->sinfo.F_b.size()), out, cap, [&](int64_t i) { return B->sinfo.F_b[i]; });
    case OBO_F_STRESS:
      return emit(h->spec.E * 12, out, cap, [&](int64_t i) {
        const FaceStress& f = B->stress.element_stress[size_t(i / 12)][size_t((i % 12) / 6)];
        const double v[6] = {f.N, f.Vy, f.Vz, f.Mx, f.My, f.Mz};
        return v[i % 6];
      });
    case OBO_F_ELEM_R:
      return emit(h->spec.E * 9, out, cap, [&](int64_t i) {
        const Matrix33& R = B->P.getElement(size_t(i / 9))->getGlobalOrientation().getRot();
        return R((i % 9) % 3, (i % 9) / 3);
      });
    case OBO_F_ELEM_KGLOBAL: {
      std::vector<double> v(size_t(h->spec.E) * 108);
      for (int64_t e = 0; e < h->spec.E; e++) {
        std::vector<TStiffnessSubmatrix> mats;
        B->P.getElement(size_t(e))->getGlobalStiffnessMatrices(mats);
        for (size_t m = 0; m < mats.size() && m < 3; m++)
          for (int j = 0; j < 6; j++)
            for (int i = 0; i < 6; i++) v[size_t(e) * 108 + m * 36 + size_t(j) * 6 + size_t(i)] = mats[m].matrix(i, j);
      }
      return emit(int64_t(v.size()), out, cap, [&](int64_t i) { return v[size_t(i)]; });
    }
    default: return -1;
  }
}

double obo_run_cases(const obo_problem* tmpl, int64_t n_cases, const double* nl_val, const double* el_q, const double* eai,
                     int n_threads, double* U_f, double* F_b, double* stress) {
  if (n_threads < 1) n_threads = 1;
  const Spec& base = tmpl->spec;
  const size_t NL = base.nl_node.size(), EL = base.el_elem.size();
  const int64_t E = base.E;
  std::atomic<int> failed{0};
  std::vector<std::unique_ptr<Built>> per;
  per.resize(size_t(n_threads));
  std::vector<Spec> specs(size_t(n_threads), base);
  for (auto& b : per) {
    b = build(base);
    if (!b) return -1.0;
  }
  const auto t0 = std::chrono::steady_clock::now();
  std::vector<std::thread> th;
  for (int t = 0; t < n_threads; t++)
    th.emplace_back([&, t] {
      Built& B = *per[size_t(t)];
      Spec& s = specs[size_t(t)];
      try {
        for (int64_t c = t; c < n_cases; c += n_threads) {
          if (nl_val) s.nl_val.assign(nl_val + c * NL, nl_val + (c + 1) * NL);
          if (el_q) s.el_q.assign(el_q + c * EL, el_q + (c + 1) * EL);
          if (eai) s.eai.assign(eai + c * 3 * E, eai + (c + 1) * 3 * E);
          apply_values(s, B);
          B.P.solveStatic(B.sinfo);
          B.P.postProcCalcStress(B.stress, B.sinfo);
          const int64_t nf = B.sinfo.U_f.size(), nb = B.sinfo.F_b.size();
          if (U_f) for (int64_t i = 0; i < nf; i++) U_f[c * nf + i] = B.sinfo.U_f[i];
          if (F_b) for (int64_t i = 0; i < nb; i++) F_b[c * nb + i] = B.sinfo.F_b[i];
          if (stress)
            for (int64_t e = 0; e < E; e++)
              for (int f = 0; f < 2; f++) {
                const FaceStress& fs = B.stress.element_stress[size_t(e)][size_t(f)];
                double* o = stress + ((c * E + e) * 2 + f) * 6;
                o[0] = fs.N; o[1] = fs.Vy; o[2] = fs.Vz; o[3] = fs.Mx; o[4] = fs.My; o[5] = fs.Mz;
              }
        }
      } catch (...) {
        failed = 1;
      }
    });
  for (auto& x : th) x.join();
  const double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  return failed ? -1.0 : dt;
}

}  // extern "C"

#include "ob_batch.ipp"
