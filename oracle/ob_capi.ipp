// ob_capi.ipp — extern "C" surface of the CPU oracle (test infrastructure; see ob_oracle.h).
struct obo_problem {
  std::unique_ptr<obo::Problem> p;
  std::string err;
};

template <class V, class O>
static int64_t copy_out(const V& v, O* out, int64_t cap) {
  const int64_t n = int64_t(v.size());
  if (out)
    for (int64_t i = 0; i < n && i < cap; i++) out[i] = O(v[i]);
  return n;
}

extern "C" {

obo_problem* obo_create(int64_t n_nodes, const double* node_xy, const double* node_rotz, int64_t n_elems,
                        const uint32_t* elem_nodes, const uint8_t* elem_type, const double* elem_EAI, int64_t n_cons,
                        const uint32_t* con_node, const uint8_t* con_dof, const double* con_val, int64_t n_nloads,
                        const uint32_t* nl_node, const uint8_t* nl_dof, const double* nl_val, int64_t n_eloads,
                        const uint32_t* el_elem, const double* el_q, const double* el_dir) {
  using namespace obo;
  auto* h = new obo_problem;
  h->p = std::make_unique<Problem>();
  Problem& P = *h->p;
  for (int64_t i = 0; i < n_nodes; i++) {
    TRotationTrans3D pose;
    pose.t[0] = node_xy[2 * i];
    pose.t[1] = node_xy[2 * i + 1];
    pose.t[2] = 0;
    // parse_file.cpp:457-477: TRotationTrans3D(x,y,z, rot_x,rot_y,rot_z) -> TRotation3D(roll,pitch,yaw)
    pose.r = TRotation3D(0, 0, node_rotz ? node_rotz[i] : 0.0);
    P.m_node_poses.push_back(pose);
  }
  for (int64_t e = 0; e < n_elems; e++) {
    auto el = createElementByType(elem_type[e]);
    if (!el) {
      delete h;
      return nullptr;
    }
    el->type = elem_type[e];
    el->conected_nodes_ids = {size_t(elem_nodes[2 * e]), size_t(elem_nodes[2 * e + 1])};
    el->E = elem_EAI[3 * e];
    el->A = elem_EAI[3 * e + 1];
    el->Iz = elem_EAI[3 * e + 2];
    el->m_parent = &P;
    P.m_elements.push_back(el);
  }
  for (int64_t i = 0; i < n_cons; i++) {
    P.raw_constraints_nd.push_back({size_t(con_node[i]), size_t(con_dof[i])});
    P.raw_constraints_val.push_back(con_val[i]);
  }
  for (int64_t i = 0; i < n_nloads; i++) {
    P.raw_nloads_nd.push_back({size_t(nl_node[i]), size_t(nl_dof[i])});
    P.raw_nloads_val.push_back(nl_val[i]);
  }
  for (int64_t i = 0; i < n_eloads; i++) {
    ElemLoad l;
    l.q = el_q[i];
    l.dir[0] = el_dir[3 * i];
    l.dir[1] = el_dir[3 * i + 1];
    l.dir[2] = el_dir[3 * i + 2];
    P.m_loads_on_beams.insert({size_t(el_elem[i]), l});
    P.raw_eload_elem.push_back(size_t(el_elem[i]));
  }
  return h;
}

void obo_destroy(obo_problem* h) { delete h; }

int obo_set_eload_kinds(obo_problem* h, const uint8_t* kind, const double* q2) {
  obo::Problem& P = *h->p;
  if (P.raw_eload_elem.empty()) return 0;
  // multimap iteration order = stable sort of insertion order by element index
  std::vector<size_t> order(P.raw_eload_elem.size());
  for (size_t i = 0; i < order.size(); i++) order[i] = i;
  std::stable_sort(order.begin(), order.end(), [&](size_t a, size_t b) { return P.raw_eload_elem[a] < P.raw_eload_elem[b]; });
  size_t k = 0;
  for (auto& kv : P.m_loads_on_beams) {
    const size_t i = order[k++];
    if (kind[i] > 3) return -1;
    kv.second.kind = kind[i];
    kv.second.q2 = q2 ? q2[i] : 0.0;
  }
  return 0;
}

static void set_loads(obo::Problem& P, const double* nl_val, const double* el_q) {
  if (nl_val)
    for (size_t i = 0; i < P.raw_nloads_val.size(); i++) P.raw_nloads_val[i] = nl_val[i];
  if (el_q && !P.raw_eload_elem.empty()) {
    // multimap iteration order = stable sort of insertion order by element index
    std::vector<size_t> order(P.raw_eload_elem.size());
    for (size_t i = 0; i < order.size(); i++) order[i] = i;
    std::stable_sort(order.begin(), order.end(),
                     [&](size_t a, size_t b) { return P.raw_eload_elem[a] < P.raw_eload_elem[b]; });
    size_t k = 0;
    for (auto& kv : P.m_loads_on_beams) kv.second.q = el_q[order[k++]];
  }
}
static void set_EAI(obo::Problem& P, const double* EAI) {
  for (size_t e = 0; e < P.m_elements.size(); e++) {
    P.m_elements[e]->E = EAI[3 * e];
    P.m_elements[e]->A = EAI[3 * e + 1];
    P.m_elements[e]->Iz = EAI[3 * e + 2];
  }
}

int obo_set_load_values(obo_problem* h, const double* nl_val, const double* el_q) {
  set_loads(*h->p, nl_val, el_q);
  return 0;
}
int obo_set_elem_EAI(obo_problem* h, const double* EAI) {
  set_EAI(*h->p, EAI);
  return 0;
}

#define OBO_TRY(stmt)          \
  try {                        \
    stmt;                      \
    h->err.clear();            \
    return 0;                  \
  } catch (std::exception & e) { \
    h->err = e.what();         \
    return 1;                  \
  }
int obo_assemble(obo_problem* h, int skip_subK) { OBO_TRY(h->p->assembleProblem(skip_subK != 0)) }
int obo_solve_static(obo_problem* h, int algorithm) { OBO_TRY(h->p->solveStatic(algorithm)) }
int obo_stress(obo_problem* h) { OBO_TRY(h->p->postProcCalcStress()) }
const char* obo_last_error(const obo_problem* h) { return h->err.c_str(); }

void obo_get_sizes(const obo_problem* h, int64_t out[8]) {
  const obo::Problem& P = *h->p;
  out[0] = int64_t(P.m_node_poses.size());
  out[1] = int64_t(P.m_elements.size());
  out[2] = int64_t(P.m_problem_DoFs.size());
  out[3] = int64_t(P.free_dof_indices.size());
  out[4] = int64_t(P.bounded_dof_indices.size());
  out[5] = int64_t(P.tri_r.size());
  out[6] = int64_t(P.K.val.size());
  out[7] = 0;
}


int64_t obo_get_i64(const obo_problem* h, int which, int64_t* out, int64_t cap) {
  const obo::Problem& P = *h->p;
  switch (which) {
    case OBO_I_DOF_NODE:
    case OBO_I_DOF_ID: {
      const int64_t n = int64_t(P.m_problem_DoFs.size());
      if (out)
        for (int64_t i = 0; i < n && i < cap; i++) out[i] = int64_t(P.m_problem_DoFs[i][which == OBO_I_DOF_NODE ? 0 : 1]);
      return n;
    }
    case OBO_I_FREE: return copy_out(P.free_dof_indices, out, cap);
    case OBO_I_BOUNDED: return copy_out(P.bounded_dof_indices, out, cap);
    case OBO_I_DOF_INDEX: {
      const int64_t n = int64_t(P.m_problem_DoFs_inverse_list.size()) * 6;
      if (out)
        for (int64_t i = 0; i < n && i < cap; i++) out[i] = P.m_problem_DoFs_inverse_list[i / 6][i % 6];
      return n;
    }
    case OBO_I_TRI_ROW: return copy_out(P.tri_r, out, cap);
    case OBO_I_TRI_COL: return copy_out(P.tri_c, out, cap);
    case OBO_I_K_OUTER: return copy_out(P.K.outer, out, cap);
    case OBO_I_K_INNER: return copy_out(P.K.inner, out, cap);
  }
  return -1;
}

int64_t obo_get_f64(const obo_problem* h, int which, double* out, int64_t cap) {
  const obo::Problem& P = *h->p;
  switch (which) {
    case OBO_F_TRI_VAL: return copy_out(P.tri_v, out, cap);
    case OBO_F_K_VAL: return copy_out(P.K.val, out, cap);
    case OBO_F_KFF_DENSE: return copy_out(P.K_ff, out, cap);
    case OBO_F_KBF_DENSE: return copy_out(P.K_bf, out, cap);
    case OBO_F_KBB_DENSE: return copy_out(P.K_bb, out, cap);
    case OBO_F_UB: return copy_out(P.U_b, out, cap);
    case OBO_F_FF: return copy_out(P.F_f, out, cap);
    case OBO_F_UF: return copy_out(P.U_f, out, cap);
    case OBO_F_FB: return copy_out(P.F_b, out, cap);
    case OBO_F_STRESS: {
      const int64_t n = int64_t(P.element_stress.size()) * 12;
      if (out)
        for (size_t e = 0; e < P.element_stress.size(); e++)
          for (size_t f = 0; f < 2 && f < P.element_stress[e].size(); f++) {
            const obo::FaceStress& s = P.element_stress[e][f];
            const double v[6] = {s.N, s.Vy, s.Vz, s.Mx, s.My, s.Mz};
            for (int k = 0; k < 6; k++) {
              const int64_t idx = int64_t(e) * 12 + int64_t(f) * 6 + k;
              if (idx < cap) out[idx] = v[k];
            }
          }
      return n;
    }
    case OBO_F_ELEM_R: {
      const int64_t n = int64_t(P.m_elements.size()) * 9;
      if (out)
        for (int64_t i = 0; i < n && i < cap; i++) out[i] = P.m_elements[i / 9]->m_global_orientation.rot.m[i % 9];
      return n;
    }
    case OBO_F_ELEM_KGLOBAL: {
      const int64_t n = int64_t(P.m_elements.size()) * 108;
      if (out)
        for (size_t e = 0; e < P.m_elements.size(); e++) {
          std::vector<obo::TStiffnessSubmatrix> mats;
          try {
            P.m_elements[e]->getGlobalStiffnessMatrices(mats);
          } catch (...) { return -1; }
          for (int j = 0; j < 3; j++)
            for (int k = 0; k < 36; k++) {
              const int64_t idx = int64_t(e) * 108 + j * 36 + k;
              if (idx < cap) out[idx] = mats[j].matrix.m[k];
            }
        }
      return n;
    }
  }
  return -1;
}

double obo_run_cases(const obo_problem* tmpl, int64_t n_cases, const double* nl_val, const double* el_q,
                     const double* elem_EAI, int n_threads, double* U_f, double* F_b, double* stress) {
  using namespace obo;
  if (n_threads < 1) n_threads = 1;
  const size_t nnl = tmpl->p->raw_nloads_val.size(), nel = tmpl->p->raw_eload_elem.size();
  const size_t nE = tmpl->p->m_elements.size();
  std::vector<std::unique_ptr<Problem>> clones;
  for (int t = 0; t < n_threads; t++) clones.push_back(tmpl->p->clone());
  std::vector<int> failed(n_threads, 0);
  const auto t0 = std::chrono::steady_clock::now();
  auto work = [&](int t) {
    Problem& P = *clones[t];
    const int64_t lo = n_cases * t / n_threads, hi = n_cases * (t + 1) / n_threads;
    try {
      for (int64_t c = lo; c < hi; c++) {
        set_loads(P, nl_val ? nl_val + c * nnl : nullptr, el_q ? el_q + c * nel : nullptr);
        if (elem_EAI) set_EAI(P, elem_EAI + c * 3 * nE);
        P.solveStatic(OBO_DENSE_LLT);  // ob-solve_main.cpp:348-349
        P.postProcCalcStress();        // ob-solve_main.cpp:889-891
        const size_t nf = P.U_f.size(), nb = P.F_b.size();
        if (U_f) std::memcpy(U_f + c * nf, P.U_f.data(), nf * sizeof(double));
        if (F_b) std::memcpy(F_b + c * nb, P.F_b.data(), nb * sizeof(double));
        if (stress)
          for (size_t e = 0; e < nE; e++)
            for (size_t f = 0; f < 2; f++) {
              const FaceStress& s = P.element_stress[e][f];
              double* o = stress + (size_t(c) * nE + e) * 12 + f * 6;
              o[0] = s.N; o[1] = s.Vy; o[2] = s.Vz; o[3] = s.Mx; o[4] = s.My; o[5] = s.Mz;
            }
      }
    } catch (...) { failed[t] = 1; }
  };
  if (n_threads == 1) {
    work(0);
  } else {
    std::vector<std::thread> th;
    for (int t = 0; t < n_threads; t++) th.emplace_back(work, t);
    for (auto& x : th) x.join();
  }
  const double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  for (int f : failed)
    if (f) return -1.0;
  return dt;
}

}  // extern "C"
