// ob_solve.ipp — part of the CPU oracle (test infrastructure; see ob_oracle.h).
// Restates src/CFiniteElementProblem_solver_static.cpp:29-219 (Dense_LLT route, linear, one iteration)
// and src/CFiniteElementProblem_postproc_stress.cpp:32-183.
namespace obo {

// Eigen::LLT<MatrixXd, Lower> unblocked kernel (Eigen/src/Cholesky/LLT.h, llt_inplace<Lower>::unblocked):
//   for k: x = A(k,k) - |A(k,0:k)|^2 ; if (x <= 0) return k ; A(k,k) = x = sqrt(x)
//          A(k+1:,k) -= A(k+1:,0:k) * A(k,0:k)^T ; A(k+1:,k) /= x
// Eigen switches to a blocked variant for size >= 32; the arithmetic differs only in rounding order.
// Third-party dependency: Eigen (header-only, not vendored, version unpinned — SURVEY.md §8c).
static int llt_inplace_lower(std::vector<num_t>& A, int n) {  // col-major
  for (int k = 0; k < n; k++) {
    num_t x = A[size_t(k) * n + k];
    num_t sq = 0;
    for (int j = 0; j < k; j++) sq += A[size_t(j) * n + k] * A[size_t(j) * n + k];
    x -= sq;
    if (x <= num_t(0)) return k;
    x = std::sqrt(x);
    A[size_t(k) * n + k] = x;
    const int rs = n - k - 1;
    if (k > 0 && rs > 0)
      for (int j = 0; j < k; j++) {
        const num_t akj = A[size_t(j) * n + k];
        const num_t* col = &A[size_t(j) * n + k + 1];
        num_t* dst = &A[size_t(k) * n + k + 1];
        for (int i = 0; i < rs; i++) dst[i] -= col[i] * akj;
      }
    if (rs > 0) {
      num_t* dst = &A[size_t(k) * n + k + 1];
      for (int i = 0; i < rs; i++) dst[i] /= x;
    }
  }
  return -1;
}
static void llt_solve(const std::vector<num_t>& Lm, int n, std::vector<num_t>& b) {
  for (int j = 0; j < n; j++) {  // L y = b (column-oriented forward substitution)
    b[j] /= Lm[size_t(j) * n + j];
    const num_t bj = b[j];
    for (int i = j + 1; i < n; i++) b[i] -= Lm[size_t(j) * n + i] * bj;
  }
  for (int j = n - 1; j >= 0; j--) {  // L^T x = y
    num_t s = b[j];
    for (int i = j + 1; i < n; i++) s -= Lm[size_t(j) * n + i] * b[i];
    b[j] = s / Lm[size_t(j) * n + j];
  }
}

// src/CFiniteElementProblem_solver_static.cpp:50-219 (+ solveStaticInternal :29-44).
// algorithm bit 1 (value 2) = StaticSolverOptions::nonLinearIterative (:55-58, :63-66, :139-193): 40 passes, each on the
// node poses moved by the previous pass's U_f; the poses are restored afterwards while the elements keep the LAST pass's
// cached orientation and the last pass's equivalent loads / fixed-end stresses (postProcCalcStress then reads lengths from
// the restored poses and orientations from the cache).
void Problem::solveStatic(int algorithm) {
  const bool nonLinearIterative = (algorithm & 2) != 0;
  algorithm &= ~2;
  if (algorithm != OBO_DENSE_LLT) throw std::runtime_error("Invalid solver algorithm value");
  std::deque<TRotationTrans3D> orig_nodes;
  if (nonLinearIterative) orig_nodes = m_node_poses;  // :57-58
  bool U_b_all_zeros = true;
  int nf = 0, nb = 0;
  const unsigned int nMaxIters = 40;  // :63
  unsigned int nIter = 0;
  while (nIter < nMaxIters) {
    nIter++;
    assembleProblem(false);  // :74
    U_b_all_zeros = true;    // :82
    for (num_t v : U_b)
      if (v != 0) U_b_all_zeros = false;
    nf = int(free_dof_indices.size());
    nb = int(bounded_dof_indices.size());

    std::vector<num_t> Kff = K_ff;  // :89 dense copy of the (lower-stored) sparse K_ff
    const int bad = llt_inplace_lower(Kff, nf);  // :97
    if (bad >= 0)
      throw std::runtime_error("[CFiniteElementProblem::solveStatic] Ill-conditioned matrix. Too few restrictions?");
    U_f = F_f;  // :35
    llt_solve(Kff, nf, U_f);
    if (!U_b_all_zeros) {  // :37-43  U_f -= Kff^-1 (Kbf^T U_b)
      std::vector<num_t> t(nf, 0);
      for (int c = 0; c < nf; c++) {
        num_t s = 0;
        for (int r = 0; r < nb; r++) s += K_bf[size_t(c) * nb + r] * U_b[r];
        t[c] = s;
      }
      llt_solve(Kff, nf, t);
      for (int c = 0; c < nf; c++) U_f[c] -= t[c];
    }
    if (!nonLinearIterative) break;  // :179-183
    for (size_t i = 0; i < free_dof_indices.size(); i++) {  // :146-176
      const num_t delta_val = U_f[i];
      const auto& D = m_problem_DoFs[free_dof_indices[i]];
      if (D[1] < 3) {
        m_node_poses[D[0]].t[D[1]] += delta_val;
      } else {
        const num_t rx = D[1] == 3 ? delta_val : 0, ry = D[1] == 4 ? delta_val : 0, rz = D[1] == 5 ? delta_val : 0;
        const Matrix33 inc = TRotation3D(rx, ry, rz).rot;
        const Matrix33 cur = m_node_poses[D[0]].r.rot;
        Matrix33 prod;
        for (int r = 0; r < 3; r++)
          for (int c = 0; c < 3; c++) {
            num_t acc = cur(r, 0) * inc(0, c);
            acc += cur(r, 1) * inc(1, c);
            acc += cur(r, 2) * inc(2, c);
            prod(r, c) = acc;
          }
        m_node_poses[D[0]].r.setRot(prod);
      }
    }
  }
  if (nonLinearIterative) m_node_poses = orig_nodes;  // :190-193
  // :196-200 reactions.  K_bb holds only r<=c (quirk Q1) and is used as a general matrix.
  F_b.assign(nb, 0);
  if (!U_b_all_zeros)
    for (int c = 0; c < nb; c++)
      for (int r = 0; r <= c; r++) F_b[r] += K_bb[size_t(c) * nb + r] * U_b[c];
  for (int c = 0; c < nf; c++) {
    const num_t u = U_f[c];
    for (int r = 0; r < nb; r++) F_b[r] += K_bf[size_t(c) * nb + r] * u;
  }
  for (const auto& kv : m_loads_at_each_dof_equivs) {  // :208-218
    const size_t idx_restricted = dof_types[kv.first].bounded_index;
    if (idx_restricted != NPOS) F_b[idx_restricted] -= kv.second;
  }
}

// src/CFiniteElementProblem_postproc_stress.cpp:32-183
void Problem::postProcCalcStress() {
  const size_t nE = m_elements.size();
  {
    std::vector<ElementStress> dumm;
    element_stress.swap(dumm);
  }
  element_stress.resize(nE);
  const size_t nNodes = m_node_poses.size();
  std::vector<std::array<num_t, 6>> U(nNodes);
  for (size_t n = 0; n < nNodes; n++)
    for (size_t k = 0; k < 6; k++)
      if (m_problem_DoFs_inverse_list[n][k] < 0) U[n][k] = 0;
  for (size_t i = 0; i < bounded_dof_indices.size(); i++) {
    const auto& D = m_problem_DoFs[bounded_dof_indices[i]];
    U[D[0]][D[1]] = U_b[i];
  }
  for (size_t i = 0; i < free_dof_indices.size(); i++) {
    const auto& D = m_problem_DoFs[free_dof_indices[i]];
    U[D[0]][D[1]] = U_f[i];
  }
  for (size_t i = 0; i < nE; i++) {
    auto el = m_elements[i];
    const size_t nFaces = el->conected_nodes_ids.size();
    std::vector<std::array<num_t, 6>> Uel(nFaces);
    const Matrix33& R = el->m_global_orientation.rot;
    for (size_t f = 0; f < nFaces; f++) {
      const auto& Uf = U[el->conected_nodes_ids[f]];
      for (int r = 0; r < 3; r++) {  // R^T * Uf(0:3)
        num_t s = R(0, r) * Uf[0];
        s += R(1, r) * Uf[1];
        s += R(2, r) * Uf[2];
        Uel[f][r] = s;
      }
      Uel[f][3] = Uf[3];  // :114-119 planar-only shortcut (quirk Q6)
      Uel[f][4] = 0;
      Uel[f][5] = Uf[5];
    }
    std::vector<TStiffnessSubmatrix> subMats;
    el->getLocalStiffnessMatrices(subMats);
    ElementStress& es = element_stress[i];
    es.assign(nFaces, FaceStress());
    for (size_t k = 0; k < subMats.size(); k++) {
      const TStiffnessSubmatrix& ss = subMats[k];
      num_t v[6];
      for (int r = 0; r < 6; r++) v[r] = 0;
      for (int c = 0; c < 6; c++)  // matrix * Uel[edge_out]  (column-wise gemv)
        for (int r = 0; r < 6; r++) v[r] += ss.matrix(r, c) * Uel[ss.edge_out][c];
      es[ss.edge_in].addVec(v);
      if (ss.edge_in != ss.edge_out) {
        for (int r = 0; r < 6; r++) {  // matrix^T * Uel[edge_in]
          num_t s = 0;
          for (int c = 0; c < 6; c++) s += ss.matrix(c, r) * Uel[ss.edge_in][c];
          v[r] = s;
        }
        es[ss.edge_out].addVec(v);
      }
    }
  }
  for (size_t i = 0; i < nE; i++) {  // :152-165
    ElementStress& es = element_stress[i];
    if (!es.empty()) {
      es[0].N = -es[0].N; es[0].Vy = -es[0].Vy; es[0].Vz = -es[0].Vz;
      es[0].Mx = -es[0].Mx; es[0].My = -es[0].My; es[0].Mz = -es[0].Mz;
    }
  }
  for (auto it = m_extra_stress_for_each_element.begin(); it != m_extra_stress_for_each_element.end(); ++it) {
    ElementStress& es = element_stress[it->first];  // :169-180
    const ElementStress& es2add = it->second;
    for (size_t i = 0; i < es.size(); ++i) es[i] += es2add[i];
  }
}

std::unique_ptr<Problem> Problem::clone() const {
  auto p = std::make_unique<Problem>();
  p->m_node_poses = m_node_poses;
  for (const auto& e : m_elements) {
    auto c = e->clone();
    c->m_parent = p.get();
    p->m_elements.push_back(c);
  }
  p->raw_constraints_nd = raw_constraints_nd;
  p->raw_constraints_val = raw_constraints_val;
  p->raw_nloads_nd = raw_nloads_nd;
  p->raw_nloads_val = raw_nloads_val;
  p->m_loads_on_beams = m_loads_on_beams;
  p->raw_eload_elem = raw_eload_elem;
  return p;
}

}  // namespace obo
