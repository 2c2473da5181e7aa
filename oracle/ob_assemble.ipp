// ob_assemble.ipp — part of the CPU oracle (test infrastructure; see ob_oracle.h).
// Restates src/CFiniteElementProblem.cpp:179-245, src/CFiniteElementProblem_assemble.cpp:37-437,
// src/CStructureProblem.cpp:59-113 and src/CLoadDistributedUniform.cpp:33-159.
namespace obo {

// src/CFiniteElementProblem.cpp:219-245
void Problem::updateNodeConnections() {
  const size_t nElements = m_elements.size();
  const size_t nNodes = m_node_poses.size();
  m_node_connections.assign(nNodes, TNodeConnections());
  for (size_t i = 0; i < nElements; i++) {
    const auto el = m_elements[i];
    const size_t nElementFaces = el->conected_nodes_ids.size();
    std::vector<used_DoFs_t> el_dofs;
    el->getGlobalDoFs(el_dofs);
    for (size_t k = 0; k < nElementFaces; k++) {
      if (!(el->conected_nodes_ids[k] < nNodes)) throw std::runtime_error("ASSERT_(node id < nNodes)");
      TNodeConnections& nc = m_node_connections[el->conected_nodes_ids[k]];
      nc[i].elementFaceId = static_cast<unsigned char>(k);
      nc[i].dofs = el_dofs[k];
    }
  }
}

// src/CFiniteElementProblem.cpp:179-216
void Problem::updateListDoFs() {
  updateNodeConnections();
  {
    std::vector<std::array<size_t, 2>> dum;
    m_problem_DoFs.swap(dum);
  }
  const size_t nNodes = m_node_poses.size();
  std::array<int, 6> unused;
  unused.fill(-1);
  m_problem_DoFs_inverse_list.assign(nNodes, unused);
  for (size_t nodeIdx = 0; nodeIdx < nNodes; nodeIdx++) {
    used_DoFs_t dofs;
    for (unsigned char d = 0; d < 6; d++) dofs[d] = false;
    for (const auto& nc : m_node_connections[nodeIdx])
      for (int d = 0; d < 6; d++)
        if (nc.second.dofs[d]) dofs[d] = true;
    for (uint8_t d = 0; d < 6; d++) {
      if (!dofs[d]) continue;
      const int new_dof_index = static_cast<int>(m_problem_DoFs.size());
      m_problem_DoFs.push_back({nodeIdx, size_t(d)});
      m_problem_DoFs_inverse_list[nodeIdx][d] = new_dof_index;
    }
  }
}

// What the parser does after updateAll(): parse_file.cpp:219-228,658-662,753 —
// (node,dof) -> getDOFIndex (CFiniteElementProblem.cpp:308-317); unknown DOFs are skipped;
// insertConstraint overwrites (CFiniteElementProblem.cpp:68-77), addLoadAtDOF accumulates (:92-99).
void Problem::applyConstraintsAndNodeLoads() {
  m_DoF_constraints.clear();
  m_loads_at_each_dof.clear();
  for (size_t i = 0; i < raw_constraints_nd.size(); i++) {
    const int idx = m_problem_DoFs_inverse_list.at(raw_constraints_nd[i][0])[raw_constraints_nd[i][1]];
    if (idx >= 0) m_DoF_constraints[size_t(idx)] = raw_constraints_val[i];
  }
  for (size_t i = 0; i < raw_nloads_nd.size(); i++) {
    const int idx = m_problem_DoFs_inverse_list.at(raw_nloads_nd[i][0])[raw_nloads_nd[i][1]];
    if (idx >= 0) m_loads_at_each_dof[size_t(idx)] += raw_nloads_val[i];
  }
}

// src/CLoadDistributedUniform.cpp:33-159
static void uniformLoad_computeStressAndEquivalentLoads(const ElemLoad& ld, const CElement* el, ElementStress& stress,
                                                        std::vector<array6>& loads) {
  const num_t* dir = ld.dir;
  const num_t q = ld.q;
  if (std::abs(dir[0] * dir[0] + dir[1] * dir[1] + dir[2] * dir[2] - 1.0) > 1e-6)
    throw std::runtime_error("ASSERT_NEAR_(|dir|^2, 1.0, 1e-6)");
  const TRotationTrans3D& node0 = el->m_parent->getNodePose(el->conected_nodes_ids[0]);
  const TRotationTrans3D& node1 = el->m_parent->getNodePose(el->conected_nodes_ids[1]);
  num_t Ax = node1.t[0] - node0.t[0];
  num_t Ay = node1.t[1] - node0.t[1];
  num_t Az = node1.t[2] - node0.t[2];
  const num_t L2 = Ax * Ax + Ay * Ay + Az * Az;
  const num_t L = std::sqrt(L2);
  if (!(L > 0)) throw std::runtime_error("ASSERT_(L > 0)");
  const num_t _1_L = 1 / L;
  Ax *= _1_L;
  Ay *= _1_L;
  Az *= _1_L;
  const num_t cross_prod[3] = {Ay * dir[2] - Az * dir[1], -Ax * dir[2] + Az * dir[0], Ax * dir[1] - Ay * dir[0]};
  const num_t proy_long = (Ax * dir[0] + Ay * dir[1] + Az * dir[2]);
  const num_t proy_perp = std::sqrt(1 - proy_long * proy_long) * (cross_prod[2] < 0 ? +1. : -1.);
  const num_t q_t = q * proy_long;
  const num_t q_n = q * proy_perp;
  num_t f1[3] = {0, 0, 0}, f2[3] = {0, 0, 0};
  num_t m1, m2;
  switch (el->type) {  // dynamic_cast chain order AA, RR, RA, AR (:85-123)
    case OBO_AA:
      f1[0] = -q_t * L * 0.5; f1[1] = q_n * L * 0.5; m1 = 0;
      f2[0] = -q_t * L * 0.5; f2[1] = q_n * L * 0.5; m2 = 0;
      break;
    case OBO_RR:
      f1[0] = -q_t * L * 0.5; f1[1] = q_n * L * 0.5; m1 = +q_n * L * L / 12;
      f2[0] = -q_t * L * 0.5; f2[1] = q_n * L * 0.5; m2 = -q_n * L * L / 12;
      break;
    case OBO_RA:
      f1[0] = -q_t * L * 0.5; f1[1] = q_n * L * (5. / 8); m1 = +q_n * L * L / 8;
      f2[0] = -q_t * L * 0.5; f2[1] = q_n * L * (3. / 8); m2 = 0;
      break;
    case OBO_AR:
      f1[0] = -q_t * L * 0.5; f1[1] = q_n * L * (3. / 8); m1 = 0;
      f2[0] = -q_t * L * 0.5; f2[1] = q_n * L * (5. / 8); m2 = -q_n * L * L / 8;
      break;
    default:
      throw std::runtime_error("Unsupported element type for load 'CLoadDistributedUniform'");
  }
  stress.assign(2, FaceStress());
  stress[0].N = -f1[0]; stress[0].Vy = -f1[1]; stress[0].Mz = -m1;
  stress[1].N = f2[0];  stress[1].Vy = f2[1];  stress[1].Mz = m2;
  const Matrix33& R = el->m_global_orientation.rot;
  num_t F1[3], F2[3];
  for (int r = 0; r < 3; r++) {  // F = -R * f   (negate R first, as Eigen's unary minus on R does)
    num_t a = (-R(r, 0)) * f1[0]; a += (-R(r, 1)) * f1[1]; a += (-R(r, 2)) * f1[2];
    num_t b = (-R(r, 0)) * f2[0]; b += (-R(r, 1)) * f2[1]; b += (-R(r, 2)) * f2[2];
    F1[r] = a; F2[r] = b;
  }
  loads.resize(2);
  loads[0] = {F1[0], F1[1], F1[2], 0, 0, -m1};
  loads[1] = {F2[0], F2[1], F2[2], 0, 0, -m2};
}

// src/CLoadDistributedTriangular.cpp:31-169, src/CLoadConcentratedForce.cpp:31-154 (both: RR beams only)
static void triOrPointLoad_computeStressAndEquivalentLoads(const ElemLoad& ld, const CElement* el, ElementStress& stress,
                                                           std::vector<array6>& loads) {
  const num_t* dir = ld.dir;
  if (!(std::abs(1 - (dir[0] * dir[0] + dir[1] * dir[1] + dir[2] * dir[2])) < 1e-6))
    throw std::runtime_error("ASSERT_(|1 - |dir|^2| < 1e-6)");
  const TRotationTrans3D& node0 = el->m_parent->getNodePose(el->conected_nodes_ids[0]);
  const TRotationTrans3D& node1 = el->m_parent->getNodePose(el->conected_nodes_ids[1]);
  num_t Ax = node1.t[0] - node0.t[0];
  num_t Ay = node1.t[1] - node0.t[1];
  num_t Az = node1.t[2] - node0.t[2];
  const num_t L2 = Ax * Ax + Ay * Ay + Az * Az;
  const num_t L = std::sqrt(L2);
  if (!(L > 0)) throw std::runtime_error("ASSERT_(L > 0)");
  const num_t _1_L = 1 / L;
  Ax *= _1_L;
  Ay *= _1_L;
  Az *= _1_L;
  const num_t cross_prod[3] = {Ay * dir[2] - Az * dir[1], -Ax * dir[2] + Az * dir[0], Ax * dir[1] - Ay * dir[0]};
  const num_t proy_long = (Ax * dir[0] + Ay * dir[1] + Az * dir[2]);
  const num_t proy_perp = std::sqrt(1 - proy_long * proy_long) * (cross_prod[2] < 0 ? +1. : -1.);
  num_t f1[3] = {0, 0, 0}, f2[3] = {0, 0, 0};
  num_t m1, m2;
  if (ld.kind == OBO_LOAD_TRIANGULAR) {  // CLoadDistributedTriangular.cpp:72-130
    const num_t q_ini = ld.q, q_end = ld.q2;
    const num_t q0_t = q_ini * proy_long, q1_t = q_end * proy_long;
    const num_t q0_n = q_ini * proy_perp, q1_n = q_end * proy_perp;
    if (el->type == OBO_RR) {
      f1[1] = L * (q0_n * 0.5 + (q1_n - q0_n) * 3. / 20.);
      m1 = +L * L * (q0_n * 1. / 12. + (q1_n - q0_n) * 1. / 30.);
      f2[1] = L * (q0_n * 0.5 + (q1_n - q0_n) * 7. / 20.);
      m2 = -L * L * (q0_n * 1. / 12. + (q1_n - q0_n) * 1. / 20.);
    } else if (el->type == OBO_AA || el->type == OBO_RA || el->type == OBO_AR) {
      throw std::runtime_error("TODO");
    } else {
      throw std::runtime_error("Unsupported element type for load 'CLoadDistributedTriangular'");
    }
    f1[0] = -0.5 * L * (q0_t + (q1_t - q0_t) / 3.0);
    f2[0] = -0.5 * L * (q0_t + 2.0 * (q1_t - q0_t) / 3.0);
  } else {  // CLoadConcentratedForce.cpp:72-116
    const num_t P_n = ld.q * proy_perp;
    if (std::abs(proy_long) > 1e-5) throw std::runtime_error("Only perpendicular forces supported by now!");
    const num_t a = ld.q2;
    const num_t b = L - a;
    if (a < 0 || b < 0) throw std::runtime_error("Error: Point of application is out of the beam!");
    const num_t a2 = a * a;
    const num_t b2 = b * b;
    const num_t L3 = L * L2;
    if (el->type == OBO_RR) {
      f1[0] = 0;
      f1[1] = P_n * (b2 / L3) * (3 * a + b);
      m1 = P_n * a * b2 / L2;
      f2[0] = 0;
      f2[1] = P_n * (a2 / L3) * (a + 3 * b);
      m2 = -P_n * a2 * b / L2;
    } else {
      throw std::runtime_error("Unsupported element type for load 'CLoadConstTemperature'");  // sic (:114-116)
    }
  }
  stress.assign(2, FaceStress());
  stress[0].N = -f1[0]; stress[0].Vy = -f1[1]; stress[0].Mz = -m1;
  stress[1].N = f2[0];  stress[1].Vy = f2[1];  stress[1].Mz = m2;
  const Matrix33& R = el->m_global_orientation.rot;
  num_t F1[3], F2[3];
  for (int r = 0; r < 3; r++) {  // F = -R * f
    num_t a = (-R(r, 0)) * f1[0]; a += (-R(r, 1)) * f1[1]; a += (-R(r, 2)) * f1[2];
    num_t b = (-R(r, 0)) * f2[0]; b += (-R(r, 1)) * f2[1]; b += (-R(r, 2)) * f2[2];
    F1[r] = a; F2[r] = b;
  }
  loads.resize(2);
  loads[0] = {F1[0], F1[1], F1[2], 0, 0, -m1};
  loads[1] = {F2[0], F2[1], F2[2], 0, 0, -m2};
}

// src/CLoadConstTemperature.cpp:31-93 (alpha = 12e-6 is hard-wired there)
static void temperatureLoad_computeStressAndEquivalentLoads(const ElemLoad& ld, const CElement* el, ElementStress& stress,
                                                            std::vector<array6>& loads) {
  if (el->type > OBO_RD)  // dynamic_cast<const CBaseElementBeam*> fails for the springs (:40-49)
    throw std::runtime_error("Unsupported element type for load 'CLoadConstTemperature'");
  const num_t E = el->E, A = el->A, alpha = 12e-6;
  const num_t N = E * A * alpha * ld.q;
  stress.assign(2, FaceStress());
  stress[0].N = -N;
  stress[1].N = -N;
  loads.resize(2);
  for (int i = 0; i < 6; i++) {
    loads[0][i] = 0;
    loads[1][i] = 0;
  }
  const TRotationTrans3D& node0 = el->m_parent->getNodePose(el->conected_nodes_ids[0]);
  const TRotationTrans3D& node1 = el->m_parent->getNodePose(el->conected_nodes_ids[1]);
  num_t Ax = node1.t[0] - node0.t[0];
  num_t Ay = node1.t[1] - node0.t[1];
  num_t Az = node1.t[2] - node0.t[2];
  const num_t L2 = Ax * Ax + Ay * Ay + Az * Az;
  const num_t L = std::sqrt(L2);
  if (!(L > 0)) throw std::runtime_error("ASSERT_(L > 0)");
  const num_t _1_L = 1 / L;
  Ax *= _1_L;
  Ay *= _1_L;
  Az *= _1_L;
  loads[0][0] = -N * Ax; loads[0][1] = -N * Ay; loads[0][2] = -N * Az;
  loads[1][0] = N * Ax;  loads[1][1] = N * Ay;  loads[1][2] = N * Az;
}

// src/CStructureProblem.cpp:59-113
void Problem::internalComputeStressAndEquivalentLoads() {
  m_loads_at_each_dof_equivs.clear();
  m_extra_stress_for_each_element.clear();
  for (const auto& elIdxLoadPair : m_loads_on_beams) {
    auto el = m_elements.at(elIdxLoadPair.first);
    ElementStress stress;
    std::vector<array6> loads;
    const ElemLoad& ld = elIdxLoadPair.second;  // virtual CLoadOnBeam::computeStressAndEquivalentLoads
    if (ld.kind == OBO_LOAD_UNIFORM) uniformLoad_computeStressAndEquivalentLoads(ld, el.get(), stress, loads);
    else if (ld.kind == OBO_LOAD_TEMPERATURE) temperatureLoad_computeStressAndEquivalentLoads(ld, el.get(), stress, loads);
    else triOrPointLoad_computeStressAndEquivalentLoads(ld, el.get(), stress, loads);
    for (size_t p = 0; p < loads.size(); p++) {
      const size_t idx_node = el->conected_nodes_ids[p];
      for (int d = 0; d < 6; d++) {
        const int dof_idx = m_problem_DoFs_inverse_list[idx_node][d];
        if (dof_idx >= 0) {
          m_loads_at_each_dof_equivs[dof_idx] += loads[p][d];
        } else if (!(std::abs(loads[p][d]) < 1e-9)) {
          throw std::runtime_error("ASSERT_(std::abs(loads[p][d]) < 1e-9): load on a DOF not in the problem");
        }
      }
    }
    ElementStress& v = m_extra_stress_for_each_element[elIdxLoadPair.first];
    if (v.size() != stress.size()) {
      v = stress;
    } else {
      for (size_t p = 0; p < v.size(); p++) v[p] += stress[p];
    }
  }
}

// Eigen::SparseMatrix::setFromTriplets as in Eigen 3.3.x / 3.4.0 (internal::set_from_triplets):
// row-major temporary filled in triplet order, collapseDuplicates (sum into first occurrence,
// in order), transposed copy -> col-major with sorted inner indices.  Explicit zeros are kept.
static void setFromTriplets(SparseCSC& K, int n, const std::vector<int>& tr, const std::vector<int>& tc,
                            const std::vector<num_t>& tv) {
  const size_t T = tr.size();
  std::vector<int> rowStart(n + 1, 0);
  for (size_t t = 0; t < T; t++) rowStart[tr[t] + 1]++;
  for (int i = 0; i < n; i++) rowStart[i + 1] += rowStart[i];
  std::vector<int> fill(rowStart.begin(), rowStart.end() - 1), trInner(T);
  std::vector<num_t> trVal(T);
  for (size_t t = 0; t < T; t++) {
    const int p = fill[tr[t]]++;
    trInner[p] = tc[t];
    trVal[p] = tv[t];
  }
  // collapseDuplicates
  std::vector<int> wi(n, -1), newStart(n + 1, 0);
  int count = 0;
  for (int j = 0; j < n; j++) {
    const int start = count;
    newStart[j] = start;
    for (int k = rowStart[j]; k < rowStart[j + 1]; k++) {
      const int i = trInner[k];
      if (wi[i] >= start) {
        trVal[wi[i]] += trVal[k];
      } else {
        trVal[count] = trVal[k];
        trInner[count] = i;
        wi[i] = count;
        ++count;
      }
    }
  }
  newStart[n] = count;
  // transposed copy (row-major -> col-major)
  K.rows = K.cols = n;
  K.outer.assign(n + 1, 0);
  for (int p = 0; p < count; p++) K.outer[trInner[p] + 1]++;
  for (int i = 0; i < n; i++) K.outer[i + 1] += K.outer[i];
  K.inner.resize(count);
  K.val.resize(count);
  std::vector<int> pos(K.outer.begin(), K.outer.end() - 1);
  for (int j = 0; j < n; j++)
    for (int p = newStart[j]; p < newStart[j + 1]; p++) {
      const int q = pos[trInner[p]]++;
      K.inner[q] = j;
      K.val[q] = trVal[p];
    }
}

static void mul33T_left(Matrix66& M, int r0, int c0, const Matrix33& R) {  // block <- R^T * block
  num_t O[3][3];
  for (int r = 0; r < 3; r++)
    for (int c = 0; c < 3; c++) {
      num_t s = R(0, r) * M(r0 + 0, c0 + c);
      s += R(1, r) * M(r0 + 1, c0 + c);
      s += R(2, r) * M(r0 + 2, c0 + c);
      O[r][c] = s;
    }
  for (int r = 0; r < 3; r++)
    for (int c = 0; c < 3; c++) M(r0 + r, c0 + c) = O[r][c];
}
static void mul33_right(Matrix66& M, int r0, int c0, const Matrix33& R) {  // block <- block * R
  num_t O[3][3];
  for (int r = 0; r < 3; r++)
    for (int c = 0; c < 3; c++) {
      num_t s = M(r0 + r, c0 + 0) * R(0, c);
      s += M(r0 + r, c0 + 1) * R(1, c);
      s += M(r0 + r, c0 + 2) * R(2, c);
      O[r][c] = s;
    }
  for (int r = 0; r < 3; r++)
    for (int c = 0; c < 3; c++) M(r0 + r, c0 + c) = O[r][c];
}
static void rotT3(const Matrix33& R, num_t* v) {  // v <- R^T v
  const num_t a = R(0, 0) * v[0] + R(1, 0) * v[1] + R(2, 0) * v[2];
  const num_t b = R(0, 1) * v[0] + R(1, 1) * v[1] + R(2, 1) * v[2];
  const num_t c = R(0, 2) * v[0] + R(1, 2) * v[1] + R(2, 2) * v[2];
  v[0] = a; v[1] = b; v[2] = c;
}

// src/CFiniteElementProblem_assemble.cpp:37-437
void Problem::assembleProblem(bool skip_subK) {
  updateElementsOrientation();
  updateNodeConnections();
  updateListDoFs();
  applyConstraintsAndNodeLoads();  // (done by the caller/parser in the reference, after updateAll())

  const size_t nDOFs = m_problem_DoFs.size();
  const size_t nNodes = m_node_poses.size();
  const size_t nElements = m_elements.size();
  if (!(nDOFs > 0)) throw std::runtime_error("ASSERT_(nDOFs > 0)");

  std::map<size_t, size_t> problem_dof2bounded_dof_indices, problem_dof2free_dof_indices;
  free_dof_indices.clear();
  bounded_dof_indices.clear();
  dof_types.resize(nDOFs);
  for (size_t k = 0; k < nDOFs; k++) {  // :74-99
    if (auto it = m_DoF_constraints.find(k); it == m_DoF_constraints.end()) {
      const size_t new_idx = free_dof_indices.size();
      problem_dof2free_dof_indices.insert(problem_dof2free_dof_indices.end(), std::make_pair(k, new_idx));
      free_dof_indices.push_back(k);
      dof_types[k].bounded_index = NPOS;
      dof_types[k].free_index = new_idx;
    } else {
      const size_t new_idx = bounded_dof_indices.size();
      problem_dof2bounded_dof_indices.insert(problem_dof2bounded_dof_indices.end(), std::make_pair(k, new_idx));
      bounded_dof_indices.push_back(k);
      dof_types[k].bounded_index = new_idx;
      dof_types[k].free_index = NPOS;
    }
  }

  tri_r.clear(); tri_c.clear(); tri_v.clear();
  const size_t est = nNodes * 6 * 6 * (1 + 1);
  tri_r.reserve(est); tri_c.reserve(est); tri_v.reserve(est);

  for (size_t e = 0; e < nElements; e++) {  // :123-223
    auto el = m_elements[e];
    std::vector<TStiffnessSubmatrix> mats;
    el->getGlobalStiffnessMatrices(mats);
    for (size_t j = 0; j < mats.size(); j++) {
      const Matrix66& K_element_org = mats[j].matrix;
      const size_t i_node_idx = el->conected_nodes_ids[mats[j].edge_in];
      const size_t j_node_idx = el->conected_nodes_ids[mats[j].edge_out];
      Matrix66 KeNodal;
      const Matrix66* K_element = nullptr;
      const TRotationTrans3D& rt_i = getNodePose(i_node_idx);
      const TRotationTrans3D& rt_j = getNodePose(j_node_idx);
      if (rt_i.r.isIdentity() && rt_j.r.isIdentity()) {
        K_element = &K_element_org;
      } else {  // :144-179
        K_element = &KeNodal;
        KeNodal = K_element_org;
        if (!rt_i.r.isIdentity())
          for (int br = 0; br < 2; br++)
            for (int bc = 0; bc < 2; bc++) mul33T_left(KeNodal, 3 * br, 3 * bc, rt_i.r.rot);
        if (!rt_j.r.isIdentity())
          for (int br = 0; br < 2; br++)
            for (int bc = 0; bc < 2; bc++) mul33_right(KeNodal, 3 * br, 3 * bc, rt_j.r.rot);
      }
      const TNodeConnections& nc_i = m_node_connections[i_node_idx];
      const TNodeConnections& nc_j = m_node_connections[j_node_idx];
      auto it_nod_i = nc_i.find(e);
      auto it_nod_j = nc_j.find(e);
      if (nc_i.end() == it_nod_i || nc_j.end() == it_nod_j) throw std::runtime_error("ASSERT_(node connection)");
      const std::array<int, 6>& mat_rowidx2DoFidx = m_problem_DoFs_inverse_list[i_node_idx];
      const std::array<int, 6>& mat_colidx2DoFidx = m_problem_DoFs_inverse_list[j_node_idx];
      for (int r = 0; r < 6; r++) {
        const int r_idx_in_K = mat_rowidx2DoFidx[r];
        if (r_idx_in_K < 0) continue;
        for (int c = (i_node_idx == j_node_idx ? r : 0); c < 6; c++) {
          const int c_idx_in_K = mat_colidx2DoFidx[c];
          if (c_idx_in_K < 0) continue;
          if (r_idx_in_K > c_idx_in_K) {
            tri_r.push_back(c_idx_in_K); tri_c.push_back(r_idx_in_K);
          } else {
            tri_r.push_back(r_idx_in_K); tri_c.push_back(c_idx_in_K);
          }
          tri_v.push_back((*K_element)(r, c));
        }
      }
    }
  }

  setFromTriplets(K, int(nDOFs), tri_r, tri_c, tri_v);  // :231-233

  const int nDOFs_f = int(free_dof_indices.size());
  const int nDOFs_b = int(bounded_dof_indices.size());
  K_ff.clear(); K_bf.clear(); K_bb.clear();
  if (!skip_subK) {  // :249-297 — structurally dense sub-matrices, every entry via coeff()
    K_bb.assign(size_t(nDOFs_b) * nDOFs_b, 0);
    K_ff.assign(size_t(nDOFs_f) * nDOFs_f, 0);
    K_bf.assign(size_t(nDOFs_b) * nDOFs_f, 0);
    for (int r = 0; r < nDOFs_b; r++)
      for (int c = r; c < nDOFs_b; c++)
        K_bb[size_t(c) * nDOFs_b + r] = K.coeff(int(bounded_dof_indices[r]), int(bounded_dof_indices[c]));
    for (int r = 0; r < nDOFs_f; r++)
      for (int c = r; c < nDOFs_f; c++)  // stored at (c,r): LOWER triangle
        K_ff[size_t(r) * nDOFs_f + c] = K.coeff(int(free_dof_indices[r]), int(free_dof_indices[c]));
    for (int r = 0; r < nDOFs_b; r++)
      for (int c = 0; c < nDOFs_f; c++) {
        const int i_b = int(bounded_dof_indices[r]);
        const int i_f = int(free_dof_indices[c]);
        K_bf[size_t(c) * nDOFs_b + r] = (i_f >= i_b) ? K.coeff(i_b, i_f) : K.coeff(i_f, i_b);
      }
  }

  {  // :304-313
    int idx = 0;
    U_b.assign(m_DoF_constraints.size(), 0);
    for (auto it = m_DoF_constraints.begin(); it != m_DoF_constraints.end(); ++it) {
      if (it->first != bounded_dof_indices[idx]) throw std::runtime_error("ASSERT_(constraint order)");
      U_b[idx++] = it->second;
    }
  }

  internalComputeStressAndEquivalentLoads();  // :321
  F_f.assign(nDOFs_f, 0);
  for (int i = 0; i < nDOFs_f; i++) {  // :326-341
    const size_t dofIdx = free_dof_indices[i];
    if (auto it = m_loads_at_each_dof.find(dofIdx); it != m_loads_at_each_dof.end()) F_f[i] = it->second;
    if (auto it = m_loads_at_each_dof_equivs.find(dofIdx); it != m_loads_at_each_dof_equivs.end())
      F_f[i] += it->second;
  }

  for (size_t i = 0; i < nNodes; i++) {  // :349-435 nodal coordinates
    const TRotationTrans3D& rt_i = getNodePose(i);
    if (rt_i.r.isIdentity()) continue;
    num_t* Fs[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    num_t* Us[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    num_t Fi[3] = {0, 0, 0}, Mi[3] = {0, 0, 0}, Ui[3] = {0, 0, 0}, URi[3] = {0, 0, 0};
    const std::array<int, 6>& dof_i = m_problem_DoFs_inverse_list[i];
    for (int k = 0; k < 6; k++) {
      if (dof_i[k] != -1) {
        if (auto it = problem_dof2free_dof_indices.find(dof_i[k]); it != problem_dof2free_dof_indices.end()) {
          Fs[k] = &F_f[it->second];
        } else {
          auto it2 = problem_dof2bounded_dof_indices.find(dof_i[k]);
          if (it2 == problem_dof2bounded_dof_indices.end()) throw std::runtime_error("ASSERT_(bounded)");
          Us[k] = &U_b[it2->second];
        }
      }
    }
    for (int k = 0; k < 6; k++) {
      if (Fs[k]) (k < 3 ? Fi[k] : Mi[k - 3]) = *Fs[k];
      if (Us[k]) (k < 3 ? Ui[k] : URi[k - 3]) = *Us[k];
    }
    rotT3(rt_i.r.rot, Fi); rotT3(rt_i.r.rot, Mi); rotT3(rt_i.r.rot, Ui); rotT3(rt_i.r.rot, URi);
    for (int k = 0; k < 6; k++) {
      if (Fs[k]) *Fs[k] = (k < 3 ? Fi[k] : Mi[k - 3]);
      if (Us[k]) *Us[k] = (k < 3 ? Ui[k] : URi[k - 3]);
    }
  }
}

}  // namespace obo
