// ob_oracle.cpp — CPU restatement of libopenbeam's static-solve path.
// TEST INFRASTRUCTURE ONLY (see ob_oracle.h).  PARITY UNPINNED by reference tests (none exist).
//
// Every function cites the reference lines it restates (paths relative to
// /root/reference/libopenbeam/).  Data structures deliberately mirror the reference's
// (std::map / std::deque / shared_ptr / virtual dispatch / heap vectors per element /
// O(n^2) coeff() extraction / dense lower Cholesky) so that timing this code is a fair
// stand-in for timing ob-solve's own CPU path ("faithful mode").
// Build: g++ -O3 -ffp-contract=off -std=c++17 -shared -fPIC  (reference: -O3 only, CMakeLists.txt:158-160)
#include "ob_oracle.h"

#include <algorithm>
#include <array>
#include <chrono>
#include <cmath>
#include <cstring>
#include <deque>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

namespace obo {

using num_t = double;  // include/openbeam/config.h:25

// ---- tiny fixed-size dense helpers standing in for Eigen Matrix33 / Matrix66 (col-major) ----
struct Matrix33 {
  num_t m[9];
  num_t& operator()(int r, int c) { return m[c * 3 + r]; }
  num_t operator()(int r, int c) const { return m[c * 3 + r]; }
  void setIdentity() {
    for (int i = 0; i < 9; i++) m[i] = 0;
    m[0] = m[4] = m[8] = 1;
  }
};
struct Matrix66 {
  num_t m[36];
  num_t& operator()(int r, int c) { return m[c * 6 + r]; }
  num_t operator()(int r, int c) const { return m[c * 6 + r]; }
  void setZero() {
    for (int i = 0; i < 36; i++) m[i] = 0;
  }
};

// include/openbeam/types.h:186-230 ; src/CElement.cpp:81-109,167-176
struct TRotation3D {
  Matrix33 rot;
  bool m_is_pure_identity = true;
  TRotation3D() { rot.setIdentity(); }
  TRotation3D(num_t roll, num_t pitch, num_t yaw) {
    if (roll == 0 && pitch == 0 && yaw == 0) {
      m_is_pure_identity = true;
      rot.setIdentity();
    } else {
      m_is_pure_identity = false;
      const num_t cy = cos(yaw), sy = sin(yaw), cp = cos(pitch), sp = sin(pitch), cr = cos(roll), sr = sin(roll);
      rot(0, 0) = cy * cp;
      rot(0, 1) = cy * sp * sr - sy * cr;
      rot(0, 2) = cy * sp * cr + sy * sr;
      rot(1, 0) = sy * cp;
      rot(1, 1) = sy * sp * sr + cy * cr;
      rot(1, 2) = sy * sp * cr - cy * sr;
      rot(2, 0) = -sp;
      rot(2, 1) = cp * sr;
      rot(2, 2) = cp * cr;
    }
  }
  void setRot(const Matrix33& r) {
    rot = r;
    m_is_pure_identity = false;
  }
  bool isIdentity() const {  // CElement.cpp:167-176 (diagonal only, quirk Q5)
    if (m_is_pure_identity) return true;
    return rot(0, 0) == 1 && rot(1, 1) == 1 && rot(2, 2) == 1;
  }
};
struct TRotationTrans3D {
  num_t t[3] = {0, 0, 0};
  TRotation3D r;
};

struct FaceStress {  // types.h:78-157 ; operator+=(vec) maps 0..5 -> N,Vy,Vz,Mx,My,Mz
  num_t N = 0, Vy = 0, Vz = 0, Mx = 0, My = 0, Mz = 0;
  FaceStress& operator+=(const FaceStress& o) {
    N += o.N; Vy += o.Vy; Vz += o.Vz; Mx += o.Mx; My += o.My; Mz += o.Mz;
    return *this;
  }
  void addVec(const num_t* o) {
    N += o[0]; Vy += o[1]; Vz += o[2]; Mx += o[3]; My += o[4]; Mz += o[5];
  }
};
using ElementStress = std::vector<FaceStress>;
using used_DoFs_t = std::array<bool, 6>;
using array6 = std::array<num_t, 6>;

struct TStiffnessSubmatrix {  // CElement.h:37-41
  size_t edge_in, edge_out;
  Matrix66 matrix;
};

struct Problem;

// src/internals.cpp:30-123
static void computeOrientationFromTwoPoints(const num_t* t1, const num_t* t2, const num_t rotation_around_x,
                                            TRotation3D& out_R) {
  const num_t p12[3] = {t2[0] - t1[0], t2[1] - t1[1], t2[2] - t1[2]};
  const num_t len = std::sqrt(p12[0] * p12[0] + p12[1] * p12[1] + p12[2] * p12[2]);  // TPoint3D::norm()
  Matrix33 R;
  if (len > 0) {
    const num_t len_inv = 1 / len;
    num_t X[3], Y[3], Z[3];
    X[0] = p12[0] * len_inv;
    X[1] = p12[1] * len_inv;
    X[2] = p12[2] * len_inv;
    if (X[2] == 0) {
      Z[0] = 0; Z[1] = 0; Z[2] = 1;
    } else {
      if (X[0] != 0 || X[1] != 0) {
        Z[0] = -X[0];
        Z[1] = -X[1];
        Z[2] = (X[0] * X[0] + X[1] * X[1]) / X[2];
        const num_t norm_fact = num_t(1) / std::sqrt(Z[0] * Z[0] + Z[1] * Z[1] + Z[2] * Z[2]);
        Z[0] *= norm_fact; Z[1] *= norm_fact; Z[2] *= norm_fact;
      } else {
        Z[0] = -1; Z[1] = 0; Z[2] = 0;
      }
    }
    Y[0] = Z[1] * X[2] - Z[2] * X[1];
    Y[1] = Z[2] * X[0] - Z[0] * X[2];
    Y[2] = Z[0] * X[1] - Z[1] * X[0];
    R(0, 0) = X[0]; R(0, 1) = Y[0]; R(0, 2) = Z[0];
    R(1, 0) = X[1]; R(1, 1) = Y[1]; R(1, 2) = Z[1];
    R(2, 0) = X[2]; R(2, 1) = Y[2]; R(2, 2) = Z[2];
  } else {
    R.setIdentity();
  }
  if (rotation_around_x != 0) {  // internals.cpp:106-120 (never taken from YAML input)
    const num_t c = cos(rotation_around_x), s = sin(rotation_around_x);
    for (int col = 0; col < 3; col++) {
      const num_t a = R(1, col), b = R(2, col);
      R(1, col) = c * a - s * b;
      R(2, col) = s * a + c * b;
    }
  }
  out_R.setRot(R);
}

// ---- elements: CElement (src/CElement.cpp) + CElementBeam_2D_{RR,RA,AR,AA} ----
struct CElement {
  std::vector<size_t> conected_nodes_ids;
  Problem* m_parent = nullptr;
  num_t m_design_rotation_around_linear_axis = 0;
  TRotation3D m_global_orientation;
  num_t E = 0, A = 0, Iz = 0, G = 0, J = 0;  // CBaseElementBeam.h:45-46 : G,J default 0
  int type = 0;
  virtual ~CElement() = default;
  virtual void getLocalStiffnessMatrices(std::vector<TStiffnessSubmatrix>& outSubMats) const = 0;
  virtual void getLocalDoFs(std::vector<used_DoFs_t>& dofs) const = 0;
  virtual std::shared_ptr<CElement> clone() const = 0;
  void updateOrientationFromNodePositions();                               // CElement.cpp:178-188
  void getGlobalDoFs(std::vector<used_DoFs_t>& dofs) const;                // CElement.cpp:192-247
  void getGlobalStiffnessMatrices(std::vector<TStiffnessSubmatrix>&) const;  // CElement.cpp:53-79
  void lengths(num_t& L, num_t& L2, num_t& L3) const;
};

// loads.h: CLoadDistributedUniform (q, dir) | CLoadDistributedTriangular (q_ini = q, q_end = q2, dir) |
// CLoadConcentratedForce (P = q, dist = q2, dir) | CLoadConstTemperature (m_incr_temp = q)
enum { OBO_LOAD_UNIFORM = 0, OBO_LOAD_TRIANGULAR = 1, OBO_LOAD_CONCENTRATED = 2, OBO_LOAD_TEMPERATURE = 3 };
struct ElemLoad {
  num_t q;
  num_t dir[3];
  int kind = OBO_LOAD_UNIFORM;
  num_t q2 = 0;
};

struct TDoFType {
  size_t bounded_index, free_index;
};
static const size_t NPOS = std::string::npos;

struct NodeConn {
  unsigned char elementFaceId;
  used_DoFs_t dofs;
};
using TNodeConnections = std::map<size_t, NodeConn>;

struct SparseCSC {  // Eigen::SparseMatrix<double> (col-major, compressed)
  int rows = 0, cols = 0;
  std::vector<int> outer, inner;
  std::vector<num_t> val;
  num_t coeff(int r, int c) const {  // binary search inside column c, 0 if absent
    const int* b = inner.data() + outer[c];
    const int* e = inner.data() + outer[c + 1];
    const int* p = std::lower_bound(b, e, r);
    return (p != e && *p == r) ? val[p - inner.data()] : num_t(0);
  }
};

struct Problem {
  // ---- model (CFiniteElementProblem.h:475-560) ----
  std::deque<TRotationTrans3D> m_node_poses;
  std::deque<std::shared_ptr<CElement>> m_elements;
  std::vector<std::array<size_t, 2>> raw_constraints_nd;  // (node,dof) as given by the caller
  std::vector<num_t> raw_constraints_val;
  std::vector<std::array<size_t, 2>> raw_nloads_nd;
  std::vector<num_t> raw_nloads_val;
  std::multimap<size_t, ElemLoad> m_loads_on_beams;
  std::vector<size_t> raw_eload_elem;  // insertion order, to re-set q values

  std::vector<std::array<size_t, 2>> m_problem_DoFs;  // (nodeId, dof)
  std::vector<std::array<int, 6>> m_problem_DoFs_inverse_list;
  std::vector<TNodeConnections> m_node_connections;
  std::map<size_t, num_t> m_DoF_constraints, m_loads_at_each_dof, m_loads_at_each_dof_equivs;
  std::map<size_t, ElementStress> m_extra_stress_for_each_element;

  // ---- BuildProblemInfo / StaticSolveProblemInfo / StressInfo ----
  std::vector<size_t> free_dof_indices, bounded_dof_indices;
  std::vector<TDoFType> dof_types;
  std::vector<int> tri_r, tri_c;
  std::vector<num_t> tri_v;
  SparseCSC K;
  std::vector<num_t> K_ff, K_bf, K_bb;  // dense col-major views of the structurally dense sparse matrices
  std::vector<num_t> U_b, F_f, U_f, F_b;
  std::vector<ElementStress> element_stress;
  std::string err;

  const TRotationTrans3D& getNodePose(size_t i) const { return m_node_poses.at(i); }
  void updateElementsOrientation() {  // CFiniteElementProblem.cpp:145-151
    for (auto& e : m_elements) e->updateOrientationFromNodePositions();
  }
  void updateNodeConnections();
  void updateListDoFs();
  void applyConstraintsAndNodeLoads();
  void internalComputeStressAndEquivalentLoads();
  void assembleProblem(bool skip_subK);
  void solveStatic(int algorithm);
  void postProcCalcStress();
  std::unique_ptr<Problem> clone() const;
};

void CElement::updateOrientationFromNodePositions() {
  const TRotationTrans3D& p1 = m_parent->getNodePose(conected_nodes_ids[0]);
  const TRotationTrans3D& p2 = m_parent->getNodePose(conected_nodes_ids[1]);
  computeOrientationFromTwoPoints(p1.t, p2.t, m_design_rotation_around_linear_axis, m_global_orientation);
}

void CElement::lengths(num_t& L, num_t& L2, num_t& L3) const {  // CElementBeam_2D_RR.cpp:54-61
  const TRotationTrans3D& node0 = m_parent->getNodePose(conected_nodes_ids[0]);
  const TRotationTrans3D& node1 = m_parent->getNodePose(conected_nodes_ids[1]);
  const num_t d[3] = {node1.t[0] - node0.t[0], node1.t[1] - node0.t[1], node1.t[2] - node0.t[2]};
  L2 = d[0] * d[0] + d[1] * d[1] + d[2] * d[2];
  L = std::sqrt(L2);
  L3 = L2 * L;
  if (!(L > 0)) throw std::runtime_error("ASSERT_(L > 0)");
}

void CElement::getGlobalDoFs(std::vector<used_DoFs_t>& dofs) const {  // CElement.cpp:192-247
  std::vector<used_DoFs_t> local_dofs;
  getLocalDoFs(local_dofs);
  const size_t N = local_dofs.size();
  dofs.resize(N);
  const Matrix33& R = m_global_orientation.rot;
  bool RnonZero[3][3];
  for (int r = 0; r < 3; r++)
    for (int c = 0; c < 3; c++) RnonZero[r][c] = std::abs(R(r, c)) > 1e-10;
  for (size_t i = 0; i < N; i++) {
    for (int blk = 0; blk < 2; blk++) {
      const int o = 3 * blk;
      if (local_dofs[i][o + 0] && local_dofs[i][o + 1] && local_dofs[i][o + 2]) {
        dofs[i][o + 0] = dofs[i][o + 1] = dofs[i][o + 2] = true;
      } else {
        for (int g = 0; g < 3; g++)
          dofs[i][o + g] = (local_dofs[i][o + 0] && RnonZero[g][0]) || (local_dofs[i][o + 1] && RnonZero[g][1]) ||
                           (local_dofs[i][o + 2] && RnonZero[g][2]);
      }
    }
  }
}

static bool anyNonZero3x3(const Matrix66& M, int r0, int c0) {
  for (int c = 0; c < 3; c++)
    for (int r = 0; r < 3; r++)
      if (M(r0 + r, c0 + c) != num_t(0)) return true;
  return false;
}
// B <- R * B * R^T on one 3x3 sub-block (Eigen evaluates (R*B) into a temporary, then *R^T)
static void rotateBlock(Matrix66& M, int r0, int c0, const Matrix33& R) {
  num_t T[3][3], O[3][3];
  for (int r = 0; r < 3; r++)
    for (int c = 0; c < 3; c++) {
      num_t s = R(r, 0) * M(r0 + 0, c0 + c);
      s += R(r, 1) * M(r0 + 1, c0 + c);
      s += R(r, 2) * M(r0 + 2, c0 + c);
      T[r][c] = s;
    }
  for (int r = 0; r < 3; r++)
    for (int c = 0; c < 3; c++) {
      num_t s = T[r][0] * R(c, 0);
      s += T[r][1] * R(c, 1);
      s += T[r][2] * R(c, 2);
      O[r][c] = s;
    }
  for (int r = 0; r < 3; r++)
    for (int c = 0; c < 3; c++) M(r0 + r, c0 + c) = O[r][c];
}
void CElement::getGlobalStiffnessMatrices(std::vector<TStiffnessSubmatrix>& outSubMats) const {  // CElement.cpp:53-79
  this->getLocalStiffnessMatrices(outSubMats);
  const Matrix33& R = m_global_orientation.rot;
  for (size_t i = 0; i < outSubMats.size(); i++) {
    Matrix66& M = outSubMats[i].matrix;
    if (anyNonZero3x3(M, 0, 0)) rotateBlock(M, 0, 0, R);
    if (anyNonZero3x3(M, 0, 3)) rotateBlock(M, 0, 3, R);
    if (anyNonZero3x3(M, 3, 0)) rotateBlock(M, 3, 0, R);
    if (anyNonZero3x3(M, 3, 3)) rotateBlock(M, 3, 3, R);
  }
}

static const num_t UNINITIALIZED_VALUE = 1.7976931348623157e308;  // types.h:56 (DBL_MAX)

static void initSub(std::vector<TStiffnessSubmatrix>& o) {
  o.resize(3);
  o[0].edge_in = 0; o[0].edge_out = 0;
  o[1].edge_in = 1; o[1].edge_out = 1;
  o[2].edge_in = 0; o[2].edge_out = 1;
  for (auto& s : o) s.matrix.setZero();
}
static void checkInit(const CElement* e, const char* cls, bool needIz) {
  if (e->E == UNINITIALIZED_VALUE) throw std::runtime_error(std::string(cls) + ": Uninitialized parameter E");
  if (e->A == UNINITIALIZED_VALUE) throw std::runtime_error(std::string(cls) + ": Uninitialized parameter A");
  if (needIz && e->Iz == UNINITIALIZED_VALUE) throw std::runtime_error(std::string(cls) + ": Uninitialized parameter Iz");
}

struct CElementBeam_2D_RR : CElement {  // src/CElementBeam_2D_RR.cpp:41-139
  void getLocalStiffnessMatrices(std::vector<TStiffnessSubmatrix>& o) const override {
    initSub(o);
    num_t L, L2, L3;
    lengths(L, L2, L3);
    checkInit(this, "CElementBeam_2D_RR", true);
    Matrix66 &K11 = o[0].matrix, &K22 = o[1].matrix, &K12 = o[2].matrix;
    K11(0, 0) = E * A / L;
    K11(1, 1) = 12 * E * Iz / L3;
    K11(1, 5) = K11(5, 1) = 6 * E * Iz / L2;
    K11(3, 3) = G * J / L;
    K11(5, 5) = 4 * E * Iz / L;
    K22(0, 0) = K11(0, 0);
    K22(1, 1) = K11(1, 1);
    K22(1, 5) = K22(5, 1) = -K11(5, 1);
    K22(3, 3) = K11(3, 3);
    K22(5, 5) = K11(5, 5);
    K12(0, 0) = -K11(0, 0);
    K12(1, 1) = -K11(1, 1);
    K12(1, 5) = K11(1, 5);
    K12(5, 1) = -K11(5, 1);
    K12(3, 3) = -K11(3, 3);
    K12(5, 5) = 0.5 * K11(5, 5);
  }
  void getLocalDoFs(std::vector<used_DoFs_t>& d) const override {
    d.resize(2);
    d[0] = d[1] = used_DoFs_t{true, true, false, false, false, true};
  }
  std::shared_ptr<CElement> clone() const override { return std::make_shared<CElementBeam_2D_RR>(*this); }
};
struct CElementBeam_2D_RA : CElement {  // src/CElementBeam_2D_RA.cpp:41-119
  void getLocalStiffnessMatrices(std::vector<TStiffnessSubmatrix>& o) const override {
    initSub(o);
    num_t L, L2, L3;
    lengths(L, L2, L3);
    checkInit(this, "CElementBeam_2D_RA", true);
    Matrix66 &K11 = o[0].matrix, &K22 = o[1].matrix, &K12 = o[2].matrix;
    K11(0, 0) = E * A / L;
    K11(1, 1) = 3 * E * Iz / L3;
    K11(1, 5) = K11(5, 1) = 3 * E * Iz / L2;
    K11(5, 5) = 3 * E * Iz / L;
    K22(0, 0) = K11(0, 0);
    K22(1, 1) = K11(1, 1);
    K12(0, 0) = -K11(0, 0);
    K12(1, 1) = -K11(1, 1);
    K12(5, 1) = -K11(1, 5);
  }
  void getLocalDoFs(std::vector<used_DoFs_t>& d) const override {
    d.resize(2);
    d[0] = used_DoFs_t{true, true, false, false, false, true};
    d[1] = used_DoFs_t{true, true, false, false, false, false};
  }
  std::shared_ptr<CElement> clone() const override { return std::make_shared<CElementBeam_2D_RA>(*this); }
};
struct CElementBeam_2D_AR : CElement {  // src/CElementBeam_2D_AR.cpp:41-119
  void getLocalStiffnessMatrices(std::vector<TStiffnessSubmatrix>& o) const override {
    initSub(o);
    num_t L, L2, L3;
    lengths(L, L2, L3);
    checkInit(this, "CElementBeam_2D_AR", true);
    Matrix66 &K11 = o[0].matrix, &K22 = o[1].matrix, &K12 = o[2].matrix;
    K11(0, 0) = E * A / L;
    K11(1, 1) = 3 * E * Iz / L3;
    K22(0, 0) = K11(0, 0);
    K22(1, 1) = K11(1, 1);
    K22(1, 5) = K22(5, 1) = -3 * E * Iz / L2;
    K22(5, 5) = 3 * E * Iz / L;
    K12(0, 0) = -K11(0, 0);
    K12(1, 1) = -K11(1, 1);
    K12(1, 5) = -K22(1, 5);
  }
  void getLocalDoFs(std::vector<used_DoFs_t>& d) const override {
    d.resize(2);
    d[0] = used_DoFs_t{true, true, false, false, false, false};
    d[1] = used_DoFs_t{true, true, false, false, false, true};
  }
  std::shared_ptr<CElement> clone() const override { return std::make_shared<CElementBeam_2D_AR>(*this); }
};
struct CElementBeam_2D_AA : CElement {  // src/CElementBeam_2D_AA.cpp:41-106
  void getLocalStiffnessMatrices(std::vector<TStiffnessSubmatrix>& o) const override {
    initSub(o);
    num_t L, L2, L3;
    lengths(L, L2, L3);
    checkInit(this, "CElementBeam_2D_AA", false);
    Matrix66 &K11 = o[0].matrix, &K22 = o[1].matrix, &K12 = o[2].matrix;
    K11(0, 0) = E * A / L;
    K22 = K11;
    K12(0, 0) = -K11(0, 0);
  }
  void getLocalDoFs(std::vector<used_DoFs_t>& d) const override {
    d.resize(2);
    d[0] = d[1] = used_DoFs_t{true, false, false, false, false, false};
  }
  std::shared_ptr<CElement> clone() const override { return std::make_shared<CElementBeam_2D_AA>(*this); }
};
struct CElementBeam_2D_RD : CElement {  // src/CElementBeam_2D_RD.cpp:41-125 (rigid end - slider end)
  void getLocalStiffnessMatrices(std::vector<TStiffnessSubmatrix>& o) const override {
    initSub(o);
    num_t L, L2, L3;
    lengths(L, L2, L3);
    checkInit(this, "CElementBeam_2D_RD", true);
    Matrix66 &K11 = o[0].matrix, &K22 = o[1].matrix, &K12 = o[2].matrix;
    K11(0, 0) = E * A / L;
    K11(5, 5) = E * Iz / L;
    K22(0, 0) = K11(0, 0);
    K22(5, 5) = K11(5, 5);
    K12(0, 0) = -K11(0, 0);
    K12(5, 5) = -K11(5, 5);
  }
  void getLocalDoFs(std::vector<used_DoFs_t>& d) const override {
    d.resize(2);
    d[0] = used_DoFs_t{true, true, false, false, false, true};
    d[1] = used_DoFs_t{true, false, false, false, false, true};
  }
  std::shared_ptr<CElement> clone() const override { return std::make_shared<CElementBeam_2D_RD>(*this); }
};
// The spring elements keep their constants in the generic parameter triple: E = K | Kx, A = Ky, Iz = KRz.
struct CElementSpring : CElement {  // src/CElementSpring.cpp:41-80 (SPRING_1D: along the element's local x)
  void getLocalStiffnessMatrices(std::vector<TStiffnessSubmatrix>& o) const override {
    initSub(o);
    if (E == UNINITIALIZED_VALUE) throw std::runtime_error("CElementSpring: Uninitialized parameter K");
    o[0].matrix(0, 0) = E;
    o[1].matrix = o[0].matrix;
    o[2].matrix(0, 0) = -E;
  }
  void getLocalDoFs(std::vector<used_DoFs_t>& d) const override {
    d.resize(2);
    d[0] = d[1] = used_DoFs_t{true, false, false, false, false, false};
  }
  std::shared_ptr<CElement> clone() const override { return std::make_shared<CElementSpring>(*this); }
};
struct CElementSpringXY : CElement {  // src/CElementSpringXY.cpp:31-66
  void getLocalStiffnessMatrices(std::vector<TStiffnessSubmatrix>& o) const override {
    initSub(o);
    o[0].matrix(0, 0) = E;
    o[0].matrix(1, 1) = A;
    o[1].matrix = o[0].matrix;
    o[2].matrix(0, 0) = -E;
    o[2].matrix(1, 1) = -A;
  }
  void getLocalDoFs(std::vector<used_DoFs_t>& d) const override {
    d.resize(2);
    d[0] = d[1] = used_DoFs_t{E != 0, A != 0, false, false, false, false};
  }
  std::shared_ptr<CElement> clone() const override { return std::make_shared<CElementSpringXY>(*this); }
};
struct CElementSpringDXDYRZ : CElement {  // src/CElementSpringDXDYRZ.cpp:31-70
  void getLocalStiffnessMatrices(std::vector<TStiffnessSubmatrix>& o) const override {
    initSub(o);
    Matrix66 &K11 = o[0].matrix, &K22 = o[1].matrix, &K12 = o[2].matrix;
    K11(0, 0) = E;
    K11(1, 1) = A;
    K11(5, 5) = Iz;
    K22 = K11;
    K12(0, 0) = -E;
    K12(1, 1) = -A;
    K11(5, 5) = -Iz;  // sic (:62): the reference assigns K11 here, not K12 -> K11(5,5) = -KRz, K12(5,5) stays 0
  }
  void getLocalDoFs(std::vector<used_DoFs_t>& d) const override {
    d.resize(2);
    d[0] = d[1] = used_DoFs_t{E != 0, A != 0, false, false, false, Iz != 0};
  }
  std::shared_ptr<CElement> clone() const override { return std::make_shared<CElementSpringDXDYRZ>(*this); }
};
struct CElementTorsionSpring : CElement {  // src/CElementTorsionSpring.cpp:41-75
  void getLocalStiffnessMatrices(std::vector<TStiffnessSubmatrix>& o) const override {
    initSub(o);
    if (E == UNINITIALIZED_VALUE) throw std::runtime_error("CElementTorsionSpring: Uninitialized parameter K");
    o[0].matrix(5, 5) = E;
    o[1].matrix = o[0].matrix;
    o[2].matrix(5, 5) = -E;
  }
  void getLocalDoFs(std::vector<used_DoFs_t>& d) const override {
    d.resize(2);
    d[0] = d[1] = used_DoFs_t{false, false, false, false, false, true};
  }
  std::shared_ptr<CElement> clone() const override { return std::make_shared<CElementTorsionSpring>(*this); }
};
static std::shared_ptr<CElement> createElementByType(int t) {
  switch (t) {
    case OBO_RR: return std::make_shared<CElementBeam_2D_RR>();
    case OBO_RA: return std::make_shared<CElementBeam_2D_RA>();
    case OBO_AR: return std::make_shared<CElementBeam_2D_AR>();
    case OBO_AA: return std::make_shared<CElementBeam_2D_AA>();
    case OBO_RD: return std::make_shared<CElementBeam_2D_RD>();
    case OBO_SPRING_1D: return std::make_shared<CElementSpring>();
    case OBO_SPRING_XY: return std::make_shared<CElementSpringXY>();
    case OBO_SPRING_DXDYRZ: return std::make_shared<CElementSpringDXDYRZ>();
    case OBO_SPRING_TORSION: return std::make_shared<CElementTorsionSpring>();
  }
  return nullptr;
}

}  // namespace obo

// The rest of the restatement lives in three implementation fragments (same translation unit):
#include "ob_assemble.ipp"  // updateNodeConnections / updateListDoFs / assembleProblem / equivalent loads
#include "ob_solve.ipp"     // solveStatic / postProcCalcStress
#include "ob_capi.ipp"      // extern "C" surface

#include "ob_batch.ipp"
