// obhost.hpp — header-only C++ mirror of libopenbeam's static-solve interface over the C ABI (obgpu.h).
//
// Same method names, argument meaning and error behaviour as the reference (paths under
// /root/reference/libopenbeam/include/openbeam/):
//   BuildProblemInfo          CFiniteElementProblem.h:76-102
//   StaticSolveProblemInfo    CFiniteElementProblem.h:105-123
//   StaticSolverAlgorithm     CFiniteElementProblem.h:69-73    (+ GPU_* cases)
//   StaticSolverOptions       CFiniteElementProblem.h:165-171
//   StressInfo, FaceStress    CFiniteElementProblem.h:173-183, types.h:78-157
//   assembleProblem / solveStatic / postProcCalcStress   CFiniteElementProblem.h:435,444-446,453-454
// Errors: std::runtime_error with the reference's message for a non-positive-definite K_ff
// (CFiniteElementProblem_solver_static.cpp:98-105).  No arithmetic happens in this header.
#pragma once
#include <array>
#include <cstddef>
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

#include "obgpu.h"

namespace obhost {

using num_t = double;  // config.h:25

enum class StaticSolverAlgorithm : uint8_t {
  Dense_LLT = 0,   // reference CPU algorithms: refused here (libobgpu has no CPU fallback)
  Sparse_LLT,
  GPU_Batched_LDLT,
  GPU_PCG_Jacobi,
  GPU_PCG_MG,
};

struct StaticSolverOptions {
  StaticSolverAlgorithm algorithm = StaticSolverAlgorithm::GPU_Batched_LDLT;
  bool nonLinearIterative = false;
  double pcg_rtol = 1e-12;
  int pcg_max_iter = 0;
};

struct TDoFType {  // BuildProblemInfo::dof_types entries (assemble.cpp:83-97)
  size_t bounded_index = std::string::npos, free_index = std::string::npos;
};

struct BuildProblemInfo {
  std::vector<size_t> free_dof_indices, bounded_dof_indices;
  std::vector<TDoFType> dof_types;
  // K: upper triangle in Eigen's compressed column order (== CSR of the lower triangle)
  std::vector<int32_t> K_outer, K_inner;
  std::vector<num_t> K_values;
  // CFiniteElementProblem.h:86-91, dense-structured and column-major as the reference fills them (assemble.cpp:249-297):
  // K_ff [n_f x n_f] LOWER triangle, K_bf [n_b x n_f], K_bb [n_b x n_b] UPPER triangle (quirk Q1).  Filled for problems
  // of up to 4096 DOFs; the true-sparse K_ff is always in K_ff_rowptr / K_ff_col / K_ff_val (CSR, lower, free indices).
  std::vector<num_t> K_ff, K_bf, K_bb;
  std::vector<int32_t> K_ff_rowptr, K_ff_col;
  std::vector<num_t> K_ff_val;
  std::vector<num_t> U_b, F_f;
};

struct StaticSolveProblemInfo {
  BuildProblemInfo build_info;
  std::vector<num_t> U_f, F_b;
};

struct FaceStress {
  num_t N = 0, Vy = 0, Vz = 0, Mx = 0, My = 0, Mz = 0;
};
using ElementStress = std::vector<FaceStress>;
struct StressInfo {
  std::vector<ElementStress> element_stress;
};

struct NodeDoF {  // CFiniteElementProblem.h:53-67
  size_t nodeId;
  uint8_t dof;
};

// One structure (2D beam elements RR/RA/AR/AA) bound to one GPU context.
struct MeshParams {  // CStructureProblem_mesh.cpp:34
  num_t max_element_length = 50e-3;
};
struct MeshOutputInfo {  // CFiniteElementProblem.h MeshOutputInfo
  size_t num_original_nodes = 0, num_original_elements = 0;
  std::vector<std::vector<size_t>> element2nodes, element2elements;
};

class CStructureProblem {
 public:
  explicit CStructureProblem(int device = 0) {
    const int rc = obgpu_create(device, &ctx_);
    if (rc != OBGPU_OK) throw std::runtime_error(std::string("obgpu_create: ") + obgpu_status_string(rc));
  }
  ~CStructureProblem() { obgpu_destroy(ctx_); }
  CStructureProblem(const CStructureProblem&) = delete;
  CStructureProblem& operator=(const CStructureProblem&) = delete;

  // ---- model building (names follow CFiniteElementProblem.h:196-330) ----
  // rot_z: yaw of the nodal frame in radians (parse_file.cpp:467; TRotationTrans3D of the node)
  size_t insertNode(num_t x, num_t y, num_t rot_z = 0) {
    node_xy_.push_back(x);
    node_xy_.push_back(y);
    node_rotz_.push_back(rot_z);
    dirty_ = true;
    return node_xy_.size() / 2 - 1;
  }
  // type: OBGPU_BEAM2D_{RR,RA,AR,AA,RD} with E, A, Iz as CBaseElementBeam; OBGPU_SPRING_* with the spring constants in
  // the same triple (include/obgpu.h)
  size_t insertElement(uint8_t type, size_t n0, size_t n1, num_t E, num_t A, num_t Iz) {
    elem_nodes_.push_back(uint32_t(n0));
    elem_nodes_.push_back(uint32_t(n1));
    elem_type_.push_back(type);
    elem_EAI_.insert(elem_EAI_.end(), {E, A, Iz});
    dirty_ = true;
    return elem_type_.size() - 1;
  }
  void insertConstraint(size_t node, uint8_t dof, num_t value = 0) {
    con_node_.push_back(uint32_t(node)); con_dof_.push_back(dof); con_val_.push_back(value);
    dirty_ = true;
  }
  void addLoadAtDOF(size_t node, uint8_t dof, num_t f) {
    nl_node_.push_back(uint32_t(node)); nl_dof_.push_back(dof); nl_val_.push_back(f);
    dirty_ = true;
  }
  // CLoadDistributedUniform(q, dir)
  void addLoadAtBeam(size_t element, num_t q, num_t dx, num_t dy, num_t dz) {
    addLoadAtBeam(element, OBGPU_LOAD_UNIFORM, q, 0, dx, dy, dz);
  }
  // the other load classes of loads.h: kind = OBGPU_LOAD_TRIANGULAR (q = q_ini, q2 = q_end), OBGPU_LOAD_CONCENTRATED
  // (q = P, q2 = dist), OBGPU_LOAD_TEMPERATURE (q = temperature increment, direction unused)
  void addLoadAtBeam(size_t element, uint8_t kind, num_t q, num_t q2, num_t dx, num_t dy, num_t dz) {
    el_elem_.push_back(uint32_t(element)); el_q_.push_back(q); el_q2_.push_back(q2); el_kind_.push_back(kind);
    el_dir_.insert(el_dir_.end(), {dx, dy, dz});
    dirty_ = true;
  }

  // CStructureProblem::mesh (CStructureProblem_mesh.cpp:37-132) through obgpu_mesh_create
  void mesh(CStructureProblem& out_fem, MeshOutputInfo& mo, const MeshParams& params = MeshParams()) {
    Offsets off(*this);
    obgpu_batch_desc d = describe(off);
    obgpu_mesh* m = nullptr;
    int rc = obgpu_mesh_create(&d, params.max_element_length, &m);
    if (rc != OBGPU_OK) throw std::runtime_error(std::string("obgpu_mesh_create: ") + obgpu_status_string(rc));
    obgpu_batch_desc o{};
    obgpu_mesh_desc(m, &o);
    const size_t N = size_t(o.node_off[1]), E = size_t(o.elem_off[1]), C = size_t(o.con_off[1]);
    const size_t NL = size_t(o.nload_off[1]), EL = size_t(o.eload_off[1]);
    out_fem.node_xy_.assign(o.node_xy, o.node_xy + 2 * N);
    if (o.node_rotz) out_fem.node_rotz_.assign(o.node_rotz, o.node_rotz + N);
    else out_fem.node_rotz_.assign(N, 0.0);
    out_fem.elem_nodes_.assign(o.elem_nodes, o.elem_nodes + 2 * E);
    out_fem.elem_type_.assign(o.elem_type, o.elem_type + E);
    out_fem.elem_EAI_.assign(o.elem_EAI, o.elem_EAI + 3 * E);
    out_fem.con_node_.assign(o.con_node, o.con_node + C); out_fem.con_dof_.assign(o.con_dof, o.con_dof + C);
    out_fem.con_val_.assign(o.con_val, o.con_val + C);
    out_fem.nl_node_.assign(o.nload_node, o.nload_node + NL); out_fem.nl_dof_.assign(o.nload_dof, o.nload_dof + NL);
    out_fem.nl_val_.assign(o.nload_val, o.nload_val + NL);
    out_fem.el_elem_.assign(o.eload_elem, o.eload_elem + EL); out_fem.el_q_.assign(o.eload_q, o.eload_q + EL);
    out_fem.el_dir_.assign(o.eload_dir, o.eload_dir + 3 * EL);
    if (o.eload_type) { out_fem.el_kind_.assign(o.eload_type, o.eload_type + EL); out_fem.el_q2_.assign(o.eload_q2, o.eload_q2 + EL); }
    else { out_fem.el_kind_.assign(EL, uint8_t(OBGPU_LOAD_UNIFORM)); out_fem.el_q2_.assign(EL, 0.0); }
    out_fem.dirty_ = true;
    const int64_t *e2e_off, *e2n_off;
    const uint32_t *e2e, *e2n;
    obgpu_mesh_info(m, &e2e_off, &e2e, &e2n_off, &e2n);
    mo = MeshOutputInfo();
    mo.num_original_nodes = getNumberOfNodes();
    mo.num_original_elements = getNumberOfElements();
    for (size_t e = 0; e < getNumberOfElements(); e++) {
      mo.element2elements.emplace_back(e2e + e2e_off[e], e2e + e2e_off[e + 1]);
      mo.element2nodes.emplace_back(e2n + e2n_off[e], e2n + e2n_off[e + 1]);
    }
    obgpu_mesh_destroy(m);
  }
  size_t getNumberOfNodes() const { return node_xy_.size() / 2; }
  size_t getNumberOfElements() const { return elem_type_.size(); }

  void updateAll() {
    if (stale_) { dirty_ = true; stale_ = false; }
    upload();
    check(obgpu_symbolic(ctx_));
  }
  std::vector<NodeDoF> getProblemDoFs() {
    updateAll();
    const auto z = sizes();
    std::vector<int32_t> nd(z[4]), id(z[4]);
    check(obgpu_get_symbolic(ctx_, OBGPU_SYM_DOF_NODE, nd.data()));
    check(obgpu_get_symbolic(ctx_, OBGPU_SYM_DOF_ID, id.data()));
    std::vector<NodeDoF> out(z[4]);
    for (size_t i = 0; i < out.size(); i++) out[i] = NodeDoF{size_t(nd[i]), uint8_t(id[i])};
    return out;
  }

  // ---- the hot path ----
  void assembleProblem(BuildProblemInfo& out) {
    updateAll();
    check(obgpu_assemble(ctx_));
    fill_lists(out);
    const auto z = sizes();
    out.K_values.resize(z[8]);
    check(obgpu_download(ctx_, OBGPU_RES_KVAL, out.K_values.data()));
    out.U_b.resize(z[6]);
    check(obgpu_download(ctx_, OBGPU_RES_UB, out.U_b.data()));
    out.F_f.resize(z[5]);
    check(obgpu_download(ctx_, OBGPU_RES_FF, out.F_f.data()));
    int64_t cnt = 0;
    check(obgpu_symbolic_count(ctx_, OBGPU_SYM_KFF_ROWPTR, &cnt));
    out.K_ff_rowptr.resize(size_t(cnt));
    check(obgpu_get_symbolic(ctx_, OBGPU_SYM_KFF_ROWPTR, out.K_ff_rowptr.data()));
    check(obgpu_symbolic_count(ctx_, OBGPU_SYM_KFF_COL, &cnt));
    out.K_ff_col.resize(size_t(cnt));
    out.K_ff_val.resize(size_t(cnt));
    if (cnt) {
      check(obgpu_get_symbolic(ctx_, OBGPU_SYM_KFF_COL, out.K_ff_col.data()));
      check(obgpu_download(ctx_, OBGPU_RES_KFF_CSR_VAL, out.K_ff_val.data()));
    }
    if (z[4] <= 4096) {
      out.K_ff.assign(size_t(z[5] * z[5]), 0); out.K_bf.assign(size_t(z[6] * z[5]), 0); out.K_bb.assign(size_t(z[6] * z[6]), 0);
      if (!out.K_ff.empty()) check(obgpu_download(ctx_, OBGPU_RES_KFF_DENSE, out.K_ff.data()));
      if (!out.K_bf.empty()) check(obgpu_download(ctx_, OBGPU_RES_KBF_DENSE, out.K_bf.data()));
      if (!out.K_bb.empty()) check(obgpu_download(ctx_, OBGPU_RES_KBB_DENSE, out.K_bb.data()));
    }
  }

  void solveStatic(StaticSolveProblemInfo& out, const StaticSolverOptions& opts = StaticSolverOptions()) {
    if (opts.algorithm == StaticSolverAlgorithm::Dense_LLT || opts.algorithm == StaticSolverAlgorithm::Sparse_LLT)
      throw std::runtime_error("Dense_LLT/Sparse_LLT are the reference's CPU algorithms; libobgpu has no CPU fallback");
    if (!opts.nonLinearIterative) {
      solve_pass(opts);
    } else {
      // solver_static.cpp:55-58,63-66,139-193: the loop sits above assemble+solve, so above the C ABI.  Every pass
      // re-assembles and re-solves on the GPU with the node poses moved by the previous pass's U_f (a translation adds
      // to the global coordinate as it is, RZ composes into the node's frame angle); no convergence test, all 40
      // passes run, the poses are restored, and the results are those of the last pass.
      const std::vector<double> xy0 = node_xy_, rz0 = node_rotz_;
      try {
        for (unsigned nIter = 0; nIter < 40; nIter++) {
          // solver_static.cpp:190-193 restores the poses BEFORE the caller's postProcCalcStress: the last pass's stresses
          // take their element lengths from the original poses, everything else from the pass itself
          solve_pass(opts, nIter == 39 ? xy0.data() : nullptr);
          const auto z = sizes();
          std::vector<int32_t> f(z[5]), nd(z[4]), id(z[4]);
          std::vector<num_t> U(z[5]);
          check(obgpu_get_symbolic(ctx_, OBGPU_SYM_FREE, f.data()));
          check(obgpu_get_symbolic(ctx_, OBGPU_SYM_DOF_NODE, nd.data()));
          check(obgpu_get_symbolic(ctx_, OBGPU_SYM_DOF_ID, id.data()));
          check(obgpu_download(ctx_, OBGPU_RES_UF, U.data()));
          if (nIter == 39) break;       // the device keeps the last pass for the downloads below and for stress
          for (size_t i = 0; i < f.size(); i++) {
            const int node = nd[f[i]], dof = id[f[i]];
            if (dof < 2) node_xy_[2 * node + dof] += U[i];
            else if (dof == 5) node_rotz_[node] += U[i];
          }
          dirty_ = true;
        }
      } catch (...) {
        node_xy_ = xy0; node_rotz_ = rz0; dirty_ = true;
        throw;
      }
      node_xy_ = xy0; node_rotz_ = rz0;
      stale_ = true;                    // the next updateAll() uploads the original poses again
    }
    const auto z = sizes();
    fill_lists(out.build_info);
    out.build_info.U_b.resize(z[6]); out.build_info.F_f.resize(z[5]);
    out.U_f.resize(z[5]); out.F_b.resize(z[6]);
    check(obgpu_download(ctx_, OBGPU_RES_UB, out.build_info.U_b.data()));
    check(obgpu_download(ctx_, OBGPU_RES_FF, out.build_info.F_f.data()));
    check(obgpu_download(ctx_, OBGPU_RES_UF, out.U_f.data()));
    check(obgpu_download(ctx_, OBGPU_RES_FB, out.F_b.data()));
  }

  void postProcCalcStress(StressInfo& out, const StaticSolveProblemInfo& /*solver_info*/) {
    check(obgpu_stress(ctx_));
    const size_t E = getNumberOfElements();
    // compact records (N, Vy, Mz per face): Vz, Mx, My are structurally zero for 2D elements
    std::vector<num_t> s(E * 6);
    check(obgpu_download(ctx_, OBGPU_RES_STRESS_COMPACT, s.data()));
    out.element_stress.assign(E, ElementStress(2));
    for (size_t e = 0; e < E; e++)
      for (int f = 0; f < 2; f++) {
        const num_t* v = &s[e * 6 + f * 3];
        out.element_stress[e][f] = FaceStress{v[0], v[1], 0, 0, 0, v[2]};
      }
  }

  obgpu_ctx* ctx() { return ctx_; }

 private:
  void solve_pass(const StaticSolverOptions& opts, const double* stress_length_poses = nullptr) {
    updateAll();
    if (stress_length_poses) check(obgpu_set_stress_length_poses(ctx_, stress_length_poses));
    obgpu_solve_opts o;
    o.method = opts.algorithm == StaticSolverAlgorithm::GPU_PCG_Jacobi ? OBGPU_PCG_JACOBI
             : opts.algorithm == StaticSolverAlgorithm::GPU_PCG_MG   ? OBGPU_PCG_MG
                                                                      : OBGPU_BATCHED_LDLT;
    o.with_stress = 1; o.rtol = opts.pcg_rtol; o.max_iter = opts.pcg_max_iter; o.flags = OBGPU_SOLVE_COMPACT_STRESS;
    check(obgpu_solve_static(ctx_, &o));
    const auto z = sizes();
    std::vector<int32_t> info(z[0] * z[1]);
    if (obgpu_download(ctx_, OBGPU_RES_INFO, info.data()) == OBGPU_ESINGULAR)
      throw std::runtime_error(obgpu_status_string(OBGPU_ESINGULAR));  // solver_static.cpp:102-104
  }
  void check(int rc) {
    if (rc != OBGPU_OK) {
      const char* m = obgpu_last_error(ctx_);
      throw std::runtime_error((m && *m) ? m : obgpu_status_string(rc));
    }
  }
  std::array<int64_t, OBGPU_NSIZES> sizes() {
    std::array<int64_t, OBGPU_NSIZES> z{};
    check(obgpu_get_sizes(ctx_, z.data()));
    return z;
  }
  struct Offsets {  // the [S+1] offset arrays of a one-structure batch
    int64_t n[2], e[2], c[2], l[2], q[2];
    explicit Offsets(const CStructureProblem& p)
        : n{0, int64_t(p.getNumberOfNodes())}, e{0, int64_t(p.getNumberOfElements())}, c{0, int64_t(p.con_node_.size())},
          l{0, int64_t(p.nl_node_.size())}, q{0, int64_t(p.el_elem_.size())} {}
  };
  obgpu_batch_desc describe(const Offsets& off) const {
    obgpu_batch_desc d{};
    d.n_structures = 1;
    d.node_off = off.n; d.elem_off = off.e; d.con_off = off.c; d.nload_off = off.l; d.eload_off = off.q;
    d.node_xy = node_xy_.data();
    bool rot = false, kinds = false;
    for (num_t r : node_rotz_) rot = rot || r != 0;
    for (uint8_t k : el_kind_) kinds = kinds || k != OBGPU_LOAD_UNIFORM;
    d.node_rotz = rot ? node_rotz_.data() : nullptr;
    d.elem_nodes = elem_nodes_.data(); d.elem_type = elem_type_.data(); d.elem_EAI = elem_EAI_.data();
    d.con_node = con_node_.data(); d.con_dof = con_dof_.data(); d.con_val = con_val_.data();
    d.nload_node = nl_node_.data(); d.nload_dof = nl_dof_.data(); d.nload_val = nl_val_.data();
    d.eload_elem = el_elem_.data(); d.eload_q = el_q_.data(); d.eload_dir = el_dir_.data();
    d.eload_type = kinds ? el_kind_.data() : nullptr;
    d.eload_q2 = kinds ? el_q2_.data() : nullptr;
    d.n_cases = 1;
    return d;
  }
  void upload() {
    if (!dirty_) return;
    Offsets off(*this);
    obgpu_batch_desc d = describe(off);
    check(obgpu_batch_upload(ctx_, &d));
    dirty_ = false;
  }
  void fill_lists(BuildProblemInfo& bi) {
    const auto z = sizes();
    std::vector<int32_t> f(z[5]), b(z[6]);
    check(obgpu_get_symbolic(ctx_, OBGPU_SYM_FREE, f.data()));
    check(obgpu_get_symbolic(ctx_, OBGPU_SYM_BOUNDED, b.data()));
    bi.free_dof_indices.assign(f.begin(), f.end());
    bi.bounded_dof_indices.assign(b.begin(), b.end());
    bi.dof_types.assign(z[4], TDoFType());
    for (size_t i = 0; i < f.size(); i++) bi.dof_types[f[i]].free_index = i;
    for (size_t i = 0; i < b.size(); i++) bi.dof_types[b[i]].bounded_index = i;
    bi.K_outer.resize(z[4] + 1);
    bi.K_inner.resize(z[8]);
    check(obgpu_get_symbolic(ctx_, OBGPU_SYM_K_ROWPTR, bi.K_outer.data()));
    check(obgpu_get_symbolic(ctx_, OBGPU_SYM_K_COL, bi.K_inner.data()));
  }

  obgpu_ctx* ctx_ = nullptr;
  bool dirty_ = true, stale_ = false;
  std::vector<double> node_xy_, node_rotz_, elem_EAI_, con_val_, nl_val_, el_q_, el_q2_, el_dir_;
  std::vector<uint32_t> elem_nodes_, con_node_, nl_node_, el_elem_;
  std::vector<uint8_t> elem_type_, con_dof_, nl_dof_, el_kind_;
};

}  // namespace obhost
