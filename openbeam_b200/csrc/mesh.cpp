// mesh.cpp — CStructureProblem::mesh() for a batch description: the step right before the path,
// and how real inputs reach large n (the web app meshes at 0.10 m).  Host-side on purpose: it is O(N) integer and
// coordinate generation done once per topology, and its product feeds the host symbolic phase (K3) anyway.
//
// Follows, per structure:
//   CStructureProblem_mesh.cpp:37-132   original nodes first, then the intermediate nodes element by element;
//                                       constraints and nodal loads are copied (they sit on original nodes);
//                                       element loads are split by the load's own meshLoad
//   CBaseElementBeam.cpp:317-422        nElements = ceil(L / max_element_length), nodes at node0 + v01 * k / nElements
//                                       with node0's frame; piece types: first AR for AR/AA else RR, last RA for RA/AA,
//                                       RD for RD, else RR, middle RR
//   CElementSpring*.cpp do_mesh         springs mesh to themselves
//   CLoadDistributedUniform.cpp:174-185, CLoadConstTemperature.cpp:104-115   identical load on every piece
//   CLoadDistributedTriangular.cpp:186-202                                    trapezoid slices q_ini + i*Aq
//   CLoadConcentratedForce.cpp:170-193                                        one piece, dist reduced by idx * Al
// Reference behaviours kept as they are:
//   QM1  a beam that is not subdivided (one piece) takes the FIRST-piece rule only: RA and RD become RR, AA becomes AR
//        (CBaseElementBeam.cpp:381-395: `idx_el == 0` is tested before `idx_el == nElements - 1`).
//   QM2  the concentrated load is attached to meshed element number `idx_affected_element`, the index of the piece
//        WITHIN its beam, not meshed_element_idxs[idx] (CLoadConcentratedForce.cpp:192) — right only for beam 0.
#include <cmath>
#include <cstdint>
#include <new>
#include <vector>

#include "../../include/obgpu.h"

struct obgpu_mesh {
  std::vector<int64_t> node_off, elem_off, con_off, nload_off, eload_off, e2e_off, e2n_off;
  std::vector<double> node_xy, node_rotz, elem_EAI, con_val, nload_val, eload_q, eload_dir, eload_q2;
  std::vector<uint32_t> elem_nodes, con_node, nload_node, eload_elem, e2e, e2n;
  std::vector<uint8_t> elem_type, con_dof, nload_dof, eload_type;
  bool any_rot = false, any_kind = false;
};

extern "C" {

int obgpu_mesh_create(const obgpu_batch_desc* in, double max_element_length, obgpu_mesh** out) {
  if (!in || !out || in->n_structures < 1 || !(max_element_length > 0)) return OBGPU_EINVAL;
  obgpu_mesh* m = new (std::nothrow) obgpu_mesh;
  if (!m) return OBGPU_EINVAL;
  const int64_t S = in->n_structures;
  m->node_off.push_back(0); m->elem_off.push_back(0); m->con_off.push_back(0);
  m->nload_off.push_back(0); m->eload_off.push_back(0); m->e2e_off.push_back(0); m->e2n_off.push_back(0);
  for (int64_t s = 0; s < S; s++) {
    const int64_t n0 = in->node_off[s], N = in->node_off[s + 1] - n0;
    const int64_t e0 = in->elem_off[s], E = in->elem_off[s + 1] - e0;
    const size_t node_base = m->node_xy.size() / 2, elem_base = m->elem_type.size();
    // (1) original nodes keep their ids
    for (int64_t i = 0; i < N; i++) {
      m->node_xy.push_back(in->node_xy[2 * (n0 + i)]);
      m->node_xy.push_back(in->node_xy[2 * (n0 + i) + 1]);
      const double r = in->node_rotz ? in->node_rotz[n0 + i] : 0.0;
      m->node_rotz.push_back(r);
      if (r != 0.0) m->any_rot = true;
    }
    // (2) intermediate nodes and pieces, element by element
    std::vector<int64_t> piece_off(size_t(E) + 1, 0);
    for (int64_t e = 0; e < E; e++) {
      const uint32_t na = in->elem_nodes[2 * (e0 + e)], nb = in->elem_nodes[2 * (e0 + e) + 1];
      const uint8_t type = in->elem_type[e0 + e];
      if (int64_t(na) >= N || int64_t(nb) >= N) { delete m; return OBGPU_EINVAL; }
      const double x0 = in->node_xy[2 * (n0 + na)], y0 = in->node_xy[2 * (n0 + na) + 1];
      const double vx = in->node_xy[2 * (n0 + nb)] - x0, vy = in->node_xy[2 * (n0 + nb) + 1] - y0;
      size_t pieces = 1;
      if (type <= OBGPU_BEAM2D_RD) {
        const double L = std::sqrt(vx * vx + vy * vy + 0.0);  // TPoint3D::norm with z = 0
        if (!(L > 0)) { delete m; return OBGPU_EINVAL; }      // ASSERT_(L > 0)
        pieces = static_cast<size_t>(std::ceil(L / max_element_length));
      }
      std::vector<uint32_t> my_nodes;
      my_nodes.push_back(na);
      for (size_t k = 1; k < pieces; k++) {
        const double f = double(k) / double(pieces);
        my_nodes.push_back(uint32_t(m->node_xy.size() / 2 - node_base));
        m->node_xy.push_back(x0 + vx * f);
        m->node_xy.push_back(y0 + vy * f);
        const double r = in->node_rotz ? in->node_rotz[n0 + na] : 0.0;  // p.r = node0.r
        m->node_rotz.push_back(r);
      }
      my_nodes.push_back(nb);
      for (size_t k = 0; k < pieces; k++) {
        uint8_t t = type;
        if (type <= OBGPU_BEAM2D_RD) {
          if (k == 0) t = (type == OBGPU_BEAM2D_AR || type == OBGPU_BEAM2D_AA) ? OBGPU_BEAM2D_AR : OBGPU_BEAM2D_RR;  // QM1
          else if (k == pieces - 1)
            t = (type == OBGPU_BEAM2D_RA || type == OBGPU_BEAM2D_AA) ? OBGPU_BEAM2D_RA
                : type == OBGPU_BEAM2D_RD                            ? OBGPU_BEAM2D_RD
                                                                     : OBGPU_BEAM2D_RR;
          else t = OBGPU_BEAM2D_RR;
        }
        m->e2e.push_back(uint32_t(m->elem_type.size() - elem_base));
        m->elem_nodes.push_back(my_nodes[k]);
        m->elem_nodes.push_back(my_nodes[k + 1]);
        m->elem_type.push_back(t);
        for (int j = 0; j < 3; j++) m->elem_EAI.push_back(in->elem_EAI[3 * (e0 + e) + j]);  // copyCommonBeamParamsFrom
      }
      for (uint32_t v : my_nodes) m->e2n.push_back(v);
      piece_off[size_t(e) + 1] = piece_off[size_t(e)] + int64_t(pieces);
      m->e2e_off.push_back(int64_t(m->e2e.size()));
      m->e2n_off.push_back(int64_t(m->e2n.size()));
    }
    // (3) constraints and nodal loads: same node ids
    for (int64_t i = in->con_off[s]; i < in->con_off[s + 1]; i++) {
      m->con_node.push_back(in->con_node[i]); m->con_dof.push_back(in->con_dof[i]); m->con_val.push_back(in->con_val[i]);
    }
    for (int64_t i = in->nload_off[s]; i < in->nload_off[s + 1]; i++) {
      m->nload_node.push_back(in->nload_node[i]); m->nload_dof.push_back(in->nload_dof[i]);
      m->nload_val.push_back(in->nload_val[i]);
    }
    // (4) element loads, in m_loads_on_beams order (multimap keyed by the element index: ascending, stable)
    const int64_t l0 = in->eload_off[s], l1 = in->eload_off[s + 1];
    std::vector<int64_t> order;
    for (int64_t e = 0; e < E; e++)
      for (int64_t i = l0; i < l1; i++)
        if (int64_t(in->eload_elem[i]) == e) order.push_back(i);
    auto push_load = [&](uint32_t elem, uint8_t kind, double q, double q2, const double* dir) {
      m->eload_elem.push_back(elem); m->eload_type.push_back(kind); m->eload_q.push_back(q); m->eload_q2.push_back(q2);
      for (int j = 0; j < 3; j++) m->eload_dir.push_back(dir[j]);
      if (kind != OBGPU_LOAD_UNIFORM) m->any_kind = true;
    };
    for (int64_t i : order) {
      const int64_t e = in->eload_elem[i];
      const uint8_t kind = in->eload_type ? in->eload_type[i] : uint8_t(OBGPU_LOAD_UNIFORM);
      const double q = in->eload_q[i], q2 = in->eload_q2 ? in->eload_q2[i] : 0.0;
      const double* dir = in->eload_dir + 3 * i;
      const int64_t p0 = piece_off[size_t(e)], nE = piece_off[size_t(e) + 1] - p0;
      if (kind == OBGPU_LOAD_UNIFORM || kind == OBGPU_LOAD_TEMPERATURE) {
        for (int64_t k = 0; k < nE; k++) push_load(uint32_t(p0 + k), kind, q, q2, dir);
      } else if (kind == OBGPU_LOAD_TRIANGULAR) {
        const double Aq = (q2 - q) / double(nE);
        for (int64_t k = 0; k < nE; k++) {
          const double qi = q + double(k) * Aq;
          push_load(uint32_t(p0 + k), kind, qi, qi + Aq, dir);
        }
      } else if (kind == OBGPU_LOAD_CONCENTRATED) {
        const uint32_t na = in->elem_nodes[2 * (e0 + e)], nb = in->elem_nodes[2 * (e0 + e) + 1];
        const double dx = in->node_xy[2 * (n0 + nb)] - in->node_xy[2 * (n0 + na)];
        const double dy = in->node_xy[2 * (n0 + nb) + 1] - in->node_xy[2 * (n0 + na) + 1];
        const double Al = std::sqrt(dx * dx + dy * dy) / double(nE);
        const double idx = std::fmin(std::floor(q2 / Al), double(nE - 1));
        push_load(uint32_t(idx) /* QM2: piece index used as the meshed element index */, kind, q, q2 - idx * Al, dir);
      } else {
        delete m;
        return OBGPU_EINVAL;
      }
    }
    m->node_off.push_back(int64_t(m->node_xy.size() / 2));
    m->elem_off.push_back(int64_t(m->elem_type.size()));
    m->con_off.push_back(int64_t(m->con_node.size()));
    m->nload_off.push_back(int64_t(m->nload_node.size()));
    m->eload_off.push_back(int64_t(m->eload_elem.size()));
  }
  *out = m;
  return OBGPU_OK;
}

int obgpu_mesh_desc(const obgpu_mesh* m, obgpu_batch_desc* d) {
  if (!m || !d) return OBGPU_EINVAL;
  *d = obgpu_batch_desc{};
  d->n_structures = int64_t(m->node_off.size()) - 1;
  d->node_off = m->node_off.data(); d->elem_off = m->elem_off.data(); d->con_off = m->con_off.data();
  d->nload_off = m->nload_off.data(); d->eload_off = m->eload_off.data();
  d->node_xy = m->node_xy.data(); d->node_rotz = m->any_rot ? m->node_rotz.data() : nullptr;
  d->elem_nodes = m->elem_nodes.data(); d->elem_type = m->elem_type.data(); d->elem_EAI = m->elem_EAI.data();
  d->con_node = m->con_node.data(); d->con_dof = m->con_dof.data(); d->con_val = m->con_val.data();
  d->nload_node = m->nload_node.data(); d->nload_dof = m->nload_dof.data(); d->nload_val = m->nload_val.data();
  d->eload_elem = m->eload_elem.data(); d->eload_q = m->eload_q.data(); d->eload_dir = m->eload_dir.data();
  d->n_cases = 1;
  d->eload_type = m->any_kind ? m->eload_type.data() : nullptr;
  d->eload_q2 = m->any_kind ? m->eload_q2.data() : nullptr;
  return OBGPU_OK;
}

int obgpu_mesh_info(const obgpu_mesh* m, const int64_t** e2e_off, const uint32_t** e2e, const int64_t** e2n_off,
                    const uint32_t** e2n) {
  if (!m) return OBGPU_EINVAL;
  if (e2e_off) *e2e_off = m->e2e_off.data();
  if (e2e) *e2e = m->e2e.data();
  if (e2n_off) *e2n_off = m->e2n_off.data();
  if (e2n) *e2n = m->e2n.data();
  return OBGPU_OK;
}

void obgpu_mesh_destroy(obgpu_mesh* m) { delete m; }

}  // extern "C"
