// symbolic.cpp — K3, the integer "symbolic phase" of libobgpu (host, once per topology).
//
// Produces what the reference rebuilds on every assembleProblem() call:
//   * DOF numbering      (updateNodeConnections + updateListDoFs, src/CFiniteElementProblem.cpp:179-245,
//                         with the per-end global DOF masks of CElement::getGlobalDoFs, src/CElement.cpp:192-247)
//   * free/bounded split (src/CFiniteElementProblem_assemble.cpp:74-99) and U_b (:304-313)
//   * the pattern of K   (upper triangle, Eigen CSC order == CSR of the lower triangle; :123-233)
//   * the element->nnz map that turns the reference's triplet scatter into a deterministic
//     segmented reduction (sources listed per nnz in the reference's triplet order).
// Integer results must match the reference bit-exactly (tests compare them with the oracle).
#include <algorithm>
#include <cmath>
#include <cstring>

#include "obgpu_internal.h"

namespace obgpu {

// 6x6 index -> compact index inside {DX,DY,RZ} (0,1,5 -> 0,1,2), -1 otherwise
static inline int c3(int d) { return d == 0 ? 0 : d == 1 ? 1 : d == 5 ? 2 : -1; }
// compact id inside ELEM_K of entry (r,c) of block j (0=K11, 1=K22, 2=K12); -1 = structural zero
static inline int elemk_cid(int j, int r, int c) {
  const int a = c3(r), b = c3(c);
  if (a < 0 || b < 0) return -1;
  if (j == 2) return 12 + a * 3 + b;
  static const int up[3][3] = {{0, 1, 2}, {1, 3, 4}, {2, 4, 5}};
  return (j == 0 ? 0 : 6) + up[a][b];
}

// Global DOF masks of the two element ends (bit d = DOF d used), CElement.cpp:192-247 specialised to
// the four beam types' local masks (getLocalDoFs: RR.cpp:129-139, RA.cpp:110-119, AR.cpp:110-119, AA.cpp:97-106).
// Orientation as internals.cpp:30-123 with z = 0: X = p12 * (1/len), Z = (0,0,1), Y = Z x X.
static bool element_end_masks(const double* p1, const double* p2, int type, const double* par, unsigned& m0, unsigned& m1) {
  const double dx = p2[0] - p1[0], dy = p2[1] - p1[1], dz = 0.0;
  const double len = std::sqrt(dx * dx + dy * dy + dz * dz);
  double R[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
  if (len > 0) {
    const double inv = 1 / len;
    const double X0 = dx * inv, X1 = dy * inv, X2 = dz * inv;
    const double Y0 = 0 * X2 - 1 * X1, Y1 = 1 * X0 - 0 * X2, Y2 = 0 * X1 - 0 * X0;
    R[0][0] = X0; R[0][1] = Y0; R[0][2] = 0;
    R[1][0] = X1; R[1][1] = Y1; R[1][2] = 0;
    R[2][0] = X2; R[2][1] = Y2; R[2][2] = 1;
  }
  bool nz[3][3];
  for (int r = 0; r < 3; r++)
    for (int c = 0; c < 3; c++) nz[r][c] = std::fabs(R[r][c]) > 1e-10;
  auto glob = [&](unsigned local) {
    unsigned g = 0;
    for (int blk = 0; blk < 2; blk++) {
      const unsigned l = (local >> (3 * blk)) & 7u;
      unsigned o = 0;
      if (l == 7u) {
        o = 7u;
      } else {
        for (int gi = 0; gi < 3; gi++)
          if (((l & 1u) && nz[gi][0]) || ((l & 2u) && nz[gi][1]) || ((l & 4u) && nz[gi][2])) o |= 1u << gi;
      }
      g |= o << (3 * blk);
    }
    return g;
  };
  const unsigned R_ = 0x23 /*DX,DY,RZ*/, A2 = 0x03 /*DX,DY*/, A1 = 0x01 /*DX*/;
  unsigned l0, l1;
  switch (type) {
    case OBGPU_BEAM2D_RR: l0 = R_; l1 = R_; break;
    case OBGPU_BEAM2D_RA: l0 = R_; l1 = A2; break;
    case OBGPU_BEAM2D_AR: l0 = A2; l1 = R_; break;
    case OBGPU_BEAM2D_AA: l0 = A1; l1 = A1; break;
    case OBGPU_BEAM2D_RD: l0 = R_; l1 = 0x21; break;                                     // RD.cpp:114-125
    case OBGPU_SPRING_1D: l0 = l1 = A1; break;                                           // CElementSpring.cpp:69-75
    case OBGPU_SPRING_XY: l0 = l1 = (par[0] != 0 ? 1u : 0u) | (par[1] != 0 ? 2u : 0u); break;  // SpringXY.cpp:62-66
    case OBGPU_SPRING_DXDYRZ:
      l0 = l1 = (par[0] != 0 ? 1u : 0u) | (par[1] != 0 ? 2u : 0u) | (par[2] != 0 ? 0x20u : 0u);
      break;                                                                             // SpringDXDYRZ.cpp:65-70
    case OBGPU_SPRING_TORSION: l0 = l1 = 0x20; break;                                    // TorsionSpring.cpp:70-75
    default: return false;
  }
  m0 = glob(l0);
  m1 = glob(l1);
  return true;
}

int run_symbolic(const HostBatch& hb, Symbolic& sym, std::string& err, bool want_pattern) {
  const int64_t S = hb.S;
  sym = Symbolic();
  sym.dof_off.assign(S + 1, 0);
  sym.free_off.assign(S + 1, 0);
  sym.bnd_off.assign(S + 1, 0);
  sym.nnz_off.assign(S + 1, 0);
  sym.tri_off.assign(S + 1, 0);
  sym.dof_index.assign(size_t(6) * hb.Ntot(), -1);
  sym.seg_ptr.assign(1, 0);
  sym.node_slot.assign(size_t(3) * hb.Ntot(), SLOT_NONE);
  sym.nload_slot.assign(hb.NLtot(), SLOT_NONE);
  sym.el_ptr.assign(hb.Etot() + 1, int32_t(hb.ELtot()));
  sym.el_order.assign(hb.ELtot(), 0);
  sym.half_bw.assign(S, 0);
  sym.has_pattern = want_pattern;
  std::vector<unsigned> node_mask;
  std::vector<int32_t> t_hi, t_lo;   // per-structure triplets: CSC(upper) column / row
  std::vector<int32_t> colstart, bucket, colrows;
  std::vector<int32_t> order;

  for (int64_t s = 0; s < S; s++) {
    const int64_t n0 = hb.node_off[s], N = hb.node_off[s + 1] - n0;
    const int64_t e0 = hb.elem_off[s], E = hb.elem_off[s + 1] - e0;
    // ---- node DOF masks (updateNodeConnections) ----
    node_mask.assign(N, 0u);
    for (int64_t e = 0; e < E; e++) {
      const uint32_t a = hb.elem_nodes[2 * (e0 + e)], b = hb.elem_nodes[2 * (e0 + e) + 1];
      if (a >= N || b >= N) {
        err = "element node index out of range (ASSERT_(conected_nodes_ids[k] < nNodes))";
        return OBGPU_EINVAL;
      }
      unsigned m0, m1;
      if (!element_end_masks(&hb.node_xy[2 * (n0 + a)], &hb.node_xy[2 * (n0 + b)], hb.elem_type[e0 + e], &hb.elem_EAI[3 * (e0 + e)], m0, m1)) {
        err = "unknown element type";
        return OBGPU_EINVAL;
      }
      node_mask[a] |= m0;
      node_mask[b] |= m1;
    }
    // ---- numbering (updateListDoFs): node-major, dof-minor ----
    const int64_t k0 = sym.dof_off[s];
    int32_t n = 0;
    for (int64_t i = 0; i < N; i++)
      for (int d = 0; d < 6; d++)
        if (node_mask[i] & (1u << d)) {
          sym.dof_node.push_back(int32_t(i));
          sym.dof_id.push_back(d);
          sym.dof_index[6 * (n0 + i) + d] = n++;
        }
    if (n == 0) {
      err = "structure without DOFs (ASSERT_(nDOFs > 0))";
      return OBGPU_EINVAL;
    }
    sym.dof_off[s + 1] = k0 + n;
    // ---- constraints -> partition (assemble.cpp:74-99) and U_b (:304-313) ----
    std::vector<char> bounded(n, 0);
    std::vector<double> cval(n, 0.0);
    for (int64_t c = hb.con_off[s]; c < hb.con_off[s + 1]; c++) {
      if (hb.con_node[c] >= N || hb.con_dof[c] > 5) {
        err = "constraint node/dof out of range";
        return OBGPU_EINVAL;
      }
      const int32_t k = sym.dof_index[6 * (n0 + hb.con_node[c]) + hb.con_dof[c]];
      if (k >= 0) {
        bounded[k] = 1;
        cval[k] = hb.con_val[c];  // insertConstraint overwrites
      }
    }
    int32_t nf = 0, nb = 0;
    for (int32_t k = 0; k < n; k++) {
      if (bounded[k]) {
        sym.bnd_idx.push_back(k);
        sym.U_b.push_back(cval[k]);
        sym.dof_type.push_back(-(nb++) - 1);
      } else {
        sym.free_idx.push_back(k);
        sym.dof_type.push_back(nf++);
      }
    }
    sym.free_off[s + 1] = sym.free_off[s] + nf;
    sym.bnd_off[s + 1] = sym.bnd_off[s] + nb;
    for (int64_t l = hb.nload_off[s]; l < hb.nload_off[s + 1]; l++)
      if (hb.nload_node[l] >= N || hb.nload_dof[l] > 5) {
        err = "node load node/dof out of range";
        return OBGPU_EINVAL;
      }
    for (int64_t l = hb.eload_off[s]; l < hb.eload_off[s + 1]; l++)
      if (hb.eload_elem[l] >= E) {
        err = "element load element index out of range";
        return OBGPU_EINVAL;
      }
    // ---- compact planar slots, element-load CSR, half bandwidth (inputs of the batched-LDLt kernels) ----
    {
      static const int d3[3] = {0, 1, 5};
      for (int64_t i = 0; i < N; i++)
        for (int a = 0; a < 3; a++) {
          const int32_t k = sym.dof_index[6 * (n0 + i) + d3[a]];
          sym.node_slot[3 * (n0 + i) + a] = k < 0 ? SLOT_NONE : sym.dof_type[k0 + k];
        }
      for (int64_t l = hb.nload_off[s]; l < hb.nload_off[s + 1]; l++) {
        const int32_t k = sym.dof_index[6 * (n0 + hb.nload_node[l]) + hb.nload_dof[l]];
        sym.nload_slot[l] = k < 0 ? SLOT_NONE : sym.dof_type[k0 + k];
      }
      // assemble.cpp:349-435: prescribed displacements of a node with a rotated frame go to nodal coordinates,
      // U' = R^T (Ux, Uy, .) with only the BOUNDED components taken (free ones count as 0).  Input preparation of
      // BuildProblemInfo::U_b, once per topology, a handful of supports.
      if (hb.has_rotz)
        for (int64_t i = 0; i < N; i++) {
          const double r = hb.node_rotz[n0 + i];
          if (r == 0.0) continue;
          const int32_t tx = sym.node_slot[3 * (n0 + i)], ty = sym.node_slot[3 * (n0 + i) + 1];
          const bool bx = tx < 0 && tx != SLOT_NONE, by = ty < 0 && ty != SLOT_NONE;
          if (!bx && !by) continue;
          const size_t ub0 = size_t(sym.bnd_off[s]);
          const double ux = bx ? sym.U_b[ub0 + (-tx - 1)] : 0.0, uy = by ? sym.U_b[ub0 + (-ty - 1)] : 0.0;
          const double cr = std::cos(r), sr = std::sin(r);
          if (bx) sym.U_b[ub0 + (-tx - 1)] = cr * ux + sr * uy;
          if (by) sym.U_b[ub0 + (-ty - 1)] = -sr * ux + cr * uy;
        }
      const int64_t l0 = hb.eload_off[s], EL = hb.eload_off[s + 1] - l0;
      order.resize(EL);
      for (int64_t i = 0; i < EL; i++) order[i] = int32_t(i);
      std::stable_sort(order.begin(), order.end(),
                       [&](int32_t a, int32_t b) { return hb.eload_elem[l0 + a] < hb.eload_elem[l0 + b]; });
      int64_t p = 0;
      for (int64_t e = 0; e < E; e++) {
        sym.el_ptr[e0 + e] = int32_t(l0 + p);
        while (p < EL && int64_t(hb.eload_elem[l0 + order[p]]) == e) {
          sym.el_order[l0 + p] = int32_t(l0 + order[p]);
          p++;
        }
      }
      int32_t bw = 0;
      for (int64_t e = 0; e < E; e++) {
        int32_t lo = INT32_MAX, hi = -1;
        for (int f = 0; f < 2; f++)
          for (int a = 0; a < 3; a++) {
            const int32_t t = sym.node_slot[3 * (n0 + hb.elem_nodes[2 * (e0 + e) + f]) + a];
            if (t >= 0) { lo = std::min(lo, t); hi = std::max(hi, t); }
          }
        if (hi >= 0) bw = std::max(bw, hi - lo);
      }
      sym.half_bw[s] = bw;
    }
    if (!want_pattern) {
      sym.nnz_off[s + 1] = sym.nnz_off[s];
      sym.tri_off[s + 1] = sym.tri_off[s];
      continue;
    }
    // ---- triplets in the reference's emission order (assemble.cpp:123-223) ----
    // (column hi, row lo) of CSC(upper) per triplet; the pattern is then built by a counting sort over the columns and a
    // small sorted-unique pass inside each column (a column holds a few dozen entries at most): O(T), no global sort
    t_hi.clear();
    t_lo.clear();
    const size_t t0 = sym.tri_src.size();
    t_hi.reserve(size_t(21) * E);
    t_lo.reserve(size_t(21) * E);
    sym.tri_src.reserve(t0 + size_t(21) * E);
    sym.k_col.reserve(sym.k_col.size() + size_t(6) * N + size_t(9) * E);
    sym.seg_ptr.reserve(sym.seg_ptr.size() + size_t(6) * N + size_t(9) * E);
    for (int64_t e = 0; e < E; e++) {
      const uint32_t nd[2] = {hb.elem_nodes[2 * (e0 + e)], hb.elem_nodes[2 * (e0 + e) + 1]};
      static const int bi[3] = {0, 1, 0}, bj[3] = {0, 1, 1};
      for (int j = 0; j < 3; j++) {
        const uint32_t in = nd[bi[j]], jn = nd[bj[j]];
        const int32_t* ri = &sym.dof_index[6 * (n0 + in)];
        const int32_t* ci = &sym.dof_index[6 * (n0 + jn)];
        for (int r = 0; r < 6; r++) {
          if (ri[r] < 0) continue;
          for (int c = (in == jn ? r : 0); c < 6; c++) {
            if (ci[c] < 0) continue;
            t_lo.push_back(std::min(ri[r], ci[c]));   // CSC(upper): column = hi, row = lo
            t_hi.push_back(std::max(ri[r], ci[c]));
            const int cid = elemk_cid(j, r, c);
            sym.tri_src.push_back(cid < 0 ? -1 : int32_t(e * ELEM_K + cid));
          }
        }
      }
    }
    const size_t T = t_hi.size();
    sym.tri_off[s + 1] = sym.tri_off[s] + int64_t(T);
    // ---- pattern: per column the sorted unique rows == Eigen's compressed CSC after setFromTriplets ----
    colstart.assign(size_t(n) + 1, 0);
    for (size_t t = 0; t < T; t++) colstart[size_t(t_hi[t]) + 1]++;
    for (int32_t k = 0; k < n; k++) colstart[size_t(k) + 1] += colstart[size_t(k)];
    bucket.resize(T);
    {
      std::vector<int32_t> fill(colstart.begin(), colstart.end() - 1);
      for (size_t t = 0; t < T; t++) bucket[size_t(fill[size_t(t_hi[t])]++)] = int32_t(t);   // stable: triplet order kept
    }
    const int64_t z0 = sym.nnz_off[s];
    const size_t rp0 = sym.k_rowptr.size();
    sym.k_rowptr.resize(rp0 + n + 1, 0);
    const size_t tn0 = sym.tri_nnz.size();
    sym.tri_nnz.resize(tn0 + T);
    int32_t nnz32 = 0;
    for (int32_t hi = 0; hi < n; hi++) {
      colrows.clear();
      for (int32_t p = colstart[size_t(hi)]; p < colstart[size_t(hi) + 1]; p++) colrows.push_back(t_lo[size_t(bucket[size_t(p)])]);
      std::sort(colrows.begin(), colrows.end());
      colrows.erase(std::unique(colrows.begin(), colrows.end()), colrows.end());
      for (int32_t p = colstart[size_t(hi)]; p < colstart[size_t(hi) + 1]; p++) {
        const int32_t t = bucket[size_t(p)];
        const int32_t at = int32_t(std::lower_bound(colrows.begin(), colrows.end(), t_lo[size_t(t)]) - colrows.begin());
        sym.tri_nnz[tn0 + size_t(t)] = nnz32 + at;
      }
      for (int32_t lo : colrows) sym.k_col.push_back(lo);
      nnz32 += int32_t(colrows.size());
      sym.k_rowptr[rp0 + size_t(hi) + 1] = nnz32;
    }
    const int64_t nnz = nnz32;
    sym.nnz_off[s + 1] = z0 + nnz;
    // ---- per-slot source lists in triplet order (stable counting sort) ----
    std::vector<int64_t> cnt(nnz + 1, 0);
    for (size_t t = 0; t < T; t++) cnt[sym.tri_nnz[tn0 + t] + 1]++;
    for (int64_t g = 0; g < nnz; g++) cnt[g + 1] += cnt[g];
    const int64_t sp0 = sym.seg_ptr.back();
    for (int64_t g = 1; g <= nnz; g++) sym.seg_ptr.push_back(sp0 + cnt[g]);
    const size_t ss0 = sym.seg_src.size();
    sym.seg_src.resize(ss0 + T);
    std::vector<int64_t> fill(cnt.begin(), cnt.end() - 1);
    for (size_t t = 0; t < T; t++) {
      const int32_t src = sym.tri_src[t0 + t];
      sym.seg_src[ss0 + fill[sym.tri_nnz[tn0 + t]]++] = src < 0 ? -1 : int64_t(e0) * ELEM_K + src;
    }
  }
  sym.valid = true;
  return OBGPU_OK;
}

}  // namespace obgpu
