// symbolic_sparse.cpp — integer plan of the sparse route (host, once per topology):
//   * K4: per K slot the list of element-block entries to sum, in the reference's triplet order
//         (assemble.cpp:123-233, Eigen setFromTriplets sums duplicates in insertion order)
//   * K5: K_ff / K_bf as gathers out of K (assemble.cpp:249-297), K_ff expanded to full symmetric CSR for SpMV
//   * load vector: per DOF the nodal loads and element-load contributions to sum (assemble.cpp:317-341)
#include <algorithm>
#include <cstdlib>

#include "obgpu_internal.h"

namespace obgpu {

int build_sparse_plan(const HostBatch& hb, const Symbolic& sym, SparsePlan& P, std::string& err) {
  P = SparsePlan();
  if (!sym.valid || !sym.has_pattern) {
    err = "sparse plan needs the K pattern";
    return OBGPU_ESEQUENCE;
  }
  const int64_t S = hb.S, nnz = sym.nnz_tot(), T = sym.T_tot(), nf = sym.nf_tot(), nb = sym.nb_tot();
  if (T >= INT32_MAX || hb.Etot() * ELEM_K >= INT32_MAX || nnz * 2 >= INT32_MAX) {
    err = "batch too large for 32-bit sparse indices";
    return OBGPU_EUNSUPPORTED;
  }
  P.seg_ptr.resize(nnz + 1);
  P.seg_src.resize(T);
  for (int64_t i = 0; i <= nnz; i++) P.seg_ptr[i] = int32_t(sym.seg_ptr[i]);
  for (int64_t i = 0; i < T; i++) P.seg_src[i] = int32_t(sym.seg_src[i]);
  // ---- fused assembly plan: block-local element lists ----
  {
    const int64_t nblk = (nnz + SparsePlan::ASM_SLOTS - 1) / SparsePlan::ASM_SLOTS;
    std::vector<int32_t> stamp(hb.Etot(), -1), local(hb.Etot(), 0);
    P.blk_elem_ptr.assign(nblk + 1, 0);
    P.seg_loc.resize(T);
    bool ok = true;
    for (int64_t b = 0; b < nblk && ok; b++) {
      const int64_t g0 = b * SparsePlan::ASM_SLOTS, g1 = std::min<int64_t>(nnz, g0 + SparsePlan::ASM_SLOTS);
      int32_t nloc = 0;
      for (int64_t t = P.seg_ptr[g0]; t < P.seg_ptr[g1]; t++) {
        const int32_t src = P.seg_src[t];
        if (src < 0) {
          P.seg_loc[t] = 0xFFFF;
          continue;
        }
        const int32_t e = src / ELEM_K, cid = src - e * ELEM_K;
        if (stamp[e] != int32_t(b)) {
          stamp[e] = int32_t(b);
          local[e] = nloc++;
          P.blk_elems.push_back(e);
        }
        if (local[e] * ELEM_K + cid >= 0xFFFF) ok = false;
        P.seg_loc[t] = uint16_t(local[e] * ELEM_K + cid);
      }
      P.blk_elem_ptr[b + 1] = int32_t(P.blk_elems.size());
      P.asm_max_loc = std::max(P.asm_max_loc, nloc);
    }
    if (!ok || size_t(P.asm_max_loc) * ELEM_K * sizeof(double) > size_t(200) * 1024) {  // irregular: two-kernel path
      P.asm_max_loc = 0;
      P.blk_elem_ptr.clear(); P.blk_elems.clear(); P.seg_loc.clear();
    }
  }
  // ---- K_ff (full symmetric) and K_bf ----
  std::vector<int32_t> lowc(nf, 0), upc(nf, 0), blow(nb, 0), bup(nb, 0);
  auto for_each_entry = [&](auto&& fn) {
    for (int64_t s = 0; s < S; s++) {
      const int64_t k0 = sym.dof_off[s], n = sym.dof_off[s + 1] - k0, z0 = sym.nnz_off[s];
      const int32_t* rp = &sym.k_rowptr[k0 + s];
      const int64_t f0 = sym.free_off[s], b0 = sym.bnd_off[s];
      for (int64_t hi = 0; hi < n; hi++) {
        const int32_t thi = sym.dof_type[k0 + hi];
        for (int32_t z = rp[hi]; z < rp[hi + 1]; z++) {
          const int32_t lo = sym.k_col[z0 + z];
          const int32_t tlo = sym.dof_type[k0 + lo];
          fn(thi, tlo, hi == lo, f0, b0, int32_t(z0 + z));
        }
      }
    }
  };
  for_each_entry([&](int32_t thi, int32_t tlo, bool diag, int64_t f0, int64_t b0, int32_t) {
    if (thi >= 0 && tlo >= 0) {
      lowc[f0 + thi]++;
      if (!diag) upc[f0 + tlo]++;
    } else if (thi < 0 && tlo >= 0) {
      blow[b0 + (-thi - 1)]++;   // bounded row hi, free column lo (lo < hi)
    } else if (thi >= 0 && tlo < 0) {
      bup[b0 + (-tlo - 1)]++;    // bounded row lo, free column hi (hi > lo)
    }
  });
  P.ff_rowptr.assign(nf + 1, 0);
  for (int64_t r = 0; r < nf; r++) P.ff_rowptr[r + 1] = P.ff_rowptr[r] + lowc[r] + upc[r];
  P.bf_rowptr.assign(nb + 1, 0);
  for (int64_t r = 0; r < nb; r++) P.bf_rowptr[r + 1] = P.bf_rowptr[r] + blow[r] + bup[r];
  P.ff_col.resize(P.ff_rowptr[nf]);
  P.ff_src.resize(P.ff_rowptr[nf]);
  P.bf_col.resize(P.bf_rowptr[nb]);
  P.bf_src.resize(P.bf_rowptr[nb]);
  P.ff_diag.assign(nf, -1);
  std::vector<int32_t> curL(nf, 0), curU(nf, 0), bcurL(nb, 0), bcurU(nb, 0);
  for_each_entry([&](int32_t thi, int32_t tlo, bool diag, int64_t f0, int64_t b0, int32_t slot) {
    if (thi >= 0 && tlo >= 0) {
      const int64_t r = f0 + thi, cc = f0 + tlo;
      const int32_t p = P.ff_rowptr[r] + curL[r]++;
      P.ff_col[p] = int32_t(cc);
      P.ff_src[p] = slot;
      if (diag) {
        P.ff_diag[r] = p;
      } else {
        const int32_t q = P.ff_rowptr[cc] + lowc[cc] + curU[cc]++;
        P.ff_col[q] = int32_t(r);
        P.ff_src[q] = slot;
      }
    } else if (thi < 0 && tlo >= 0) {
      const int64_t r = b0 + (-thi - 1);
      const int32_t p = P.bf_rowptr[r] + bcurL[r]++;
      P.bf_col[p] = int32_t(f0 + tlo);
      P.bf_src[p] = slot;
    } else if (thi >= 0 && tlo < 0) {
      const int64_t r = b0 + (-tlo - 1);
      const int32_t p = P.bf_rowptr[r] + blow[r] + bcurU[r]++;
      P.bf_col[p] = int32_t(f0 + thi);
      P.bf_src[p] = slot;
    }
  });
  // ---- K_bb (upper: row = bounded index of lo, column = bounded index of hi), only read when some U_b != 0 ----
  {
    std::vector<int32_t> cnt(nb + 1, 0);
    for_each_entry([&](int32_t thi, int32_t tlo, bool, int64_t, int64_t b0, int32_t) {
      if (thi < 0 && tlo < 0) cnt[b0 + (-tlo - 1) + 1]++;
    });
    for (int64_t r = 0; r < nb; r++) cnt[r + 1] += cnt[r];
    P.bb_rowptr = cnt;
    P.bb_col.resize(cnt[nb]);
    P.bb_src.resize(cnt[nb]);
    std::vector<int32_t> cur(cnt.begin(), cnt.end() - 1);
    for_each_entry([&](int32_t thi, int32_t tlo, bool, int64_t, int64_t b0, int32_t slot) {
      if (thi < 0 && tlo < 0) {
        const int32_t p = cur[b0 + (-tlo - 1)]++;
        P.bb_col[p] = int32_t(b0 + (-thi - 1));
        P.bb_src[p] = slot;
      }
    });
  }
  for (int64_t r = 0; r < nf; r++)
    if (P.ff_diag[r] < 0) {
      err = "free DOF without a diagonal entry in K";
      return OBGPU_EINVAL;
    }
  // ---- K_ff for SpMV: SELL-32 (sliced ELLPACK).  Slice = 32 consecutive rows, entries column-major inside the
  // slice (entry k of row r at sl_ptr[r/32] + 32 k + r%32) and padded to the slice's longest row with
  // (col = own row, value 0): one warp per slice reads 256 contiguous bytes of values per step and the k loop has
  // no row-pointer dependency, so all of a row's loads are in flight together.  Order inside a row = CSR order.
  {
    const int64_t nsl = (nf + 31) / 32;
    P.sl_ptr.assign(nsl + 1, 0);
    for (int64_t sl = 0; sl < nsl; sl++) {
      int32_t w = 0;
      for (int64_t r = sl * 32; r < std::min<int64_t>(nf, sl * 32 + 32); r++) w = std::max(w, P.ff_rowptr[r + 1] - P.ff_rowptr[r]);
      const int64_t next = int64_t(P.sl_ptr[sl]) + int64_t(w) * 32;
      if (next >= INT32_MAX) {
        err = "K_ff too large for 32-bit SELL offsets";
        return OBGPU_EUNSUPPORTED;
      }
      P.sl_ptr[sl + 1] = int32_t(next);
    }
    std::vector<int32_t> col(P.sl_ptr[nsl]), src(P.sl_ptr[nsl], -1), diag(nf, -1);
    for (int64_t sl = 0; sl < nsl; sl++) {
      const int32_t base = P.sl_ptr[sl], w = (P.sl_ptr[sl + 1] - base) / 32;
      for (int l = 0; l < 32; l++) {
        const int64_t r = sl * 32 + l;
        const int32_t a0 = r < nf ? P.ff_rowptr[r] : 0, len = r < nf ? P.ff_rowptr[r + 1] - a0 : 0;
        for (int32_t k = 0; k < w; k++) {
          const int32_t q = base + 32 * k + l;
          if (k < len) {
            col[q] = P.ff_col[a0 + k];
            src[q] = P.ff_src[a0 + k];
            if (a0 + k == P.ff_diag[r]) diag[r] = q;
          } else {
            col[q] = int32_t(std::min<int64_t>(r, nf - 1));
          }
        }
      }
    }
    P.ff_nnz = int64_t(P.ff_col.size());
    P.ff_col.swap(col);
    P.ff_src.swap(src);
    P.ff_diag.swap(diag);
    // 3x3 block view for the SpMV-class kernels when every structure's free DOFs come in whole triples (frames whose
    // nodes are either free or fully fixed: the C4 / C5 lattices) and the blocks are well filled
    bool triples = nf % 3 == 0 && !getenv("OBGPU_NO_BLOCK_SELL");
    for (int64_t s = 0; s <= S && triples; s++) triples = sym.free_off[s] % 3 == 0;
    if (triples) build_block_sell(nf, P.sl_ptr, P.ff_col, &P.ff_src, P.ff_blk);
    P.ff_rowptr.clear();
    P.ff_rowptr.shrink_to_fit();
  }
  // ---- load vector gather lists ----
  const int64_t nt = nf + nb;
  std::vector<int32_t> cnt(nt + 1, 0);
  auto target = [&](int64_t s, int32_t t) -> int64_t {
    if (t == SLOT_NONE) return -1;
    return t >= 0 ? sym.free_off[s] + t : nf + sym.bnd_off[s] + (-t - 1);
  };
  auto for_each_load = [&](auto&& fn) {
    for (int64_t s = 0; s < S; s++) {
      const int64_t n0 = hb.node_off[s], e0 = hb.elem_off[s];
      for (int64_t l = hb.nload_off[s]; l < hb.nload_off[s + 1]; l++)
        if (sym.nload_slot[l] >= 0) fn(target(s, sym.nload_slot[l]), int32_t(-(l + 1)));  // free DOFs only
      for (int64_t j = hb.eload_off[s]; j < hb.eload_off[s + 1]; j++) {  // j = multimap position
        const int64_t e = e0 + hb.eload_elem[sym.el_order[j]];
        for (int f = 0; f < 2; f++)
          for (int a = 0; a < 3; a++) {
            const int64_t t = target(s, sym.node_slot[3 * (n0 + hb.elem_nodes[2 * e + f]) + a]);
            if (t >= 0) fn(t, int32_t(j * 6 + 3 * f + a));
          }
      }
    }
  };
  for_each_load([&](int64_t t, int32_t) { cnt[t + 1]++; });
  for (int64_t i = 0; i < nt; i++) cnt[i + 1] += cnt[i];
  P.fseg_ptr = cnt;
  P.fseg_src.resize(cnt[nt]);
  std::vector<int32_t> cur(cnt.begin(), cnt.end() - 1);
  for_each_load([&](int64_t t, int32_t src) { P.fseg_src[cur[t]++] = src; });
  P.valid = true;
  return OBGPU_OK;
}

}  // namespace obgpu
