// symbolic_mg.cpp — integer/coordinate plan of the multigrid preconditioner of the sparse route (host, once per
// topology).  The preconditioner replaces the fill-reducing ordering + factorisation of the reference's sparse branch
// (Eigen::SimplicialLLT, solver_static.cpp:110-129) for structures far beyond what that branch can hold: an
// aggregation multigrid V-cycle whose coarse spaces are the rigid-body modes of boxes of nodes.
//
//   level 0   rows = free DOFs of K_ff; "node" = structure node, position = node_xy
//   level l   rows = 3 per aggregate (tx, ty, rot about the centre); "node" = aggregate, position = mean of members
//   boxes     structure-aligned square cells of side 2^(l+1) * h0, h0 = median element length of the structure
//   P row     fine row of kind tx: (1, 0, -(y - yc)); ty: (0, 1, x - xc); rot: (0, 0, 1)   -> rigid-body motion
//
// A structure stops coarsening when a level has at most MgPlan::DENSE_MAX rows (solved with an explicit inverse on the
// device) or when the boxes no longer merge anything.
#include <algorithm>
#include <cmath>
#include <numeric>
#include <cstdlib>

#include "obgpu_internal.h"

namespace obgpu {

namespace {

struct LevelNodes {                 // the "nodes" of one structure at one level
  std::vector<double> x, y;         // position
  std::vector<int32_t> row[3];      // global row of kind k at this level, -1 = none
};

inline int32_t pad96(int64_t n) { return int32_t((n + 95) / 96 * 96); }

}  // namespace

bool build_block_sell(int64_t n, const std::vector<int32_t>& sl_ptr, const std::vector<int32_t>& col,
                      const std::vector<int32_t>* src, BlockSell& out) {
  out = BlockSell();
  if (n <= 0 || n % 3 != 0) return false;
  const int64_t nbr = n / 3, nsl = (nbr + 31) / 32;
  // block columns of every block row: union over its three scalar rows, ascending
  std::vector<int32_t> bptr(size_t(nbr) + 1, 0), bcols, tmp;
  int64_t scalar_entries = 0;
  for (int64_t I = 0; I < nbr; I++) {
    tmp.clear();
    for (int a = 0; a < 3; a++) {
      const int64_t r = 3 * I + a;
      const int32_t base = sl_ptr[size_t(r >> 5)], w = (sl_ptr[size_t(r >> 5) + 1] - base) >> 5;
      for (int32_t k = 0; k < w; k++) {
        const int32_t q = base + 32 * k + int32_t(r & 31);
        if (src && (*src)[size_t(q)] < 0) continue;   // padding entry of the scalar layout (level 0 marks them)
        tmp.push_back(col[size_t(q)] / 3);             // (coarse levels pad with the row's own column: same block)
        scalar_entries++;
      }
    }
    std::sort(tmp.begin(), tmp.end());
    tmp.erase(std::unique(tmp.begin(), tmp.end()), tmp.end());
    bcols.insert(bcols.end(), tmp.begin(), tmp.end());
    bptr[size_t(I) + 1] = int32_t(bcols.size());
  }
  if (double(scalar_entries) < 0.7 * 9.0 * double(bcols.size())) return false;  // poorly filled blocks: scalar SELL is smaller
  out.sl_ptr.assign(size_t(nsl) + 1, 0);
  for (int64_t sl = 0; sl < nsl; sl++) {
    int32_t w = 0;
    for (int64_t I = sl * 32; I < std::min<int64_t>(nbr, sl * 32 + 32); I++) w = std::max(w, bptr[size_t(I) + 1] - bptr[size_t(I)]);
    const int64_t next = int64_t(out.sl_ptr[size_t(sl)]) + int64_t(w) * 32;
    if (next * 9 >= INT32_MAX) return false;
    out.sl_ptr[size_t(sl) + 1] = int32_t(next);
  }
  const size_t nblk = size_t(out.sl_ptr[size_t(nsl)]);
  out.col.assign(nblk, 0);
  out.map.assign(nblk * 9, -1);
  for (int64_t sl = 0; sl < nsl; sl++) {
    const int32_t base = out.sl_ptr[size_t(sl)], w = (out.sl_ptr[size_t(sl) + 1] - base) >> 5;
    for (int l = 0; l < 32; l++) {
      const int64_t I = sl * 32 + l;
      for (int32_t k = 0; k < w; k++) out.col[size_t(base + 32 * k + l)] = int32_t(std::min<int64_t>(I, nbr - 1));  // padding: own block
      if (I >= nbr) continue;
      const int32_t nb = bptr[size_t(I) + 1] - bptr[size_t(I)];
      for (int32_t k = 0; k < nb; k++) out.col[size_t(base + 32 * k + l)] = bcols[size_t(bptr[size_t(I)] + k)];
      for (int a = 0; a < 3; a++) {
        const int64_t r = 3 * I + a;
        const int32_t sb = sl_ptr[size_t(r >> 5)], sw = (sl_ptr[size_t(r >> 5) + 1] - sb) >> 5;
        for (int32_t k = 0; k < sw; k++) {
          const int32_t q = sb + 32 * k + int32_t(r & 31);
          if (src && (*src)[size_t(q)] < 0) continue;
          const int32_t c = col[size_t(q)], J = c / 3, b = c - 3 * J;
          const int32_t* first = &bcols[size_t(bptr[size_t(I)])];
          const int32_t slot = int32_t(std::lower_bound(first, first + nb, J) - first);
          const size_t pos = size_t(9) * size_t(base) + size_t((9 * slot + 3 * a + b) * 32 + l);
          if (out.map[pos] < 0) out.map[pos] = q;   // a coarse row may list its own column twice through padding: first wins
        }
      }
    }
  }
  out.valid = true;
  return true;
}

int build_mg_plan(const HostBatch& hb, const Symbolic& sym, const SparsePlan& plan, MgPlan& mg, std::string& err) {
  mg = MgPlan();
  if (!plan.valid) {
    err = "multigrid plan needs the sparse plan";
    return OBGPU_ESEQUENCE;
  }
  const int64_t S = hb.S, nf = sym.nf_tot();
  mg.nlev.assign(S, 1);
  mg.dense.assign(S, 0);
  mg.lev.resize(1);
  // ---- level-0 nodes and cell size per structure ----
  std::vector<LevelNodes> cur(S);
  std::vector<double> h0(S, 1.0);
  std::vector<int32_t> cur_rows(S);  // rows in use at the current level
  for (int64_t s = 0; s < S; s++) {
    const int64_t n0 = hb.node_off[s], N = hb.node_off[s + 1] - n0, f0 = sym.free_off[s];
    LevelNodes& ln = cur[s];
    for (int64_t i = 0; i < N; i++) {
      int32_t r[3];
      bool any = false;
      for (int a = 0; a < 3; a++) {
        const int32_t sl = sym.node_slot[3 * (n0 + i) + a];
        r[a] = sl >= 0 ? int32_t(f0 + sl) : -1;
        any |= sl >= 0;
      }
      if (!any) continue;
      ln.x.push_back(hb.node_xy[2 * (n0 + i)]);
      ln.y.push_back(hb.node_xy[2 * (n0 + i) + 1]);
      for (int a = 0; a < 3; a++) ln.row[a].push_back(r[a]);
    }
    std::vector<double> len;
    for (int64_t e = hb.elem_off[s]; e < hb.elem_off[s + 1]; e++) {
      const int64_t a = n0 + hb.elem_nodes[2 * e], b = n0 + hb.elem_nodes[2 * e + 1];
      const double l = std::hypot(hb.node_xy[2 * b] - hb.node_xy[2 * a], hb.node_xy[2 * b + 1] - hb.node_xy[2 * a + 1]);
      if (l > 0) len.push_back(l);
    }
    if (!len.empty()) {
      std::nth_element(len.begin(), len.begin() + len.size() / 2, len.end());
      h0[s] = len[len.size() / 2];
    }
    cur_rows[s] = int32_t(sym.free_off[s + 1] - f0);
  }
  // fine-level pattern accessors (level l-1 while building level l)
  const std::vector<int32_t>* f_slptr = &plan.sl_ptr;
  const std::vector<int32_t>* f_col = &plan.ff_col;
  int64_t f_n = nf;
  std::vector<uint8_t> active(S, 1);
  for (int l = 1; l < 24; l++) {
    // which structures coarsen further?
    bool any = false;
    for (int64_t s = 0; s < S; s++) {
      if (!active[s]) continue;
      if (cur_rows[s] <= MgPlan::DENSE_MAX) {
        active[s] = 0;
        mg.dense[s] = 1;
      } else {
        any = true;
      }
    }
    if (!any) break;
    MgLevel L;
    L.row_off.assign(S + 1, 0);
    L.nrows.assign(S, 0);
    L.f_agg.assign(f_n, -1);
    L.f_kind.assign(f_n, 0);
    L.f_w.assign(f_n, 0.0);
    std::vector<LevelNodes> nxt(S);
    std::vector<std::vector<int32_t>> agg_of_node(S);
    // ---- aggregation: boxes of side 2^l h0 ----
    for (int64_t s = 0; s < S; s++) {
      L.row_off[s + 1] = L.row_off[s];
      if (!active[s]) continue;
      const LevelNodes& ln = cur[s];
      const size_t nn = ln.x.size();
      const double h = h0[s] * std::ldexp(1.0, l);
      double xmin = 1e300, ymin = 1e300, xmax = -1e300;
      for (size_t i = 0; i < nn; i++) {
        xmin = std::min(xmin, ln.x[i]); ymin = std::min(ymin, ln.y[i]); xmax = std::max(xmax, ln.x[i]);
      }
      const int64_t ncx = int64_t(std::floor((xmax - xmin) / h + 1e-9)) + 1;
      std::vector<std::pair<int64_t, int32_t>> key(nn);
      for (size_t i = 0; i < nn; i++) {
        const int64_t ix = int64_t(std::floor((ln.x[i] - xmin) / h + 1e-9)), iy = int64_t(std::floor((ln.y[i] - ymin) / h + 1e-9));
        key[i] = {iy * ncx + ix, int32_t(i)};
      }
      std::sort(key.begin(), key.end());
      std::vector<int32_t>& an = agg_of_node[s];
      an.assign(nn, 0);
      int32_t na = 0;
      for (size_t i = 0; i < nn; i++) {
        if (i > 0 && key[i].first != key[i - 1].first) na++;
        an[key[i].second] = na;
      }
      na = nn ? na + 1 : 0;
      if (double(na) > 0.8 * double(nn)) {  // the boxes merge (almost) nothing: stop here, smoother-only coarsest level
        active[s] = 0;
        an.clear();
        continue;
      }
      LevelNodes& nx = nxt[s];
      nx.x.assign(na, 0.0); nx.y.assign(na, 0.0);
      std::vector<int32_t> cnt(na, 0);
      for (size_t i = 0; i < nn; i++) { nx.x[an[i]] += ln.x[i]; nx.y[an[i]] += ln.y[i]; cnt[an[i]]++; }
      for (int32_t a = 0; a < na; a++) { nx.x[a] /= cnt[a]; nx.y[a] /= cnt[a]; }
      const int32_t r0 = L.row_off[s];
      for (int k = 0; k < 3; k++) {
        nx.row[k].resize(na);
        for (int32_t a = 0; a < na; a++) nx.row[k][a] = r0 + 3 * a + k;
      }
      L.nrows[s] = 3 * na;
      L.row_off[s + 1] = r0 + pad96(3 * int64_t(na));
      for (size_t i = 0; i < nn; i++) {
        const int32_t a = an[i];
        const double dx = ln.x[i] - nx.x[a], dy = ln.y[i] - nx.y[a];
        for (int k = 0; k < 3; k++) {
          const int32_t r = ln.row[k][i];
          if (r < 0) continue;
          L.f_agg[r] = r0 + 3 * a;
          L.f_kind[r] = uint8_t(k);
          L.f_w[r] = k == 0 ? -dy : k == 1 ? dx : 1.0;
        }
      }
      mg.nlev[s] = l + 1;
    }
    L.n = L.row_off[S];
    if (L.n == 0) break;  // nobody coarsened
    if (L.n >= INT32_MAX / 64) {
      err = "multigrid level too large for 32-bit offsets";
      return OBGPU_EUNSUPPORTED;
    }
    // ---- members of each aggregate (fine rows ascending) ----
    const int64_t nblk = L.n / 3;
    L.mem_ptr.assign(nblk + 1, 0);
    for (int64_t r = 0; r < f_n; r++)
      if (L.f_agg[r] >= 0) L.mem_ptr[L.f_agg[r] / 3 + 1]++;
    for (int64_t b = 0; b < nblk; b++) L.mem_ptr[b + 1] += L.mem_ptr[b];
    L.mem.resize(L.mem_ptr[nblk]);
    {
      std::vector<int32_t> pos(L.mem_ptr.begin(), L.mem_ptr.end() - 1);
      for (int64_t r = 0; r < f_n; r++)
        if (L.f_agg[r] >= 0) L.mem[pos[L.f_agg[r] / 3]++] = int32_t(r);
    }
    // ---- pattern: aggregate I couples to J when a fine entry (i in I, j in J) exists ----
    std::vector<int32_t> blk_ptr(nblk + 1, 0), blk_col;
    {
      std::vector<int32_t> stamp(nblk, -1), lst;
      for (int64_t I = 0; I < nblk; I++) {
        lst.clear();
        for (int32_t m = L.mem_ptr[I]; m < L.mem_ptr[I + 1]; m++) {
          const int32_t i = L.mem[m];
          const int32_t base = (*f_slptr)[i >> 5], w = ((*f_slptr)[(i >> 5) + 1] - base) >> 5;
          for (int32_t k = 0; k < w; k++) {
            const int32_t j = (*f_col)[base + 32 * k + (i & 31)];
            const int32_t ja = L.f_agg[j];
            if (ja < 0) continue;
            const int32_t J = ja / 3;
            if (stamp[J] != int32_t(I)) { stamp[J] = int32_t(I); lst.push_back(J); }
          }
        }
        std::sort(lst.begin(), lst.end());
        blk_col.insert(blk_col.end(), lst.begin(), lst.end());
        blk_ptr[I + 1] = int32_t(blk_col.size());
      }
    }
    const int64_t nsl = L.n / 32;
    L.sl_ptr.assign(nsl + 1, 0);
    for (int64_t sl = 0; sl < nsl; sl++) {
      int32_t w = 0;
      for (int64_t r = sl * 32; r < sl * 32 + 32; r++) w = std::max(w, 3 * (blk_ptr[r / 3 + 1] - blk_ptr[r / 3]));
      const int64_t next = int64_t(L.sl_ptr[sl]) + int64_t(w) * 32;
      if (next >= INT32_MAX) {
        err = "multigrid level too large for 32-bit SELL offsets";
        return OBGPU_EUNSUPPORTED;
      }
      L.sl_ptr[sl + 1] = int32_t(next);
    }
    L.col.resize(L.sl_ptr[nsl]);
    L.diag.assign(L.n, -1);
    for (int64_t r = 0; r < L.n; r++) {
      const int32_t base = L.sl_ptr[r >> 5], w = (L.sl_ptr[(r >> 5) + 1] - base) >> 5;
      const int64_t I = r / 3;
      const int32_t nb = blk_ptr[I + 1] - blk_ptr[I];
      for (int32_t k = 0; k < w; k++) {
        const int32_t q = base + 32 * k + int32_t(r & 31);
        if (k < 3 * nb) {
          const int32_t c = 3 * blk_col[blk_ptr[I] + k / 3] + k % 3;
          L.col[q] = c;
          if (c == r) L.diag[r] = q;
        } else {
          L.col[q] = int32_t(r);
        }
      }
    }
    if (!getenv("OBGPU_NO_BLOCK_SELL")) build_block_sell(L.n, L.sl_ptr, L.col, nullptr, L.blk);
    mg.lev.push_back(std::move(L));
    for (int64_t s = 0; s < S; s++)
      if (active[s]) {
        cur[s] = std::move(nxt[s]);
        cur_rows[s] = mg.lev.back().nrows[s];
      }
    f_slptr = &mg.lev.back().sl_ptr;
    f_col = &mg.lev.back().col;
    f_n = mg.lev.back().n;
  }
  mg.valid = true;
  return OBGPU_OK;
}

}  // namespace obgpu
