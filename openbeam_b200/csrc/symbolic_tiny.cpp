// symbolic_tiny.cpp — compiles the symbolic result of ONE shared topology into the flat "program"
// consumed by the thread-per-system kernel (kernels_tiny.cu).  Integer work only.
#include <algorithm>
#include <numeric>

#include "obgpu_internal.h"
#include "tiny_program.h"

namespace obgpu {

int build_tiny_program(const HostBatch& hb, const Symbolic& sym, TinyProgram& prog, std::string& err) {
  prog = TinyProgram();
  if (hb.S != 1) {
    err = "tiny program needs a single shared topology";
    return OBGPU_EUNSUPPORTED;
  }
  const int N = int(hb.Ntot()), E = int(hb.Etot()), n = int(sym.n_tot()), nf = int(sym.nf_tot()), nb = int(sym.nb_tot());
  const int NL = int(hb.NLtot()), EL = int(hb.ELtot());
  if (nf > TINY_MAX_NF || n > TINY_MAX_N || E > TINY_MAX_E || N > TINY_MAX_NODES || NL > TINY_MAX_NL ||
      EL > TINY_MAX_EL || hb.has_rotz) {
    err = "structure too large for the thread-per-system kernel";
    return OBGPU_EUNSUPPORTED;
  }
  for (double v : sym.U_b)
    if (v != 0.0) {
      err = "non-zero prescribed displacements are not handled by the thread-per-system kernel";
      return OBGPU_EUNSUPPORTED;
    }
  TinyLayout lay(N, E, n, nf, nb, NL, EL);
  prog.N = N; prog.E = E; prog.n = n; prog.nf = nf; prog.nb = nb; prog.NL = NL; prog.EL = EL;
  prog.ints.assign(lay.n_ints, -1);
  prog.dbls.assign(lay.n_dbls, 0.0);
  int32_t* I = prog.ints.data();
  double* D = prog.dbls.data();
  I[0] = N; I[1] = E; I[2] = n; I[3] = nf; I[4] = nb; I[5] = NL; I[6] = EL; I[7] = nf * (nf + 1) / 2;
  auto packed = [](int i, int j) { return i >= j ? i * (i + 1) / 2 + j : j * (j + 1) / 2 + i; };
  static const int d3[3] = {0, 1, 5};
  // element loads in the reference's iteration order: multimap<element, load> == stable sort by element
  std::vector<int> lorder(EL);
  std::iota(lorder.begin(), lorder.end(), 0);
  std::stable_sort(lorder.begin(), lorder.end(), [&](int a, int b) { return hb.eload_elem[a] < hb.eload_elem[b]; });
  for (int e = 0; e < E; e++) {
    const uint32_t nd[2] = {hb.elem_nodes[2 * e], hb.elem_nodes[2 * e + 1]};
    I[lay.o_enode + 2 * e] = int(nd[0]);
    I[lay.o_enode + 2 * e + 1] = int(nd[1]);
    I[lay.o_etype + e] = hb.elem_type[e];
    int fidx[2][3];  // free index of (end, compact dof) or -1
    for (int f = 0; f < 2; f++)
      for (int a = 0; a < 3; a++) {
        const int k = sym.dof_index[6 * nd[f] + d3[a]];
        I[lay.o_udof + 6 * e + 3 * f + a] = (k >= 0 && sym.dof_type[k] < 0) ? (k | 64) : k;  // bit 6: bounded
        fidx[f][a] = (k >= 0 && sym.dof_type[k] >= 0) ? sym.dof_type[k] : -1;
      }
    int* dst = &I[lay.o_edst + ELEM_K * e];
    static const int up[6][2] = {{0, 0}, {0, 1}, {0, 2}, {1, 1}, {1, 2}, {2, 2}};
    for (int blk = 0; blk < 2; blk++)
      for (int u = 0; u < 6; u++) {
        const int fi = fidx[blk][up[u][0]], fj = fidx[blk][up[u][1]];
        dst[6 * blk + u] = (fi >= 0 && fj >= 0) ? packed(fi, fj) : -1;
      }
    for (int a = 0; a < 3; a++)
      for (int b = 0; b < 3; b++) {
        const int fi = fidx[0][a], fj = fidx[1][b];
        dst[12 + 3 * a + b] = (fi >= 0 && fj >= 0) ? packed(fi, fj) : -1;
      }
    D[lay.o_eai + 3 * e] = hb.elem_EAI[3 * e];
    D[lay.o_eai + 3 * e + 1] = hb.elem_EAI[3 * e + 1];
    D[lay.o_eai + 3 * e + 2] = hb.elem_EAI[3 * e + 2];
  }
  {  // CSR of element loads per element, in multimap order
    int p = 0;
    for (int e = 0; e < E; e++) {
      I[lay.o_elptr + e] = p;
      while (p < EL && int(hb.eload_elem[lorder[p]]) == e) {
        I[lay.o_elslot + p] = lorder[p];
        p++;
      }
    }
    I[lay.o_elptr + E] = p;
  }
  for (int i = 0; i < NL; i++) {
    const int k = sym.dof_index[6 * hb.nload_node[i] + hb.nload_dof[i]];
    I[lay.o_nldst + i] = (k >= 0 && sym.dof_type[k] >= 0) ? k : -1;  // loads on bounded/absent DOFs never reach F_f
    D[lay.o_nlval + i] = hb.nload_val[i];
  }
  for (int i = 0; i < nf; i++) I[lay.o_freek + i] = sym.free_idx[i];
  for (int i = 0; i < nb; i++) I[lay.o_bndk + i] = sym.bnd_idx[i];
  for (int i = 0; i < 2 * N; i++) D[lay.o_xy + i] = hb.node_xy[i];
  for (int i = 0; i < EL; i++) {
    D[lay.o_elq + i] = hb.eload_q[i];
    for (int a = 0; a < 3; a++) D[lay.o_eldir + 3 * i + a] = hb.eload_dir[3 * i + a];
    I[lay.o_elkind + i] = hb.has_load_kinds ? int(hb.eload_type[i]) : int(OBGPU_LOAD_UNIFORM);
    D[lay.o_elq2 + i] = hb.has_load_kinds ? hb.eload_q2[i] : 0.0;
  }
  prog.has_load_kinds = hb.has_load_kinds;
  prog.valid = true;
  return OBGPU_OK;
}

}  // namespace obgpu
