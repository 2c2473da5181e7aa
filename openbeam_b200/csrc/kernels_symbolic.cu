// kernels_symbolic.cu — K3 on the DEVICE for batches of many small structures (BASELINE config 3: millions of different
// frames): the integer "symbolic phase" the reference rebuilds inside every assembleProblem() call,
//   per-end global DOF masks        CElement::getGlobalDoFs, CElement.cpp:192-247 (+ getLocalDoFs of each element class)
//   node masks, DOF numbering       updateNodeConnections + updateListDoFs, CFiniteElementProblem.cpp:179-245
//   free / bounded partition, U_b   assemble.cpp:74-99, :304-313 (+ nodal-frame rotation of U_b, :349-435)
// as three kernels over the batch already in HBM — thread = structure (structures here have tens of nodes):
//   k_sym_count  validates the structure, ORs the element-end masks into its nodes, marks constrained DOFs, counts n/n_f/n_b
//   k_sym_scan_* exclusive scans of the three counts over the structures (chunk sums, bases, apply): DOF / free / bounded offsets
//   k_sym_fill   per-node slots (free index >= 0, bounded < 0), U_b, nodal-load slots, the element-load multimap order,
//                half bandwidth and the per-structure record the batched-LDLt kernel reads
// What it produces is bit-identical to the host phase (symbolic.cpp), which tests/test_gpu_devsym.py checks.
#include "kernels.h"

namespace obgpu {

// same arithmetic as symbolic.cpp's element_end_masks (internals.cpp:30-123 with z = 0), with explicit round-to-nearest
// products and sums so that no FMA contraction can move a value across the 1e-10 threshold differently from the host
__device__ __forceinline__ bool dev_end_masks(double2 p1, double2 p2, int type, const double* par, unsigned& m0, unsigned& m1,
                                              bool& zero_len) {
  const double dx = __dsub_rn(p2.x, p1.x), dy = __dsub_rn(p2.y, p1.y);
  const double len = sqrt(__dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), 0.0));
  zero_len = !(len > 0);
  double X0 = 1, X1 = 0, Y0 = 0, Y1 = 1;
  if (len > 0) {
    const double inv = 1 / len;
    X0 = __dmul_rn(dx, inv); X1 = __dmul_rn(dy, inv);
    Y0 = __dsub_rn(__dmul_rn(0.0, 0.0), __dmul_rn(1.0, X1));   // Z x X with Z = (0,0,1)
    Y1 = __dsub_rn(__dmul_rn(1.0, X0), __dmul_rn(0.0, 0.0));
  }
  // R = [X Y Z] columns; nz[g][l] = |R(g,l)| > 1e-10
  const bool nz00 = fabs(X0) > 1e-10, nz01 = fabs(Y0) > 1e-10, nz10 = fabs(X1) > 1e-10, nz11 = fabs(Y1) > 1e-10;
  auto glob = [&](unsigned local) {
    unsigned g = 0;
    {  // translations
      const unsigned l = local & 7u;
      unsigned o = 0;
      if (l == 7u) o = 7u;
      else {
        if (((l & 1u) && nz00) || ((l & 2u) && nz01)) o |= 1u;
        if (((l & 1u) && nz10) || ((l & 2u) && nz11)) o |= 2u;
        if (l & 4u) o |= 4u;   // R(2,2) = 1
      }
      g |= o;
    }
    {  // rotations
      const unsigned l = (local >> 3) & 7u;
      unsigned o = 0;
      if (l == 7u) o = 7u;
      else {
        if (((l & 1u) && nz00) || ((l & 2u) && nz01)) o |= 1u;
        if (((l & 1u) && nz10) || ((l & 2u) && nz11)) o |= 2u;
        if (l & 4u) o |= 4u;
      }
      g |= o << 3;
    }
    return g;
  };
  const unsigned R_ = 0x23, A2 = 0x03, A1 = 0x01;
  unsigned l0, l1;
  switch (type) {
    case OBGPU_BEAM2D_RR: l0 = R_; l1 = R_; break;
    case OBGPU_BEAM2D_RA: l0 = R_; l1 = A2; break;
    case OBGPU_BEAM2D_AR: l0 = A2; l1 = R_; break;
    case OBGPU_BEAM2D_AA: l0 = A1; l1 = A1; break;
    case OBGPU_BEAM2D_RD: l0 = R_; l1 = 0x21; break;
    case OBGPU_SPRING_1D: l0 = l1 = A1; break;
    case OBGPU_SPRING_XY: l0 = l1 = (par[0] != 0 ? 1u : 0u) | (par[1] != 0 ? 2u : 0u); break;
    case OBGPU_SPRING_DXDYRZ: l0 = l1 = (par[0] != 0 ? 1u : 0u) | (par[1] != 0 ? 2u : 0u) | (par[2] != 0 ? 0x20u : 0u); break;
    case OBGPU_SPRING_TORSION: l0 = l1 = 0x20; break;
    default: return false;
  }
  m0 = glob(l0);
  m1 = glob(l1);
  return true;
}

// 6-bit global mask {DX,DY,DZ,RX,RY,RZ} -> 3-bit planar mask {DX,DY,RZ}; a planar element never sets DZ, RX, RY
__device__ __forceinline__ unsigned planar3(unsigned m6) { return (m6 & 3u) | ((m6 >> 3) & 4u); }

__device__ __forceinline__ void sym_error(int32_t* err, int code, int64_t s) {
  if (atomicCAS(err, 0, code) == 0) err[1] = int32_t(s);
}

__global__ void __launch_bounds__(128) k_sym_count(SymDev d) {
  const int64_t s = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (s >= d.S) return;
  const int64_t n0 = d.node_off[s], N = d.node_off[s + 1] - n0, e0 = d.elem_off[s], E = d.elem_off[s + 1] - e0;
  int32_t* cnt = d.counts + 4 * s;
  cnt[0] = cnt[1] = cnt[2] = cnt[3] = 0;
  for (int64_t i = 0; i < N; i++) d.node_mask[n0 + i] = 0;
  for (int64_t e = 0; e < E; e++) {
    const uint2 nd = d.elem_nodes[e0 + e];
    if (nd.x >= N || nd.y >= N) return sym_error(d.error, SYM_E_ELEM_NODE, s);
    const int type = d.elem_type[e0 + e];
    const double par[3] = {d.elem_EAI[3 * (e0 + e)], d.elem_EAI[3 * (e0 + e) + 1], d.elem_EAI[3 * (e0 + e) + 2]};
    if (par[0] == 1.7976931348623157e308 || par[1] == 1.7976931348623157e308 || par[2] == 1.7976931348623157e308)
      return sym_error(d.error, SYM_E_UNINIT, s);
    unsigned m0, m1;
    bool zero_len;
    if (!dev_end_masks(d.node_xy[n0 + nd.x], d.node_xy[n0 + nd.y], type, par, m0, m1, zero_len))
      return sym_error(d.error, SYM_E_ELEM_TYPE, s);
    if (zero_len && type <= OBGPU_BEAM2D_RD) return sym_error(d.error, SYM_E_ZERO_LEN, s);
    if ((m0 | m1) & 0x1Cu) return sym_error(d.error, SYM_E_NONPLANAR, s);
    d.node_mask[n0 + nd.x] |= uint8_t(planar3(m0));
    d.node_mask[n0 + nd.y] |= uint8_t(planar3(m1));
  }
  // constraints (insertConstraint overwrites: the last value wins); constraints on absent DOFs are ignored
  for (int64_t c = d.con_off[s]; c < d.con_off[s + 1]; c++) {
    const uint32_t nd = d.con_node[c];
    const int dof = d.con_dof[c];
    if (nd >= N || dof > 5) return sym_error(d.error, SYM_E_CONSTRAINT, s);
    const int a = dof == 0 ? 0 : dof == 1 ? 1 : dof == 5 ? 2 : -1;
    if (a < 0 || !((d.node_mask[n0 + nd] >> a) & 1u)) continue;
    d.node_mask[n0 + nd] |= uint8_t(8u << a);
    d.node_cval[3 * (n0 + nd) + a] = d.con_val[c];
  }
  for (int64_t l = d.nload_off[s]; l < d.nload_off[s + 1]; l++)
    if (d.nload_node[l] >= N || d.nload_dof[l] > 5) return sym_error(d.error, SYM_E_NLOAD, s);
  for (int64_t l = d.eload_off[s]; l < d.eload_off[s + 1]; l++) {
    const uint32_t el = d.eload_elem[l];
    if (el >= E) return sym_error(d.error, SYM_E_ELOAD, s);
    const int lt = d.eload_type ? d.eload_type[l] : OBGPU_LOAD_UNIFORM;
    const int et = d.elem_type[e0 + el];
    if (lt > OBGPU_LOAD_TEMPERATURE) return sym_error(d.error, SYM_E_ELOAD, s);
    if (lt != OBGPU_LOAD_TEMPERATURE) {  // ASSERT_NEAR_(|dir|^2, 1, 1e-6), CLoadDistributedUniform.cpp:37-38
      const double vx = d.eload_dir[3 * l], vy = d.eload_dir[3 * l + 1], vz = d.eload_dir[3 * l + 2];
      const double n2 = vx * vx + vy * vy + vz * vz;
      if (!(n2 > 1 - 1e-6 && n2 < 1 + 1e-6)) return sym_error(d.error, SYM_E_LOAD_DIR, s);
    }
    if ((lt == OBGPU_LOAD_UNIFORM && et > OBGPU_BEAM2D_AA) || (lt == OBGPU_LOAD_TEMPERATURE && et > OBGPU_BEAM2D_RD) ||
        ((lt == OBGPU_LOAD_TRIANGULAR || lt == OBGPU_LOAD_CONCENTRATED) && et != OBGPU_BEAM2D_RR))
      return sym_error(d.error, SYM_E_LOAD_ELEM, s);
    if (lt == OBGPU_LOAD_CONCENTRATED) {  // CLoadConcentratedForce.cpp:77-84,114-116
      const uint2 nd = d.elem_nodes[e0 + el];
      const double2 p0 = d.node_xy[n0 + nd.x], p1 = d.node_xy[n0 + nd.y];
      const double ax = p1.x - p0.x, ay = p1.y - p0.y, len = sqrt(ax * ax + ay * ay);
      const double plong = (ax * d.eload_dir[3 * l] + ay * d.eload_dir[3 * l + 1]) / len;
      if (fabs(plong) > 1e-5 || d.eload_q2[l] < 0 || len - d.eload_q2[l] < 0) return sym_error(d.error, SYM_E_LOAD_ELEM, s);
    }
  }
  int n = 0, nb = 0;
  for (int64_t i = 0; i < N; i++) {
    const unsigned m = d.node_mask[n0 + i];
    n += __popc(m & 7u);
    nb += __popc((m >> 3) & 7u);
  }
  if (n == 0) return sym_error(d.error, SYM_E_NO_DOFS, s);
  cnt[0] = n; cnt[1] = n - nb; cnt[2] = nb;
}

// exclusive scans of counts[.][0..2] over the structures: per-CTA partial sums, a one-CTA scan of the partials, then every
// CTA rescans its own chunk from its base (thread = structure inside the chunk; fixed order, integers: exact)
constexpr int SCAN_CHUNK = 1024;
__global__ void __launch_bounds__(SCAN_CHUNK) k_sym_scan_partial(int64_t S, const int32_t* __restrict__ counts,
                                                                   int64_t* __restrict__ partial /* [nchunk][3] */) {
  __shared__ int64_t red[3][SCAN_CHUNK / 32];
  const int64_t i = int64_t(blockIdx.x) * SCAN_CHUNK + threadIdx.x;
  const int4 c = i < S ? reinterpret_cast<const int4*>(counts)[i] : make_int4(0, 0, 0, 0);
  int64_t v[3] = {c.x, c.y, c.z};
#pragma unroll
  for (int k = 0; k < 3; k++) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v[k] += __shfl_xor_sync(0xffffffffu, v[k], o);
    if ((threadIdx.x & 31) == 0) red[k][threadIdx.x >> 5] = v[k];
  }
  __syncthreads();
  if (threadIdx.x < 3) {
    int64_t t = 0;
    for (int w = 0; w < SCAN_CHUNK / 32; w++) t += red[threadIdx.x][w];
    partial[3 * int64_t(blockIdx.x) + threadIdx.x] = t;
  }
}
__global__ void k_sym_scan_bases(int64_t nchunk, int64_t* __restrict__ partial) {  // in place: sums -> exclusive bases
  if (threadIdx.x < 3) {
    int64_t run = 0;
    for (int64_t k = 0; k < nchunk; k++) {
      const int64_t v = partial[3 * k + threadIdx.x];
      partial[3 * k + threadIdx.x] = run;
      run += v;
    }
    partial[3 * nchunk + threadIdx.x] = run;  // totals
  }
}
__global__ void __launch_bounds__(SCAN_CHUNK) k_sym_scan_apply(int64_t S, const int32_t* __restrict__ counts,
                                                                 const int64_t* __restrict__ partial, int64_t* __restrict__ dof_off,
                                                                 int64_t* __restrict__ free_off, int64_t* __restrict__ bnd_off) {
  __shared__ int64_t wsum[3][SCAN_CHUNK / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t i = int64_t(blockIdx.x) * SCAN_CHUNK + threadIdx.x;
  const int4 c = i < S ? reinterpret_cast<const int4*>(counts)[i] : make_int4(0, 0, 0, 0);
  int64_t v[3] = {c.x, c.y, c.z}, inc[3];
#pragma unroll
  for (int k = 0; k < 3; k++) {
    int64_t x = v[k];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int64_t y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    inc[k] = x;  // inclusive within the warp
    if (lane == 31) wsum[k][warp] = x;
  }
  __syncthreads();
  if (threadIdx.x < 3) {
    int64_t run = 0;
    for (int w = 0; w < SCAN_CHUNK / 32; w++) { const int64_t t = wsum[threadIdx.x][w]; wsum[threadIdx.x][w] = run; run += t; }
  }
  __syncthreads();
  if (i < S) {
    dof_off[i] = partial[3 * int64_t(blockIdx.x) + 0] + wsum[0][warp] + inc[0] - v[0];
    free_off[i] = partial[3 * int64_t(blockIdx.x) + 1] + wsum[1][warp] + inc[1] - v[1];
    bnd_off[i] = partial[3 * int64_t(blockIdx.x) + 2] + wsum[2][warp] + inc[2] - v[2];
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    const int64_t nchunk = gridDim.x;
    dof_off[S] = partial[3 * nchunk]; free_off[S] = partial[3 * nchunk + 1]; bnd_off[S] = partial[3 * nchunk + 2];
  }
}

__global__ void __launch_bounds__(128) k_sym_fill(SymDev d) {
  const int64_t s = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (s >= d.S || d.error[0] != 0) return;   // a malformed structure anywhere in the batch refuses the batch
  const int64_t n0 = d.node_off[s], N = d.node_off[s + 1] - n0, e0 = d.elem_off[s], E = d.elem_off[s + 1] - e0;
  const int64_t f0 = d.free_off[s], b0 = d.bnd_off[s];
  int32_t* cnt = d.counts + 4 * s;
  // node-major, dof-minor numbering; the partition keeps ascending DOF order inside each part (assemble.cpp:74-99)
  int nf = 0, nb = 0;
  bool any_ub = false;
  for (int64_t i = 0; i < N; i++) {
    const unsigned m = d.node_mask[n0 + i];
    for (int a = 0; a < 3; a++) {
      int32_t slot = SLOT_NONE;
      if ((m >> a) & 1u) {
        if ((m >> (3 + a)) & 1u) {
          const double v = d.node_cval[3 * (n0 + i) + a];
          d.U_b[b0 + nb] = v;
          any_ub |= v != 0.0;
          d.bnd_dofid[b0 + nb] = uint8_t(a == 2 ? OBGPU_RZ : a);
          slot = -(nb++) - 1;
        } else {
          slot = nf++;
        }
      }
      d.node_slot[3 * (n0 + i) + a] = slot;
    }
  }
  if (any_ub) d.error[2] = 1;
  // assemble.cpp:349-435: prescribed displacements of a node with a rotated frame go to nodal coordinates, only the
  // bounded components taken (free ones count as 0)
  if (d.node_rotz)
    for (int64_t i = 0; i < N; i++) {
      const double r = d.node_rotz[n0 + i];
      if (r == 0.0) continue;
      const int32_t tx = d.node_slot[3 * (n0 + i)], ty = d.node_slot[3 * (n0 + i) + 1];
      const bool bx = tx < 0 && tx != SLOT_NONE, by = ty < 0 && ty != SLOT_NONE;
      if (!bx && !by) continue;
      const double ux = bx ? d.U_b[b0 + (-tx - 1)] : 0.0, uy = by ? d.U_b[b0 + (-ty - 1)] : 0.0;
      const double cr = cos(r), sr = sin(r);
      if (bx) d.U_b[b0 + (-tx - 1)] = __dadd_rn(__dmul_rn(cr, ux), __dmul_rn(sr, uy));
      if (by) d.U_b[b0 + (-ty - 1)] = __dadd_rn(__dmul_rn(-sr, ux), __dmul_rn(cr, uy));
    }
  for (int64_t l = d.nload_off[s]; l < d.nload_off[s + 1]; l++) {
    const int dof = d.nload_dof[l];
    const int a = dof == 0 ? 0 : dof == 1 ? 1 : dof == 5 ? 2 : -1;
    d.nload_slot[l] = a < 0 ? SLOT_NONE : d.node_slot[3 * (n0 + d.nload_node[l]) + a];
  }
  // element-load multimap order (m_loads_on_beams: by element, insertion order within an element): counting sort
  {
    const int64_t l0 = d.eload_off[s], EL = d.eload_off[s + 1] - l0;
    for (int64_t e = 0; e < E; e++) d.el_cursor[e0 + e] = 0;
    for (int64_t l = 0; l < EL; l++) d.el_cursor[e0 + d.eload_elem[l0 + l]]++;
    int32_t run = int32_t(l0);
    for (int64_t e = 0; e < E; e++) {
      const int32_t c = d.el_cursor[e0 + e];
      d.el_ptr[e0 + e] = run;
      d.el_cursor[e0 + e] = run;
      run += c;
    }
    for (int64_t l = 0; l < EL; l++) d.el_order[d.el_cursor[e0 + d.eload_elem[l0 + l]]++] = int32_t(l0 + l);
    if (s == d.S - 1) d.el_ptr[d.elem_off[d.S]] = int32_t(d.eload_off[d.S]);
  }
  int32_t bw = 0;
  for (int64_t e = 0; e < E; e++) {
    const uint2 nd = d.elem_nodes[e0 + e];
    int32_t lo = INT32_MAX, hi = -1;
#pragma unroll
    for (int a = 0; a < 3; a++) {
      const int32_t t0 = d.node_slot[3 * (n0 + nd.x) + a], t1 = d.node_slot[3 * (n0 + nd.y) + a];
      if (t0 >= 0) { lo = min(lo, t0); hi = max(hi, t0); }
      if (t1 >= 0) { lo = min(lo, t1); hi = max(hi, t1); }
    }
    if (hi >= 0) bw = max(bw, hi - lo);
  }
  cnt[3] = bw;
  FrameMeta m;
  m.n0 = n0; m.e0 = e0; m.nl0 = d.nload_off[s]; m.f0 = f0; m.b0 = b0;
  m.N = int32_t(N); m.E = int32_t(E); m.NL = int32_t(d.nload_off[s + 1] - d.nload_off[s]);
  m.n = cnt[0]; m.nf = nf; m.nb = nb; m.hb = min(bw, max(nf - 1, 0)); m.pad = 0;
  d.meta[s] = m;
  // size class of the warp-per-structure kernel (frame_class_of, kernels.h) + the class's extents
  {
    const int k = frame_class_of(m.hb, nf, nb);
    d.cls[s] = uint8_t(k);
    int32_t* st = d.cls_stats + 5 * k;   // count, max_band, max_nf, max_nb, max_hb
    atomicAdd(st, 1);
    atomicMax(st + 1, frame_ld(m.hb) * nf);
    atomicMax(st + 2, nf);
    atomicMax(st + 3, nb);
    atomicMax(st + 4, m.hb);
  }
}

// structure ids grouped by size class: class k occupies order[start[k] .. start[k+1]); the position inside a class comes
// from an atomic cursor, so it varies from run to run — it only decides which warp serves which structure
__global__ void __launch_bounds__(256) k_sym_order(int64_t S, const uint8_t* __restrict__ cls, int32_t* cursor /* [NCLS] starts */,
                                                   int32_t* __restrict__ order) {
  const int64_t s = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (s >= S) return;
  order[atomicAdd(cursor + cls[s], 1)] = int32_t(s);
}
cudaError_t launch_symbolic_order(int64_t S, const uint8_t* cls, int32_t* cursor, int32_t* order, cudaStream_t st) {
  k_sym_order<<<unsigned((S + 255) / 256), 256, 0, st>>>(S, cls, cursor, order);
  return cudaGetLastError();
}

cudaError_t launch_symbolic_device(const SymDev& d, cudaStream_t st) {
  if (d.S <= 0) return cudaSuccess;
  const unsigned g = unsigned((d.S + 127) / 128);
  k_sym_count<<<g, 128, 0, st>>>(d);
  const unsigned nchunk = unsigned((d.S + SCAN_CHUNK - 1) / SCAN_CHUNK);
  k_sym_scan_partial<<<nchunk, SCAN_CHUNK, 0, st>>>(d.S, d.counts, d.scan_partial);
  k_sym_scan_bases<<<1, 32, 0, st>>>(nchunk, d.scan_partial);
  k_sym_scan_apply<<<nchunk, SCAN_CHUNK, 0, st>>>(d.S, d.counts, d.scan_partial, d.dof_off, d.free_off, d.bnd_off);
  k_sym_fill<<<g, 128, 0, st>>>(d);
  return cudaGetLastError();
}

// ---- obgpu_scan_offsets_static: the offset scans with their scratch in a file-scope device array (no allocation) ----
// Added for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
// serving loops); it has no counterpart in the reference.  The same three scan kernels as launch_symbolic_device, but the
// per-chunk partial sums live in g_scan_static instead of the context's cudaMalloc'ed d_sym_partial.
//   * the array is held by every GPU that loads the library until the process exits: 3 * (4096 + 1) * 8 B = 96 KB
//   * nothing in it has to start at zero: k_sym_scan_partial writes every entry [0, 3 * nchunk) before k_sym_scan_bases
//     reads them and writes the totals, so the contents a previous call left behind are never read
//   * every call shares the array: two calls that overlap on the GPU (different streams, or a graph replayed while
//     another call runs) race on it.  Serialise them on one stream.
constexpr int64_t SCAN_STATIC_MAX_CHUNKS = 4096;                                 // 4 194 304 structures (BASELINE config C3)
constexpr int64_t SCAN_STATIC_MAX_STRUCTURES = SCAN_STATIC_MAX_CHUNKS * SCAN_CHUNK;
static __device__ __align__(256) int64_t g_scan_static[3 * (SCAN_STATIC_MAX_CHUNKS + 1)];

// In host code `g_scan_static` is only the symbol's host-side stand-in: the device address comes from cudaGetSymbolAddress,
// once per device, with that device current.
constexpr int SCAN_STATIC_MAX_DEVICES = 64;
static int64_t* g_scan_static_addr[SCAN_STATIC_MAX_DEVICES];   // device address of g_scan_static on device d (0 = not asked yet)

static void scan_static_msg(char* err, size_t cap, const char* msg) {
  if (!err || cap == 0) return;
  size_t i = 0;
  for (; msg[i] && i + 1 < cap; i++) err[i] = msg[i];
  err[i] = 0;
}

}  // namespace obgpu

extern "C" int obgpu_scan_offsets_static(int device, int64_t n_structures, const int32_t* d_counts, int64_t* d_dof_off,
                                         int64_t* d_free_off, int64_t* d_bnd_off, void* stream, char* err, size_t err_cap) {
  using namespace obgpu;
  scan_static_msg(err, err_cap, "");
  if (n_structures < 1 || !d_counts || !d_dof_off || !d_free_off || !d_bnd_off) {
    scan_static_msg(err, err_cap, "obgpu_scan_offsets_static: null array or n_structures < 1");
    return OBGPU_EINVAL;
  }
  if (n_structures > SCAN_STATIC_MAX_STRUCTURES) {
    scan_static_msg(err, err_cap,
                    "obgpu_scan_offsets_static: n_structures exceeds the static scratch (maximum 4194304 structures); use "
                    "obgpu_batch_upload_device + obgpu_symbolic, which allocate their scratch");
    return OBGPU_EINVAL;
  }
  if (device < 0 || device >= SCAN_STATIC_MAX_DEVICES) {
    scan_static_msg(err, err_cap, "obgpu_scan_offsets_static: device index out of range");
    return OBGPU_EINVAL;
  }
  int prev = -1;
  cudaError_t e = cudaGetDevice(&prev);
  if (e == cudaSuccess && prev != device) e = cudaSetDevice(device);
  if (e == cudaSuccess && !g_scan_static_addr[device]) {  // set-up, once per device (racing first calls store the same value)
    int64_t* p = nullptr;
    e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_scan_static);
    if (e == cudaSuccess) g_scan_static_addr[device] = p;
  }
  if (e == cudaSuccess) {
    int64_t* scratch = g_scan_static_addr[device];
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const unsigned nchunk = unsigned((n_structures + SCAN_CHUNK - 1) / SCAN_CHUNK);
    k_sym_scan_partial<<<nchunk, SCAN_CHUNK, 0, st>>>(n_structures, d_counts, scratch);
    k_sym_scan_bases<<<1, 32, 0, st>>>(nchunk, scratch);
    k_sym_scan_apply<<<nchunk, SCAN_CHUNK, 0, st>>>(n_structures, d_counts, scratch, d_dof_off, d_free_off, d_bnd_off);
    e = cudaGetLastError();
  }
  if (prev >= 0 && prev != device) cudaSetDevice(prev);
  if (e != cudaSuccess) {
    scan_static_msg(err, err_cap, cudaGetErrorString(e));
    return OBGPU_ECUDA;
  }
  return OBGPU_OK;
}
