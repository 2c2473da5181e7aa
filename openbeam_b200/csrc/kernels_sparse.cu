// kernels_sparse.cu — the sparse route for large structures (BASELINE configs C4/C5):
//   K1  k_elem_stiff     lane = element: orientation + local stiffness + rotation -> compact global blocks ELEM_K
//   K2  k_elem_loads     lane = element load: equivalent nodal loads in global axes
//   K4  k_assemble_seg   thread = K slot: deterministic segmented sum of its element-block entries in the
//                        reference's triplet order (assemble.cpp:123-233) — a gather, no atomics
//   K5  k_gather         K_ff / K_bf values as gathers out of K through precomputed maps (assemble.cpp:249-297)
//       k_load_vector    F_f and the equivalent loads on bounded DOFs as per-DOF gathers (assemble.cpp:317-341)
//   K7  k_spmv_dot       y = A p (CSR, 8 lanes per row) fused with the p.Ap partial dot
//   K8  k_pcg_update / k_pcg_dir   fused axpy / preconditioner / dot kernels of Jacobi-PCG
//   K10 k_stress         lane = element: end forces from U (postproc_stress.cpp:32-183)
// Every kernel is HBM-bound; reductions finish in-kernel in a fixed order (last-block pattern) so results are
// bit-reproducible run to run.
#include "beam_math.cuh"
#include "kernels.h"

namespace obgpu {

__global__ void __launch_bounds__(256) k_elem_stiff(int64_t Etot, const uint2* __restrict__ elem_nodes,
                                                    const uint8_t* __restrict__ elem_type, const double* __restrict__ eai,
                                                    const double2* __restrict__ node_xy, const int32_t* __restrict__ elem_node0,
                                                    const double* __restrict__ node_rot, double* __restrict__ elem_k) {
  // a lane's 21 values are a 168-byte record: staged per warp in shared memory (odd stride: conflict-free) and
  // written as 32 consecutive doubles per instruction instead of 32 scattered 8-byte stores
  __shared__ double stg[8][32 * ELEM_K];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t e = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (e < Etot) {
    const uint2 nd = elem_nodes[e];
    const int64_t n0 = elem_node0[e];
    const double2 p0 = node_xy[n0 + nd.x], p1 = node_xy[n0 + nd.y];
    const ElemGeom g = elem_geom(p0.x, p0.y, p1.x, p1.y);
    const ElemLocalK lk = elem_local_k(elem_type[e], eai[3 * e], eai[3 * e + 1], eai[3 * e + 2], g);
    double ke[ELEM_K];
    elem_global_k(lk, g, ke);
    if (node_rot) {  // assemble.cpp:144-179: blocks of nodes with a rotated frame go to nodal coordinates
      const double r0 = node_rot[n0 + nd.x], r1 = node_rot[n0 + nd.y];
      if (r0 != 0.0 || r1 != 0.0) rotate_nodal(ke, cos(r0), sin(r0), cos(r1), sin(r1));
    }
#pragma unroll
    for (int k = 0; k < ELEM_K; k++) stg[warp][lane * ELEM_K + k] = ke[k];
  }
  __syncwarp();
  const int64_t e0 = e - lane;  // first element of this warp
  const int cnt = int(min(int64_t(32), Etot - e0));
  double* o = elem_k + e0 * ELEM_K;
  for (int i = lane; i < cnt * ELEM_K; i += 32) o[i] = stg[warp][i];
}

__global__ void __launch_bounds__(256) k_elem_loads(int64_t ELtot, const int32_t* __restrict__ el_order,
                                                    const int32_t* __restrict__ eload_elem_g, const double* __restrict__ q,
                                                    const double* __restrict__ dir, const uint2* __restrict__ elem_nodes,
                                                    const uint8_t* __restrict__ elem_type, const double2* __restrict__ node_xy,
                                                    const int32_t* __restrict__ elem_node0, const uint8_t* __restrict__ eload_type,
                                                    const double* __restrict__ eload_q2, const double* __restrict__ eai,
                                                    double* __restrict__ eload_g) {
  const int64_t j = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (j >= ELtot) return;
  const int sl = el_order[j];
  const int64_t e = eload_elem_g[sl];
  const uint2 nd = elem_nodes[e];
  const int64_t n0 = elem_node0[e];
  const double2 p0 = node_xy[n0 + nd.x], p1 = node_xy[n0 + nd.y];
  const ElemGeom g = elem_geom(p0.x, p0.y, p1.x, p1.y);
  double gl[6], ex[6];
  if (eload_type)
    elem_load(eload_type[sl], elem_type[e], g, q[sl], eload_q2[sl], dir[3 * sl], dir[3 * sl + 1], eai[3 * e] * eai[3 * e + 1], gl, ex);
  else
    uniform_load(elem_type[e], g, q[sl], dir[3 * sl], dir[3 * sl + 1], gl, ex);
#pragma unroll
  for (int k = 0; k < 6; k++) eload_g[j * 6 + k] = gl[k];
}

__global__ void __launch_bounds__(256) k_assemble_seg(int64_t nnz, const int32_t* __restrict__ seg_ptr,
                                                      const int32_t* __restrict__ seg_src, const double* __restrict__ elem_k,
                                                      double* __restrict__ kval) {
  const int64_t g = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (g >= nnz) return;
  double acc = 0.0;
  for (int32_t t = seg_ptr[g]; t < seg_ptr[g + 1]; t++) {
    const int32_t src = seg_src[t];
    acc += src >= 0 ? elem_k[src] : 0.0;  // structural zeros of the triplet list add +0 like the reference
  }
  kval[g] = acc;
}

// K1 fused into K4.  One CTA owns ASM_SLOTS consecutive K slots; the host plan lists the elements those slots draw
// from (a slot block of a node-ordered mesh is a compact patch: ~60 elements per 1024 slots on a grid frame).  The CTA
// computes their global blocks into shared memory (lane = element, records at an odd stride) and every thread then
// sums its slots from there in the reference's triplet order.  HBM sees the element inputs, 4 B per slot of segment
// pointers, 2 B per triplet of block-local sources and the K values once: no 168-byte element records are written
// and gathered back 8 bytes at a time (400 MB of reads at C4 in the two-kernel version).
constexpr int ASM_SLOTS = SparsePlan::ASM_SLOTS;
constexpr int ASM_SPT = 4;                        // consecutive slots per thread
constexpr int ASM_THREADS = ASM_SLOTS / ASM_SPT;  // 128 threads per CTA: short element phase, 16 CTAs/SM to overlap it
__global__ void __launch_bounds__(ASM_THREADS) k_assemble_fused(int64_t nnz, const int32_t* __restrict__ seg_ptr,
                                                        const uint16_t* __restrict__ seg_loc,
                                                        const int32_t* __restrict__ blk_elem_ptr, const int32_t* __restrict__ blk_elems,
                                                        const uint2* __restrict__ elem_nodes, const uint8_t* __restrict__ elem_type,
                                                        const double* __restrict__ eai, const double2* __restrict__ node_xy,
                                                        const int32_t* __restrict__ elem_node0, const double* __restrict__ node_rot,
                                                        double* __restrict__ kval) {
  extern __shared__ double kes[];  // [nloc][ELEM_K]
  constexpr int SPT = ASM_SPT;
  const int64_t g = int64_t(blockIdx.x) * ASM_SLOTS + SPT * threadIdx.x;
  int32_t p[SPT + 1];  // segment pointers first: their latency overlaps the element phase
#pragma unroll
  for (int r = 0; r <= SPT; r++) p[r] = seg_ptr[min(g + r, nnz)];
  const int32_t b0 = blk_elem_ptr[blockIdx.x], nloc = blk_elem_ptr[blockIdx.x + 1] - b0;
  for (int i = threadIdx.x; i < nloc; i += blockDim.x) {
    const int64_t e = blk_elems[b0 + i];
    const uint2 nd = elem_nodes[e];
    const int64_t n0 = elem_node0[e];
    const double2 p0 = node_xy[n0 + nd.x], p1 = node_xy[n0 + nd.y];
    const ElemGeom g = elem_geom(p0.x, p0.y, p1.x, p1.y);
    const ElemLocalK lk = elem_local_k(elem_type[e], eai[3 * e], eai[3 * e + 1], eai[3 * e + 2], g);
    double ke[ELEM_K];
    elem_global_k(lk, g, ke);
    if (node_rot) {
      const double r0 = node_rot[n0 + nd.x], r1 = node_rot[n0 + nd.y];
      if (r0 != 0.0 || r1 != 0.0) rotate_nodal(ke, cos(r0), sin(r0), cos(r1), sin(r1));
    }
#pragma unroll
    for (int k = 0; k < ELEM_K; k++) kes[i * ELEM_K + k] = ke[k];
  }
  __syncthreads();
  // each thread owns 4 CONSECUTIVE slots: one read of 5 consecutive segment pointers, one contiguous run of
  // block-local sources, 32 contiguous bytes of K values (two dependent global round trips in all)
  if (g >= nnz) return;
  double acc[SPT];
  // a slot sums 1-4 element entries on a frame (more only where many members meet): the first four sources of every
  // slot are loaded together (16 independent loads per thread in flight), the rest in a loop.  Order = triplet order;
  // structural zeros of the triplet list (0xFFFF) add +0 like the reference.
  unsigned src[SPT][4];
#pragma unroll
  for (int r = 0; r < SPT; r++)
#pragma unroll
    for (int i = 0; i < 4; i++) src[r][i] = p[r] + i < p[r + 1] ? unsigned(seg_loc[p[r] + i]) : 0xFFFFu;
#pragma unroll
  for (int r = 0; r < SPT; r++) {
    acc[r] = 0.0;
#pragma unroll
    for (int i = 0; i < 4; i++)
      if (p[r] + i < p[r + 1]) acc[r] += src[r][i] != 0xFFFFu ? kes[src[r][i]] : 0.0;
    for (int32_t t = p[r] + 4; t < p[r + 1]; t++) {
      const unsigned s2 = seg_loc[t];
      acc[r] += s2 != 0xFFFFu ? kes[s2] : 0.0;
    }
  }
  if (g + SPT <= nnz && (reinterpret_cast<uintptr_t>(kval + g) & 15) == 0) {
    double2* o = reinterpret_cast<double2*>(kval + g);
#pragma unroll
    for (int r = 0; r < SPT; r += 2) o[r / 2] = make_double2(acc[r], acc[r + 1]);
  } else {
#pragma unroll
    for (int r = 0; r < SPT; r++)
      if (g + r < nnz) kval[g + r] = acc[r];
  }
}

// values of a CSR matrix whose entries are K slots (K_bb): thread per row
__global__ void __launch_bounds__(256) k_gather_rows(int64_t nrows, const int32_t* __restrict__ rowptr, const int32_t* __restrict__ src,
                                                     const double* __restrict__ in, double* __restrict__ out) {
  const int64_t r = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (r >= nrows) return;
  for (int32_t k = rowptr[r]; k < rowptr[r + 1]; k++) out[k] = in[src[k]];
}

__global__ void __launch_bounds__(256) k_gather(int64_t n, const int32_t* __restrict__ src, const double* __restrict__ in,
                                                double* __restrict__ out) {
  const int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < n) {
    const int32_t s = src[i];
    out[i] = s >= 0 ? in[s] : 0.0;  // -1: padding entry of the SELL layout
  }
}

// FP32 copy of a 3x3 block-SELL matrix in the paired layout block_row_product_f32 reads (pcg_common.cuh): float j of the
// output belongs to block column B = j / 320, pair (j % 320) / 64, lane ((j % 320) / 2) % 32, component j & 1; the
// FP64 layout holds value v of that block at 288 * B + 32 * v + lane of `src`
__global__ void __launch_bounds__(256) k_gather_blk_f32(int64_t n_out, const int32_t* __restrict__ src,
                                                        const double* __restrict__ in, float* __restrict__ out) {
  const int64_t j = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (j >= n_out) return;
  const int64_t t = j >> 1, B = t / 160;
  const int rem = int(t - B * 160), pr = rem >> 5, lane = rem & 31, v = 2 * pr + int(j & 1);
  float r = 0.0f;
  if (v < 9) {
    const int32_t s = src[B * 288 + v * 32 + lane];
    if (s >= 0) r = float(in[s]);
  }
  out[j] = r;
}

__global__ void __launch_bounds__(256) k_load_vector(int64_t nf, int64_t nb, const int32_t* __restrict__ fseg_ptr,
                                                     const int32_t* __restrict__ fseg_src, const double* __restrict__ nload_val,
                                                     const double* __restrict__ eload_g, double* __restrict__ F_f,
                                                     double* __restrict__ Eq) {
  const int64_t t = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (t >= nf + nb) return;
  double acc = 0.0;
  for (int32_t i = fseg_ptr[t]; i < fseg_ptr[t + 1]; i++) {
    const int32_t s = fseg_src[i];
    acc += s >= 0 ? eload_g[s] : nload_val[-s - 1];
  }
  if (t < nf) F_f[t] = acc;
  else Eq[t - nf] = acc;
}

// assemble.cpp:349-435: (Fx, Fy) of a rotated node's FREE DOFs go to the nodal frame (bounded components count as 0)
__global__ void __launch_bounds__(256) k_rotate_loads(int64_t Ntot, const double* __restrict__ node_rot,
                                                      const int32_t* __restrict__ node_slot, const int32_t* __restrict__ node_f0,
                                                      double* __restrict__ F_f) {
  const int64_t n = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (n >= Ntot) return;
  const double r = node_rot[n];
  if (r == 0.0) return;
  const int32_t sx = node_slot[3 * n], sy = node_slot[3 * n + 1], f0 = node_f0[n];
  const double fx = sx >= 0 ? F_f[f0 + sx] : 0.0, fy = sy >= 0 ? F_f[f0 + sy] : 0.0;
  const double cr = cos(r), sr = sin(r);
  if (sx >= 0) F_f[f0 + sx] = cr * fx + sr * fy;
  if (sy >= 0) F_f[f0 + sy] = -sr * fx + cr * fy;
}

// rhs = F_f - K_bf^T U_b: one CTA; the bounded rows are applied one after the other (entries of a row hit distinct
// free columns), so the sum into every free DOF has a fixed order
__global__ void __launch_bounds__(256) k_rhs_prescribed(int64_t nf, int64_t nb, const int32_t* __restrict__ bf_rowptr,
                                                        const int32_t* __restrict__ bf_col, const double* __restrict__ bf_val,
                                                        const double* __restrict__ U_b, const double* __restrict__ F_f,
                                                        double* __restrict__ rhs) {
  for (int64_t i = threadIdx.x; i < nf; i += blockDim.x) rhs[i] = F_f[i];
  __syncthreads();
  for (int64_t r = 0; r < nb; r++) {
    const double ub = U_b[r];
    if (ub != 0.0)
      for (int32_t k = bf_rowptr[r] + threadIdx.x; k < bf_rowptr[r + 1]; k += blockDim.x) rhs[bf_col[k]] -= bf_val[k] * ub;
    __syncthreads();
  }
}

__global__ void __launch_bounds__(256) k_stress(int64_t Etot, const uint2* __restrict__ elem_nodes,
                                                const uint8_t* __restrict__ elem_type, const double* __restrict__ eai,
                                                const double2* __restrict__ node_xy, const int32_t* __restrict__ elem_node0,
                                                const int32_t* __restrict__ elem_f0, const int32_t* __restrict__ node_slot,
                                                const int32_t* __restrict__ el_ptr, const int32_t* __restrict__ el_order,
                                                const double* __restrict__ q, const double* __restrict__ dir,
                                                const uint8_t* __restrict__ eload_type, const double* __restrict__ eload_q2,
                                                const double* __restrict__ U_f, const double* __restrict__ U_b,
                                                const int32_t* __restrict__ elem_b0, double* __restrict__ stress) {
  __shared__ double stg[8][32 * 13];  // [lane][12] records, stride 13: conflict-free staging for coalesced stores
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t e = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (e < Etot) {
  const uint2 nd = elem_nodes[e];
  const int64_t n0 = elem_node0[e];
  const int type = elem_type[e];
  const double2 p0 = node_xy[n0 + nd.x], p1 = node_xy[n0 + nd.y];
  const ElemGeom g = elem_geom(p0.x, p0.y, p1.x, p1.y);
  const ElemLocalK lk = elem_local_k(type, eai[3 * e], eai[3 * e + 1], eai[3 * e + 2], g);
  const double* U = U_f + elem_f0[e];
  double ug[6], es[6];
#pragma unroll
  for (int k = 0; k < 3; k++) {
    const int32_t t0 = node_slot[3 * (n0 + nd.x) + k], t1 = node_slot[3 * (n0 + nd.y) + k];
    // postproc_stress.cpp:50-78: U_b on bounded DOFs, U_f on free ones (nodal frames are not undone there: quirk Q6)
    ug[k] = t0 >= 0 ? U[t0] : (U_b && t0 != SLOT_NONE) ? U_b[elem_b0[e] + (-t0 - 1)] : 0.0;
    ug[3 + k] = t1 >= 0 ? U[t1] : (U_b && t1 != SLOT_NONE) ? U_b[elem_b0[e] + (-t1 - 1)] : 0.0;
  }
  elem_end_forces(lk, g, ug, es);
  double ext[6] = {0, 0, 0, 0, 0, 0};
  for (int j = el_ptr[e]; j < el_ptr[e + 1]; j++) {
    const int sl = el_order[j];
    double gl[6], ex[6];
    if (eload_type)
      elem_load(eload_type[sl], type, g, q[sl], eload_q2[sl], dir[3 * sl], dir[3 * sl + 1], eai[3 * e] * eai[3 * e + 1], gl, ex);
    else
      uniform_load(type, g, q[sl], dir[3 * sl], dir[3 * sl + 1], gl, ex);
#pragma unroll
    for (int k = 0; k < 6; k++) ext[k] += ex[k];
  }
  double* o = stg[warp] + lane * 13;
  o[0] = -es[0] + ext[0]; o[1] = -es[1] + ext[1]; o[2] = 0.0; o[3] = 0.0; o[4] = 0.0; o[5] = -es[2] + ext[2];
  o[6] = es[3] + ext[3];  o[7] = es[4] + ext[4];  o[8] = 0.0; o[9] = 0.0; o[10] = 0.0; o[11] = es[5] + ext[5];
  }
  __syncwarp();
  const int64_t e0 = e - lane;
  const int cnt = int(min(int64_t(32), Etot - e0));
  double* g = stress + e0 * 12;
  for (int i = lane; i < cnt * 12; i += 32) g[i] = stg[warp][(i / 12) * 13 + i % 12];
}

static inline unsigned nblk(int64_t n, int b = 256) { return unsigned((n + b - 1) / b); }

cudaError_t launch_elem_stiff(const SparseDev& d, const double* eai, cudaStream_t st) {
  k_elem_stiff<<<nblk(d.Etot), 256, 0, st>>>(d.Etot, d.elem_nodes, d.elem_type, eai, d.node_xy, d.elem_node0, d.node_rot, d.elem_k);
  return cudaGetLastError();
}
cudaError_t launch_assemble(const SparseDev& d, cudaStream_t st) {
  k_assemble_seg<<<nblk(d.nnz), 256, 0, st>>>(d.nnz, d.seg_ptr, d.seg_src, d.elem_k, d.kval);
  return cudaGetLastError();
}
cudaError_t launch_assemble_fused(const SparseDev& d, const double* eai, cudaStream_t st) {
  const size_t smem = size_t(d.asm_max_loc) * ELEM_K * sizeof(double);
  // the attribute belongs to the (function, device) pair, not to the launch: always the device maximum, set on every call
  // (contexts on other devices or host threads launch the same function with other sizes)
  cudaError_t e = cudaFuncSetAttribute(k_assemble_fused, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
  if (e != cudaSuccess) return e;
  k_assemble_fused<<<nblk(d.nnz, ASM_SLOTS), ASM_THREADS, smem, st>>>(d.nnz, d.seg_ptr, d.seg_loc, d.blk_elem_ptr, d.blk_elems,
                                                              d.elem_nodes, d.elem_type, eai, d.node_xy, d.elem_node0, d.node_rot, d.kval);
  return cudaGetLastError();
}
cudaError_t launch_gather(int64_t n, const int32_t* src, const double* in, double* out, cudaStream_t st) {
  if (n > 0) k_gather<<<nblk(n), 256, 0, st>>>(n, src, in, out);
  return cudaGetLastError();
}
cudaError_t launch_gather_blk_f32(int64_t n_blocks, const int32_t* src, const double* in, float* out, cudaStream_t st) {
  if (n_blocks > 0) k_gather_blk_f32<<<nblk(10 * n_blocks), 256, 0, st>>>(10 * n_blocks, src, in, out);
  return cudaGetLastError();
}
cudaError_t launch_bc_gather(const SparseDev& d, cudaStream_t st) {
  if (d.nnz_ff) k_gather<<<nblk(d.nnz_ff), 256, 0, st>>>(d.nnz_ff, d.ff_src, d.kval, d.ff_val);
  if (d.nnz_bf) k_gather<<<nblk(d.nnz_bf), 256, 0, st>>>(d.nnz_bf, d.bf_src, d.kval, d.bf_val);
  if (d.U_b && d.bb_rowptr && d.nb) {
    // K_bb is only needed when some prescribed displacement is non-zero (solver_static.cpp:196-198)
    k_gather_rows<<<nblk(d.nb), 256, 0, st>>>(d.nb, d.bb_rowptr, d.bb_src, d.kval, d.bb_val);
  }
  return cudaGetLastError();
}
cudaError_t launch_rhs_prescribed(const SparseDev& d, const double* F_f, double* rhs, cudaStream_t st) {
  k_rhs_prescribed<<<1, 256, 0, st>>>(d.nf, d.nb, d.bf_rowptr, d.bf_col, d.bf_val, d.U_b, F_f, rhs);
  return cudaGetLastError();
}
cudaError_t launch_load_vector(const SparseDev& d, const double* nload_val, const double* q, const double* eai, double* F_f,
                               cudaStream_t st) {
  if (d.ELtot)
    k_elem_loads<<<nblk(d.ELtot), 256, 0, st>>>(d.ELtot, d.el_order, d.eload_elem_g, q, d.eload_dir, d.elem_nodes,
                                                 d.elem_type, d.node_xy, d.elem_node0, d.eload_type, d.eload_q2, eai, d.eload_g);
  k_load_vector<<<nblk(d.nf + d.nb), 256, 0, st>>>(d.nf, d.nb, d.fseg_ptr, d.fseg_src, nload_val, d.eload_g, F_f, d.Eq);
  if (d.node_rot) k_rotate_loads<<<nblk(d.Ntot), 256, 0, st>>>(d.Ntot, d.node_rot, d.node_slot, d.node_f0, F_f);
  return cudaGetLastError();
}
cudaError_t launch_stress(const SparseDev& d, const double* eai, const double* q, const double* U_f, double* stress,
                          cudaStream_t st) {
  k_stress<<<nblk(d.Etot), 256, 0, st>>>(d.Etot, d.elem_nodes, d.elem_type, eai, d.node_xy, d.elem_node0, d.elem_f0,
                                         d.node_slot, d.el_ptr, d.el_order, q, d.eload_dir, d.eload_type, d.eload_q2, U_f, d.U_b, d.elem_b0,
                                         stress);
  return cudaGetLastError();
}

}  // namespace obgpu
