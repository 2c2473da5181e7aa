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
                                                    double* __restrict__ elem_k) {
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

__global__ void __launch_bounds__(256) k_gather(int64_t n, const int32_t* __restrict__ src, const double* __restrict__ in,
                                                double* __restrict__ out) {
  const int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < n) {
    const int32_t s = src[i];
    out[i] = s >= 0 ? in[s] : 0.0;  // -1: padding entry of the SELL layout
  }
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

__global__ void __launch_bounds__(256) k_stress(int64_t Etot, const uint2* __restrict__ elem_nodes,
                                                const uint8_t* __restrict__ elem_type, const double* __restrict__ eai,
                                                const double2* __restrict__ node_xy, const int32_t* __restrict__ elem_node0,
                                                const int32_t* __restrict__ elem_f0, const int32_t* __restrict__ node_slot,
                                                const int32_t* __restrict__ el_ptr, const int32_t* __restrict__ el_order,
                                                const double* __restrict__ q, const double* __restrict__ dir,
                                                const uint8_t* __restrict__ eload_type, const double* __restrict__ eload_q2,
                                                const double* __restrict__ U_f, double* __restrict__ stress) {
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
    ug[k] = t0 >= 0 ? U[t0] : 0.0;
    ug[3 + k] = t1 >= 0 ? U[t1] : 0.0;
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
  k_elem_stiff<<<nblk(d.Etot), 256, 0, st>>>(d.Etot, d.elem_nodes, d.elem_type, eai, d.node_xy, d.elem_node0, d.elem_k);
  return cudaGetLastError();
}
cudaError_t launch_assemble(const SparseDev& d, cudaStream_t st) {
  k_assemble_seg<<<nblk(d.nnz), 256, 0, st>>>(d.nnz, d.seg_ptr, d.seg_src, d.elem_k, d.kval);
  return cudaGetLastError();
}
cudaError_t launch_bc_gather(const SparseDev& d, cudaStream_t st) {
  if (d.nnz_ff) k_gather<<<nblk(d.nnz_ff), 256, 0, st>>>(d.nnz_ff, d.ff_src, d.kval, d.ff_val);
  if (d.nnz_bf) k_gather<<<nblk(d.nnz_bf), 256, 0, st>>>(d.nnz_bf, d.bf_src, d.kval, d.bf_val);
  return cudaGetLastError();
}
cudaError_t launch_load_vector(const SparseDev& d, const double* nload_val, const double* q, const double* eai, double* F_f,
                               cudaStream_t st) {
  if (d.ELtot)
    k_elem_loads<<<nblk(d.ELtot), 256, 0, st>>>(d.ELtot, d.el_order, d.eload_elem_g, q, d.eload_dir, d.elem_nodes,
                                                 d.elem_type, d.node_xy, d.elem_node0, d.eload_type, d.eload_q2, eai, d.eload_g);
  k_load_vector<<<nblk(d.nf + d.nb), 256, 0, st>>>(d.nf, d.nb, d.fseg_ptr, d.fseg_src, nload_val, d.eload_g, F_f, d.Eq);
  return cudaGetLastError();
}
cudaError_t launch_stress(const SparseDev& d, const double* eai, const double* q, const double* U_f, double* stress,
                          cudaStream_t st) {
  k_stress<<<nblk(d.Etot), 256, 0, st>>>(d.Etot, d.elem_nodes, d.elem_type, eai, d.node_xy, d.elem_node0, d.elem_f0,
                                         d.node_slot, d.el_ptr, d.el_order, q, d.eload_dir, d.eload_type, d.eload_q2, U_f, stress);
  return cudaGetLastError();
}

}  // namespace obgpu
