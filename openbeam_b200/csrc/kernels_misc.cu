// kernels_misc.cu — small helper kernels: per-system summaries for the final NCCL gather.
#include "kernels.h"

namespace obgpu {

// One thread per system (c, s): {max|U_f|, sum of F_b over DX dofs, sum of F_b over DY dofs, info}.
// Results are laid out [L][S][4] like OBGPU_RES_INFO.  Memory-bound: every U_f/F_b value is read once.
__global__ void __launch_bounds__(256) k_summaries(SummaryArgs a) {
  const int64_t total = a.L * a.S;
  for (int64_t sys = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; sys < total; sys += int64_t(gridDim.x) * blockDim.x) {
    const int64_t c = sys / a.S, s = sys - c * a.S;
    const int64_t f0 = a.free_off[s], f1 = a.free_off[s + 1], b0 = a.bnd_off[s], b1 = a.bnd_off[s + 1];
    const double* U = a.U_f + c * a.nf_tot;
    const double* F = a.F_b + c * a.nb_tot;
    double umax = 0.0, fx = 0.0, fy = 0.0;
    for (int64_t i = f0; i < f1; i++) umax = fmax(umax, fabs(U[i]));
    for (int64_t i = b0; i < b1; i++) {
      const int d = a.bnd_dofid[i];
      const double v = F[i];
      if (d == OBGPU_DX) fx += v;
      if (d == OBGPU_DY) fy += v;
    }
    double4 o = make_double4(umax, fx, fy, double(a.info[sys]));
    reinterpret_cast<double4*>(a.out)[sys] = o;
  }
}

// The same record for LARGE systems (the sparse route: one to a few structures of 10^5..10^6 DOFs): one CTA per system, its
// threads stride over the DOFs, fixed-order reductions (lane tree, then warp 0 over the warp results) — deterministic,
// and 2 M DOFs take microseconds where a single thread of k_summaries took 125 ms.
__global__ void __launch_bounds__(256) k_summaries_block(SummaryArgs a) {
  __shared__ double red[3][8];
  const int64_t total = a.L * a.S;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int64_t sys = blockIdx.x; sys < total; sys += gridDim.x) {
    const int64_t c = sys / a.S, s = sys - c * a.S;
    const int64_t f0 = a.free_off[s], f1 = a.free_off[s + 1], b0 = a.bnd_off[s], b1 = a.bnd_off[s + 1];
    const double* U = a.U_f + c * a.nf_tot;
    const double* F = a.F_b + c * a.nb_tot;
    double v[3] = {0.0, 0.0, 0.0};  // max|U_f|, sum F_b over DX, sum F_b over DY
    for (int64_t i = f0 + threadIdx.x; i < f1; i += 256) v[0] = fmax(v[0], fabs(U[i]));
    for (int64_t i = b0 + threadIdx.x; i < b1; i += 256) {
      const int d = a.bnd_dofid[i];
      const double x = F[i];
      if (d == OBGPU_DX) v[1] += x;
      if (d == OBGPU_DY) v[2] += x;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      v[0] = fmax(v[0], __shfl_xor_sync(0xffffffffu, v[0], o));
      v[1] += __shfl_xor_sync(0xffffffffu, v[1], o);
      v[2] += __shfl_xor_sync(0xffffffffu, v[2], o);
    }
    if (lane == 0) { red[0][warp] = v[0]; red[1][warp] = v[1]; red[2][warp] = v[2]; }
    __syncthreads();
    if (threadIdx.x == 0) {
      double umax = 0.0, fx = 0.0, fy = 0.0;
      for (int w = 0; w < 8; w++) { umax = fmax(umax, red[0][w]); fx += red[1][w]; fy += red[2][w]; }
      reinterpret_cast<double4*>(a.out)[sys] = make_double4(umax, fx, fy, double(a.info[sys]));
    }
    __syncthreads();
  }
}

cudaError_t launch_summaries(const SummaryArgs& a, int sm_count, cudaStream_t st) {
  const int64_t total = a.L * a.S;
  if (a.S > 0 && a.nf_tot / a.S >= 4096) {  // large systems: a CTA each
    int64_t g = total < int64_t(sm_count) * 8 ? total : int64_t(sm_count) * 8;
    k_summaries_block<<<unsigned(g < 1 ? 1 : g), 256, 0, st>>>(a);
    return cudaGetLastError();
  }
  int64_t grid = (total + 255) / 256;
  const int64_t cap = int64_t(sm_count) * 8;
  if (grid > cap) grid = cap;
  if (grid < 1) grid = 1;
  k_summaries<<<unsigned(grid), 256, 0, st>>>(a);
  return cudaGetLastError();
}

// full element stress records [n][2][6] -> compact [n][2][3] = (N, Vy, Mz) per face
__global__ void __launch_bounds__(256) k_compact_stress(const double* __restrict__ full, double* __restrict__ out, int64_t n6) {
  for (int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n6; i += int64_t(gridDim.x) * blockDim.x) {
    const int64_t rec = i / 6;
    const int k = int(i - rec * 6);  // 0..2 face 0, 3..5 face 1
    const int src = k < 3 ? (k == 2 ? 5 : k) : (k == 5 ? 11 : k + 3);
    out[i] = full[rec * 12 + src];
  }
}
cudaError_t launch_compact_stress(const double* full, double* compact, int64_t n_elem_records, cudaStream_t st) {
  const int64_t n6 = n_elem_records * 6;
  if (n6 == 0) return cudaSuccess;
  int64_t grid = (n6 + 255) / 256;
  if (grid > 148 * 16) grid = 148 * 16;
  k_compact_stress<<<unsigned(grid), 256, 0, st>>>(full, compact, n6);
  return cudaGetLastError();
}

}  // namespace obgpu
