// kernels_pcg.cu — Jacobi-preconditioned CG for the large-structure route (replaces the SimplicialLLT branch of
// solveStatic, solver_static.cpp:110-129): memory-bound CSR SpMV fused with the p.Ap dot, and fused
// axpy / preconditioner / dot kernels.  Dots finish in-kernel (last-block pattern, fixed summation order), the
// step lengths alpha/beta are computed on the device from `scal`, so an iteration is 3 launches with no host sync.
#include <cstdlib>

#include "pcg_common.cuh"

namespace obgpu {

// MG = true: the preconditioner is the V-cycle (launched right after): only x, r, p = 0 and b.b here
template <bool MG>
__global__ void __launch_bounds__(PCG_BLOCK) k_pcg_init(PcgArgs a) {
  __shared__ double sh[PCG_BLOCK / 32];
  double v[2] = {0.0, 0.0};
  for (int64_t i = a.r0 + int64_t(blockIdx.x) * PCG_BLOCK + threadIdx.x; i < a.r1; i += int64_t(gridDim.x) * PCG_BLOCK) {
    const double b = a.b[i], z = MG ? 0.0 : a.minv[i] * b;
    a.x[i] = 0.0;
    a.r[i] = b;
    if (!MG) a.z[i] = z;
    a.p[i] = z;
    v[0] += b * z;
    v[1] += b * b;
  }
  double* out[2] = {a.scal + 0, a.scal + 4};
  finish_dots<2>(v, a.partial, a.counter, sh, out);
  if (blockIdx.x == 0 && threadIdx.x == 0) a.scal[1] = 0.0;
}

// ap = A p on rows [r0, r1); scal[2] = p.Ap.  K_ff is in SELL-32 (symbolic_sparse.cpp): one warp per slice of 32
// rows, lane = row; step k reads 32 consecutive values and columns (fully coalesced) and gathers p.  The k loop is
// unrolled by 4 with all matrix loads issued ahead of the gathers.  The matrix streams through once per
// iteration (360 MB at C4, 3x the L2): evict-first loads keep the 16 MB vectors resident in the 126 MB L2.
template <bool STREAM>
__global__ void __launch_bounds__(PCG_BLOCK) k_spmv_dot(PcgArgs a) {
  __shared__ double sh[PCG_BLOCK / 32];
  const int lane = threadIdx.x & 31;
  const int64_t nwarps = int64_t(gridDim.x) * (PCG_BLOCK / 32);
  const int64_t s1 = (int64_t(a.r1) + 31) >> 5;
  double v[1] = {0.0};
  for (int64_t sl = (a.r0 >> 5) + int64_t(blockIdx.x) * (PCG_BLOCK / 32) + (threadIdx.x >> 5); sl < s1; sl += nwarps) {
    const int32_t base = a.slptr[sl], w = (a.slptr[sl + 1] - base) >> 5;
    const double* __restrict__ pv = a.val + base + lane;
    const int32_t* __restrict__ pc = a.col + base + lane;
    double sum = 0.0;
    int k = 0;
    for (; k + 4 <= w; k += 4) {
      int32_t c[4];
      double m[4];
#pragma unroll
      for (int u = 0; u < 4; u++) {
        c[u] = STREAM ? __ldcs(pc + 32 * (k + u)) : pc[32 * (k + u)];
        m[u] = STREAM ? __ldcs(pv + 32 * (k + u)) : pv[32 * (k + u)];
      }
#pragma unroll
      for (int u = 0; u < 4; u++) sum += m[u] * a.p[c[u]];
    }
    for (; k < w; k++) {
      const int32_t c = STREAM ? __ldcs(pc + 32 * k) : pc[32 * k];
      const double m = STREAM ? __ldcs(pv + 32 * k) : pv[32 * k];
      sum += m * a.p[c];
    }
    const int64_t row = sl * 32 + lane;
    if (row >= a.r0 && row < a.r1) {
      a.ap[row] = sum;
      v[0] += sum * a.p[row];
    }
  }
  double* out[1] = {a.scal + 2};
  finish_dots<1>(v, a.partial, a.counter, sh, out);
}

// alpha = rz/pAp; x += alpha p; r -= alpha Ap; z = M^-1 r; rz_next = r.z; rr = r.r
// MG = true: z and r.z come from the V-cycle that follows; only r.r is reduced here
// the same on the 3x3 block-SELL view: warp per slice of 32 block rows (96 rows), lane = block row
template <bool ROWS>
__global__ void __launch_bounds__(PCG_BLOCK) k_spmv_dot_blk(PcgArgs a) {
  __shared__ double sh[PCG_BLOCK / 32];
  const int lane = threadIdx.x & 31;
  const int64_t nwarps = int64_t(gridDim.x) * (PCG_BLOCK / 32);
  const int64_t b0 = a.r0 / 3, b1 = a.r1 / 3;   // block rows of the structure
  const int64_t s1 = (b1 + 31) >> 5;
  double v[1] = {0.0};
  for (int64_t sl = (b0 >> 5) + int64_t(blockIdx.x) * (PCG_BLOCK / 32) + (threadIdx.x >> 5); sl < s1; sl += nwarps) {
    double sum[3];
    block_row_product(a.bslptr, a.bcol, a.bval, sl, lane, a.p, sum);
    if (ROWS) {
      double t[3];
      block_sums_to_rows(sum, lane, t);
#pragma unroll
      for (int j = 0; j < 3; j++) {
        const int64_t row = sl * 96 + 32 * j + lane;
        if (row >= a.r0 && row < a.r1) {
          a.ap[row] = t[j];
          v[0] += t[j] * a.p[row];
        }
      }
    } else {
      const int64_t I = sl * 32 + lane;
      if (I >= b0 && I < b1) {
#pragma unroll
        for (int k = 0; k < 3; k++) {
          a.ap[3 * I + k] = sum[k];
          v[0] += sum[k] * a.p[3 * I + k];
        }
      }
    }
  }
  double* out[1] = {a.scal + 2};
  finish_dots<1>(v, a.partial, a.counter, sh, out);
}

template <bool MG>
__global__ void __launch_bounds__(PCG_BLOCK) k_pcg_update(PcgArgs a, int cur) {
  __shared__ double sh[PCG_BLOCK / 32];
  const double rz = a.scal[cur], pap = a.scal[2];
  const double alpha = pap != 0.0 ? rz / pap : 0.0;
  double v[2] = {0.0, 0.0};
  for (int64_t i = a.r0 + int64_t(blockIdx.x) * PCG_BLOCK + threadIdx.x; i < a.r1; i += int64_t(gridDim.x) * PCG_BLOCK) {
    a.x[i] += alpha * a.p[i];
    const double r = a.r[i] - alpha * a.ap[i];
    a.r[i] = r;
    if (!MG) {
      const double z = a.minv[i] * r;
      a.z[i] = z;
      v[0] += r * z;
    }
    v[1] += r * r;
  }
  if (MG) {
    double w[1] = {v[1]};
    double* out[1] = {a.scal + 3};
    finish_dots<1>(w, a.partial, a.counter, sh, out);
  } else {
    double* out[2] = {a.scal + (cur ^ 1), a.scal + 3};
    finish_dots<2>(v, a.partial, a.counter, sh, out);
  }
}

// r = b - A x on the structure's rows (the TRUE residual, not the recurrence), scal[3] = r.r, and
// scal[5] = sum_i (sum_j |a_ij| |x_j|)^2: the scale of the rounding error of evaluating A x in FP64, i.e. the floor
// below which no residual of this x can be told from zero
__global__ void __launch_bounds__(PCG_BLOCK) k_pcg_true_residual(PcgArgs a) {
  __shared__ double sh[PCG_BLOCK / 32];
  const int lane = threadIdx.x & 31;
  const int64_t nwarps = int64_t(gridDim.x) * (PCG_BLOCK / 32);
  const int64_t s1 = (int64_t(a.r1) + 31) >> 5;
  double v[2] = {0.0, 0.0};
  for (int64_t sl = (a.r0 >> 5) + int64_t(blockIdx.x) * (PCG_BLOCK / 32) + (threadIdx.x >> 5); sl < s1; sl += nwarps) {
    const int32_t base = a.slptr[sl], w = (a.slptr[sl + 1] - base) >> 5;
    double sum = 0.0, asum = 0.0;
    for (int k = 0; k < w; k++) {
      const double t = a.val[base + lane + 32 * k] * a.x[a.col[base + lane + 32 * k]];
      sum += t;
      asum += fabs(t);
    }
    const int64_t row = sl * 32 + lane;
    if (row >= a.r0 && row < a.r1) {
      const double r = a.b[row] - sum;
      a.r[row] = r;
      v[0] += r * r;
      v[1] += asum * asum;
    }
  }
  double* out[2] = {a.scal + 3, a.scal + 5};
  finish_dots<2>(v, a.partial, a.counter, sh, out);
}

// restart of the Jacobi variant from the current r: z = M^-1 r, p = z, scal[cur] = r.z
__global__ void __launch_bounds__(PCG_BLOCK) k_pcg_restart_jacobi(PcgArgs a, int cur) {
  __shared__ double sh[PCG_BLOCK / 32];
  double v[1] = {0.0};
  for (int64_t i = a.r0 + int64_t(blockIdx.x) * PCG_BLOCK + threadIdx.x; i < a.r1; i += int64_t(gridDim.x) * PCG_BLOCK) {
    const double r = a.r[i], z = a.minv[i] * r;
    a.z[i] = z;
    a.p[i] = z;
    v[0] += r * z;
  }
  double* out[1] = {a.scal + cur};
  finish_dots<1>(v, a.partial, a.counter, sh, out);
}
__global__ void __launch_bounds__(PCG_BLOCK) k_pcg_set_p(PcgArgs a) {
  for (int64_t i = a.r0 + int64_t(blockIdx.x) * PCG_BLOCK + threadIdx.x; i < a.r1; i += int64_t(gridDim.x) * PCG_BLOCK)
    a.p[i] = a.z[i];
}

// beta = rz_next/rz; p = z + beta p
__global__ void __launch_bounds__(PCG_BLOCK) k_pcg_dir(PcgArgs a, int cur) {
  const double rz = a.scal[cur], rzn = a.scal[cur ^ 1];
  const double beta = rz != 0.0 ? rzn / rz : 0.0;
  for (int64_t i = a.r0 + int64_t(blockIdx.x) * PCG_BLOCK + threadIdx.x; i < a.r1; i += int64_t(gridDim.x) * PCG_BLOCK)
    a.p[i] = a.z[i] + beta * a.p[i];
}

static unsigned pcg_grid(int64_t work_items, int sm_count, int ctas_per_sm = 8) {
  int64_t g = (work_items + PCG_BLOCK - 1) / PCG_BLOCK;
  const int64_t cap = int64_t(sm_count) * ctas_per_sm;  // persistent-ish: a whole number of CTAs per SM, grid-stride inside
  if (g > cap) g = cap;
  return unsigned(g < 1 ? 1 : g);
}

cudaError_t launch_pcg_init(const PcgArgs& a, int sm_count, cudaStream_t st) {
  if (!a.mg) {
    k_pcg_init<false><<<pcg_grid(a.r1 - a.r0, sm_count), PCG_BLOCK, 0, st>>>(a);
    return cudaGetLastError();
  }
  // x = 0, r = b, p = 0, scal[1] = 0;  z = M r with r.z -> scal[0];  p = z + (scal[0]/scal[1] -> 0) p
  k_pcg_init<true><<<pcg_grid(a.r1 - a.r0, sm_count), PCG_BLOCK, 0, st>>>(a);
  cudaError_t e = launch_mg_vcycle(*a.mg, a.scal + 0, a.partial, a.counter, sm_count, a.ctas_per_sm, st);
  if (e != cudaSuccess) return e;
  k_pcg_dir<<<pcg_grid(a.r1 - a.r0, sm_count, a.ctas_per_sm > 0 ? a.ctas_per_sm : 8), PCG_BLOCK, 0, st>>>(a, 1);
  return cudaGetLastError();
}

static unsigned spmv_grid(const PcgArgs& a, int sm_count, int cps) {
  const int64_t slices = ((int64_t(a.r1) + 31) >> 5) - (a.r0 >> 5);
  return pcg_grid(slices * 32, sm_count, cps);
}

cudaError_t launch_spmv_only(const PcgArgs& a, int sm_count, cudaStream_t st) {
  static const bool blk_rows = getenv("OBGPU_BLK_ROWS") && atoi(getenv("OBGPU_BLK_ROWS")) != 0;
  const int64_t bslices = ((int64_t(a.r1) / 3 + 31) >> 5) - ((a.r0 / 3) >> 5);
  if (a.bval && blk_rows) k_spmv_dot_blk<true><<<balanced_grid(k_spmv_dot_blk<true>, bslices, sm_count, 0), PCG_BLOCK, 0, st>>>(a);
  else if (a.bval) k_spmv_dot_blk<false><<<balanced_grid(k_spmv_dot_blk<false>, bslices, sm_count, 0), PCG_BLOCK, 0, st>>>(a);
  else k_spmv_dot<false><<<spmv_grid(a, sm_count, 8), PCG_BLOCK, 0, st>>>(a);  // timed alone: the whole GPU
  return cudaGetLastError();
}

cudaError_t launch_pcg_true_residual(const PcgArgs& a, int sm_count, cudaStream_t st) {
  k_pcg_true_residual<<<spmv_grid(a, sm_count, a.ctas_per_sm > 0 ? a.ctas_per_sm : 8), PCG_BLOCK, 0, st>>>(a);
  return cudaGetLastError();
}

cudaError_t launch_pcg_restart(const PcgArgs& a, int it, int sm_count, cudaStream_t st) {
  const int cur = it & 1, cps = a.ctas_per_sm > 0 ? a.ctas_per_sm : 8;
  if (!a.mg) {
    k_pcg_restart_jacobi<<<pcg_grid(a.r1 - a.r0, sm_count, cps), PCG_BLOCK, 0, st>>>(a, cur);
    return cudaGetLastError();
  }
  cudaError_t e = launch_mg_vcycle(*a.mg, a.scal + cur, a.partial, a.counter, sm_count, a.ctas_per_sm, st);
  if (e != cudaSuccess) return e;
  k_pcg_set_p<<<pcg_grid(a.r1 - a.r0, sm_count, cps), PCG_BLOCK, 0, st>>>(a);
  return cudaGetLastError();
}

cudaError_t launch_pcg_iter(const PcgArgs& a, int it, int sm_count, cudaStream_t st) {
  const int64_t n = a.r1 - a.r0;
  const int cur = it & 1;
  static const bool stream = getenv("OBGPU_SPMV_STREAM") ? atoi(getenv("OBGPU_SPMV_STREAM")) != 0 : false;
  static const int cps_env = getenv("OBGPU_PCG_CTAS") ? atoi(getenv("OBGPU_PCG_CTAS")) : 0;
  const int cps = cps_env > 0 ? cps_env : a.ctas_per_sm > 0 ? a.ctas_per_sm : 8;
  const int64_t slices = ((int64_t(a.r1) + 31) >> 5) - (a.r0 >> 5);
  const unsigned g = pcg_grid(slices * 32, sm_count, cps);  // one warp per slice, grid-stride
  static const bool blk_rows = getenv("OBGPU_BLK_ROWS") && atoi(getenv("OBGPU_BLK_ROWS")) != 0;  // experiment: see k_mg_spmv_blk
  const int64_t bslices = ((int64_t(a.r1) / 3 + 31) >> 5) - ((a.r0 / 3) >> 5);
  if (a.bval && blk_rows) k_spmv_dot_blk<true><<<balanced_grid(k_spmv_dot_blk<true>, bslices, sm_count, cps), PCG_BLOCK, 0, st>>>(a);
  else if (a.bval) k_spmv_dot_blk<false><<<balanced_grid(k_spmv_dot_blk<false>, bslices, sm_count, cps), PCG_BLOCK, 0, st>>>(a);
  else if (stream) k_spmv_dot<true><<<g, PCG_BLOCK, 0, st>>>(a);
  else k_spmv_dot<false><<<g, PCG_BLOCK, 0, st>>>(a);
  static const int cpv_env = getenv("OBGPU_PCG_VEC_CTAS") ? atoi(getenv("OBGPU_PCG_VEC_CTAS")) : 0;
  const int cpv = cpv_env > 0 ? cpv_env : a.ctas_per_sm > 0 ? a.ctas_per_sm : 8;
  if (a.mg) {
    k_pcg_update<true><<<pcg_grid(n, sm_count, cpv), PCG_BLOCK, 0, st>>>(a, cur);
    cudaError_t e = launch_mg_vcycle(*a.mg, a.scal + (cur ^ 1), a.partial, a.counter, sm_count, a.ctas_per_sm, st);
    if (e != cudaSuccess) return e;
  } else {
    k_pcg_update<false><<<pcg_grid(n, sm_count, cpv), PCG_BLOCK, 0, st>>>(a, cur);
  }
  k_pcg_dir<<<pcg_grid(n, sm_count, cpv), PCG_BLOCK, 0, st>>>(a, cur);
  return cudaGetLastError();
}

__global__ void __launch_bounds__(256) k_inv_diag(int64_t nf, const int32_t* __restrict__ ff_diag,
                                                  const double* __restrict__ ff_val, double* __restrict__ minv) {
  const int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < nf) minv[i] = 1.0 / ff_val[ff_diag[i]];
}
cudaError_t launch_inv_diag(int64_t nf, const int32_t* ff_diag, const double* ff_val, double* minv, cudaStream_t st) {
  k_inv_diag<<<unsigned((nf + 255) / 256), 256, 0, st>>>(nf, ff_diag, ff_val, minv);
  return cudaGetLastError();
}

// F_b = K_bf U_f - equivalent loads on the bounded DOFs (solver_static.cpp:196-218); one warp per bounded row
// bb_*: K_bb U_b with K_bb holding its upper triangle only, used as a general matrix (assemble.cpp:258-262,
// solver_static.cpp:197: SURVEY quirk Q1); null when every prescribed displacement is zero
__global__ void __launch_bounds__(256) k_reactions(int64_t nb, const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col,
                                                   const double* __restrict__ val, const double* __restrict__ U,
                                                   const double* __restrict__ Eq, const int32_t* __restrict__ bb_rowptr,
                                                   const int32_t* __restrict__ bb_col, const double* __restrict__ bb_val,
                                                   const double* __restrict__ U_b, double* __restrict__ F_b) {
  const int64_t row = (int64_t(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (row >= nb) return;
  double s = 0.0;
  for (int32_t k = rowptr[row] + lane; k < rowptr[row + 1]; k += 32) s += val[k] * U[col[k]];
  if (U_b)
    for (int32_t k = bb_rowptr[row] + lane; k < bb_rowptr[row + 1]; k += 32) s += bb_val[k] * U_b[bb_col[k]];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) F_b[row] = s - Eq[row];
}
cudaError_t launch_reactions(const SparseDev& d, const double* U_f, double* F_b, cudaStream_t st) {
  if (d.nb == 0) return cudaSuccess;
  k_reactions<<<unsigned((d.nb * 32 + 255) / 256), 256, 0, st>>>(d.nb, d.bf_rowptr, d.bf_col, d.bf_val, U_f, d.Eq, d.bb_rowptr,
                                                                  d.bb_col, d.bb_val, d.U_b, F_b);
  return cudaGetLastError();
}

}  // namespace obgpu
