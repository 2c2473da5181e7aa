// kernels_mg.cu — aggregation-multigrid V-cycle used as the CG preconditioner of the sparse route (OBGPU_PCG_MG).
// It stands where the reference's sparse branch puts its factorisation (Eigen::SimplicialLLT, solver_static.cpp:110-129).
//
//   coarse spaces  rigid-body modes (tx, ty, rot) of boxes of nodes (plan: symbolic_mg.cpp), unsmoothed prolongator,
//                  Galerkin operators A_c = P^T A P built on the device from the current K_ff values
//   smoother       Chebyshev(2) on D^-1 A, interval [lambda/8, lambda], lambda = Gershgorin bound; pre = post (symmetric)
//   coarsest       explicit dense inverse (<= MgPlan::DENSE_MAX rows)
//   cycle          V, coarse correction over-relaxed by omega (compensates the unsmoothed prolongator's energy loss)
//
// Every level is a SELL-32 matrix and every smoothing step is ONE kernel: SpMV (warp per slice of 32 rows, lane = row)
// with the row-local Chebyshev update fused behind it.  All HBM-bound; no atomics on values (the only atomic is the
// order-independent max of the Gershgorin bound), so the preconditioner — hence the iterates — are reproducible.
#include <cstdlib>

#include <cooperative_groups.h>

#include "pcg_common.cuh"

namespace cg = cooperative_groups;

namespace obgpu {

// ---------------------------------------------------------------- numeric setup
// thread = coarse row R = 3 I + a: sum over the aggregate's fine rows i and their entries (i, j):
//   A_c(3I+a, 3J+b) += P_i[a] * A(i,j) * P_j[b],   P_i = e_kind (+ w on the rotation column)
__global__ void __launch_bounds__(128) k_mg_galerkin(MgGalerkinArgs g) {
  const int64_t R = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (R >= g.n_coarse) return;
  const int32_t cbase = g.c_slptr[R >> 5] + int32_t(R & 31), cw = (g.c_slptr[(R >> 5) + 1] - g.c_slptr[R >> 5]) >> 5;
  for (int32_t k = 0; k < cw; k++) g.c_val[cbase + 32 * k] = 0.0;
  const int64_t I = R / 3;
  const int a = int(R - 3 * I);
  int32_t hint = 0;  // block slot of the previous hit: consecutive entries mostly fall into the same aggregate
  for (int32_t m = g.mem_ptr[I]; m < g.mem_ptr[I + 1]; m++) {
    const int32_t i = g.mem[m];
    const int ki = g.f_kind[i];
    const double pa = a == 2 ? g.f_w[i] : (a == ki ? 1.0 : 0.0);
    if (pa == 0.0) continue;
    const int32_t fbase = g.f_slptr[i >> 5] + (i & 31), fw = (g.f_slptr[(i >> 5) + 1] - g.f_slptr[i >> 5]) >> 5;
    for (int32_t k = 0; k < fw; k++) {
      const double v = g.f_val[fbase + 32 * k];
      if (v == 0.0) continue;  // padding and structural zeros
      const int32_t j = g.f_col[fbase + 32 * k];
      const int32_t jb = g.f_agg[j];
      if (jb < 0) continue;
      if (g.c_col[cbase + 32 * hint] != jb) {
        hint = -1;
        for (int32_t q = 0; q < cw; q += 3)
          if (g.c_col[cbase + 32 * q] == jb) { hint = q; break; }
        if (hint < 0) { hint = 0; continue; }  // cannot happen: the pattern was built from these entries
      }
      const int kj = g.f_kind[j];
      const double t = pa * v;
      if (kj < 2) g.c_val[cbase + 32 * (hint + kj)] += t;
      g.c_val[cbase + 32 * (hint + 2)] += t * g.f_w[j];
    }
  }
}
cudaError_t launch_mg_galerkin(const MgGalerkinArgs& g, cudaStream_t st) {
  if (g.n_coarse == 0) return cudaSuccess;
  k_mg_galerkin<<<unsigned((g.n_coarse + 127) / 128), 128, 0, st>>>(g);
  return cudaGetLastError();
}

__global__ void __launch_bounds__(256) k_mg_diag(int64_t r0, int64_t r1, const int32_t* __restrict__ slptr,
                                                 const int32_t* __restrict__ col, const double* __restrict__ val,
                                                 double* __restrict__ minv, double* lam) {
  const int64_t r = r0 + int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  double ratio = 0.0;
  if (r < r1) {
    const int32_t base = slptr[r >> 5] + int32_t(r & 31), w = (slptr[(r >> 5) + 1] - slptr[r >> 5]) >> 5;
    double s = 0.0, d = 0.0;
    for (int32_t k = 0; k < w; k++) {
      const double v = val[base + 32 * k];
      s += fabs(v);
      if (col[base + 32 * k] == r) d += v;  // padding entries carry value 0
    }
    const bool ok = d > 0.0;
    minv[r] = ok ? 1.0 / d : 0.0;
    ratio = ok ? s / d : 0.0;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) ratio = fmax(ratio, __shfl_xor_sync(0xffffffffu, ratio, o));
  if ((threadIdx.x & 31) == 0 && ratio > 0.0)  // positive doubles order like their bit patterns
    atomicMax(reinterpret_cast<unsigned long long*>(lam), (unsigned long long)__double_as_longlong(ratio));
}
cudaError_t launch_mg_diag(int64_t r0, int64_t r1, const int32_t* slptr, const int32_t* col, const double* val, double* minv,
                           double* lam, cudaStream_t st) {
  if (r1 <= r0) return cudaSuccess;
  k_mg_diag<<<unsigned((r1 - r0 + 255) / 256), 256, 0, st>>>(r0, r1, slptr, col, val, minv, lam);
  return cudaGetLastError();
}

// Gauss-Jordan inverse of the SPD coarsest matrix (no pivoting) on MG_INV_CTAS co-resident CTAs (cooperative launch):
// rows are dealt cyclically to the CTAs and live in shared memory for the whole elimination; step k needs one thing
// from outside the CTA — the scaled pivot row — which its owner publishes to a double-buffered global row right after
// finishing step k-1 on it, so a step costs ONE grid barrier.  A pivot that is not positive relative to the largest
// original diagonal marks a null direction (an aggregate without a rotation-capable member, a mechanism): its row and
// column of the inverse are zeroed, which keeps the preconditioner symmetric positive semi-definite there.
constexpr int MG_INV_CTAS = 64, MG_INV_THREADS = 256;
__global__ void __launch_bounds__(MG_INV_THREADS) k_mg_dense_inverse(int32_t r0, int32_t nc, const int32_t* __restrict__ slptr,
                                                                     const int32_t* __restrict__ col,
                                                                     const double* __restrict__ val, double* __restrict__ inv,
                                                                     double* __restrict__ prow /* [2][nc + 2] */) {
  extern __shared__ double sA[];  // [rows of this CTA][nc]
  cg::grid_group grid = cg::this_grid();
  const int tid = threadIdx.x, nt = blockDim.x, cta = blockIdx.x, ncta = gridDim.x;
  const int my_rows = (nc - cta + ncta - 1) / ncta;  // rows cta, cta + ncta, ...
  for (int q = tid; q < my_rows * nc; q += nt) sA[q] = 0.0;
  __syncthreads();
  for (int lr = tid; lr < my_rows; lr += nt) {
    const int64_t r = int64_t(r0) + cta + int64_t(lr) * ncta;
    const int32_t base = slptr[r >> 5] + int32_t(r & 31), w = (slptr[(r >> 5) + 1] - slptr[r >> 5]) >> 5;
    for (int32_t k = 0; k < w; k++) {
      const int32_t c = col[base + 32 * k] - r0;
      const double v = val[base + 32 * k];
      if (v != 0.0 && c >= 0 && c < nc) sA[lr * nc + c] += v;
    }
  }
  __syncthreads();
  // largest original diagonal -> null-direction threshold (every CTA computes the same value from global data)
  __shared__ double s_red[MG_INV_THREADS / 32];
  __shared__ double s_tiny;
  {
    double m = 0.0;
    for (int i = tid; i < nc; i += nt) {
      const int64_t r = int64_t(r0) + i;
      const int32_t base = slptr[r >> 5] + int32_t(r & 31), w = (slptr[(r >> 5) + 1] - slptr[r >> 5]) >> 5;
      for (int32_t k = 0; k < w; k++)
        if (col[base + 32 * k] == r) m = fmax(m, val[base + 32 * k]);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((tid & 31) == 0) s_red[tid >> 5] = m;
    __syncthreads();
    if (tid == 0) {
      for (int w2 = 1; w2 < MG_INV_THREADS / 32; w2++) m = fmax(m, s_red[w2]);
      s_tiny = 1e-13 * m;
    }
    __syncthreads();
  }
  const double tiny = s_tiny;
  // publish pivot row k (scaled by 1/pivot, with [nc] = 1/pivot and [nc+1] = skip flag) from its owner's shared memory
  auto publish = [&](int k) {
    if (k >= nc || k % ncta != cta) return;
    const double* row = sA + (k / ncta) * nc;
    double* out = prow + (k & 1) * (nc + 2);
    const double p = row[k];
    const bool skip = !(p > tiny);
    const double pinv = skip ? 0.0 : 1.0 / p;
    for (int j = tid; j < nc; j += nt) out[j] = row[j] * pinv;
    if (tid == 0) { out[nc] = pinv; out[nc + 1] = skip ? 1.0 : 0.0; }
  };
  publish(0);
  __threadfence();
  grid.sync();
  for (int k = 0; k < nc; k++) {
    const double* __restrict__ pk = prow + (k & 1) * (nc + 2);
    const double pinv = __ldcg(pk + nc);
    const bool skip = __ldcg(pk + nc + 1) != 0.0;
    for (int lr = tid >> 5; lr < my_rows; lr += nt >> 5) {   // one warp per owned row
      const int i = cta + lr * ncta;
      double* row = sA + lr * nc;
      const int lane = tid & 31;
      if (skip) {
        if (i == k) { for (int j = lane; j < nc; j += 32) row[j] = 0.0; }
        else if (lane == 0) row[k] = 0.0;
        continue;
      }
      if (i == k) {
        for (int j = lane; j < nc; j += 32) row[j] = j == k ? pinv : __ldcg(pk + j);
      } else {
        const double f = row[k];
        __syncwarp();
        if (f != 0.0)
          for (int j = lane; j < nc; j += 32) row[j] = j == k ? -f * pinv : row[j] - f * __ldcg(pk + j);
      }
    }
    __syncthreads();
    publish(k + 1);
    __threadfence();
    grid.sync();
  }
  for (int lr = 0; lr < my_rows; lr++) {
    const int i = cta + lr * ncta;
    for (int j = tid; j < nc; j += nt) inv[int64_t(i) * nc + j] = sA[lr * nc + j];
  }
}
cudaError_t launch_mg_dense_inverse(int32_t r0, int32_t nc, const int32_t* slptr, const int32_t* col, const double* val,
                                    double* inv, double* prow, cudaStream_t st) {
  if (nc <= 0) return cudaSuccess;
  int ncta = MG_INV_CTAS;
  if (ncta > nc) ncta = nc;
  const size_t smem = size_t((nc + ncta - 1) / ncta) * nc * sizeof(double);
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(k_mg_dense_inverse, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    if (e != cudaSuccess) return e;
    attr_set = true;
  }
  void* args[] = {&r0, &nc, &slptr, &col, &val, &inv, &prow};
  return cudaLaunchCooperativeKernel((void*)k_mg_dense_inverse, dim3(unsigned(ncta)), dim3(MG_INV_THREADS), args, smem, st);
}

// ---------------------------------------------------------------- V-cycle kernels
// first Chebyshev step from a zero initial guess: d0 = c0 D^-1 b, x = d0
__global__ void __launch_bounds__(PCG_BLOCK) k_mg_smooth0(MgLevelDev L) {
  for (int64_t i = L.r0 + int64_t(blockIdx.x) * PCG_BLOCK + threadIdx.x; i < L.r1; i += int64_t(gridDim.x) * PCG_BLOCK) {
    const double d = L.c0 * L.minv[i] * L.b[i];
    L.d0[i] = d;
    L.x[i] = d;
  }
}

// t = A din on the structure's rows, then row-locally
//   CHEB : r' = rin - t;  d = c1 din + c2 D^-1 r';  r = r';  dout = d;  x += d      (DOT: sum x'[row] * dotv[row])
//   RESID: r  = rin - t
enum { MG_CHEB = 0, MG_RESID = 1 };
template <int MODE, bool DOT>
__global__ void __launch_bounds__(PCG_BLOCK) k_mg_spmv(MgLevelDev L, const double* __restrict__ din, const double* rin,
                                                       double* dout, double c1, double c2, const double* __restrict__ dotv,
                                                       double* dot_out, double* partial, unsigned int* counter) {
  __shared__ double sh[PCG_BLOCK / 32];
  const int lane = threadIdx.x & 31;
  const int64_t nwarps = int64_t(gridDim.x) * (PCG_BLOCK / 32);
  const int64_t s1 = (int64_t(L.r1) + 31) >> 5;
  double v[1] = {0.0};
  for (int64_t sl = (L.r0 >> 5) + int64_t(blockIdx.x) * (PCG_BLOCK / 32) + (threadIdx.x >> 5); sl < s1; sl += nwarps) {
    const int32_t base = L.slptr[sl], w = (L.slptr[sl + 1] - base) >> 5;
    const double* __restrict__ pv = L.val + base + lane;
    const int32_t* __restrict__ pc = L.col + base + lane;
    double sum = 0.0;
    int k = 0;
    for (; k + 4 <= w; k += 4) {
      int32_t c[4];
      double m[4];
#pragma unroll
      for (int u = 0; u < 4; u++) {
        c[u] = pc[32 * (k + u)];
        m[u] = pv[32 * (k + u)];
      }
#pragma unroll
      for (int u = 0; u < 4; u++) sum += m[u] * din[c[u]];
    }
    for (; k < w; k++) sum += pv[32 * k] * din[pc[32 * k]];
    const int64_t row = sl * 32 + lane;
    if (row >= L.r0 && row < L.r1) {
      const double rn = rin[row] - sum;
      L.r[row] = rn;
      if (MODE == MG_CHEB) {
        const double d = c1 * din[row] + c2 * L.minv[row] * rn;
        dout[row] = d;
        const double xn = L.x[row] + d;
        L.x[row] = xn;
        if (DOT) v[0] += xn * dotv[row];
      }
    }
  }
  if (DOT) {
    double* out[1] = {dot_out};
    finish_dots<1>(v, partial, counter, sh, out);
  }
}

// the same on the 3x3 block-SELL view of the level: warp per slice of 32 block rows, lane = block row (3 rows)
template <int MODE, bool DOT, class T>
__global__ void __launch_bounds__(PCG_BLOCK) k_mg_spmv_blk(MgLevelDev L, const double* __restrict__ din, const double* rin,
                                                           double* dout, double c1, double c2, const double* __restrict__ dotv,
                                                           double* dot_out, double* partial, unsigned int* counter) {
  __shared__ double sh[PCG_BLOCK / 32];
  const int lane = threadIdx.x & 31;
  const int64_t nwarps = int64_t(gridDim.x) * (PCG_BLOCK / 32);
  const int64_t b0 = L.r0 / 3, b1 = L.r1 / 3;
  const int64_t s1 = (b1 + 31) >> 5;
  double v[1] = {0.0};
  for (int64_t sl = (b0 >> 5) + int64_t(blockIdx.x) * (PCG_BLOCK / 32) + (threadIdx.x >> 5); sl < s1; sl += nwarps) {
    // (redistributing the sums so that lane l serves rows l, 32 + l, 64 + l — fully coalesced vector accesses — measured
    // SLOWER on B200 than letting every lane finish its own three consecutive rows: 97.6 vs 96.0 us here, 83 vs 60 us in
    // k_spmv_dot_blk)
    double sum[3];
    if (sizeof(T) == 4) block_row_product_f32(L.bslptr, L.bcol, L.bval32, sl, lane, din, sum);
    else block_row_product(L.bslptr, L.bcol, L.bval, sl, lane, din, sum);
    const int64_t I = sl * 32 + lane;
#pragma unroll
    for (int j = 0; j < 3; j++) {
      const int64_t row = 3 * I + j;
      if (I >= b0 && I < b1) {
        const double rn = rin[row] - sum[j];
        L.r[row] = rn;
        if (MODE == MG_CHEB) {
          const double d = c1 * din[row] + c2 * L.minv[row] * rn;
          dout[row] = d;
          const double xn = L.x[row] + d;
          L.x[row] = xn;
          if (DOT) v[0] += xn * dotv[row];
        }
      }
    }
  }
  if (DOT) {
    double* out[1] = {dot_out};
    finish_dots<1>(v, partial, counter, sh, out);
  }
}

// b_c = P^T r (thread = coarse row, fixed member order), optionally followed by the coarse level's first smoothing step
__global__ void __launch_bounds__(PCG_BLOCK) k_mg_restrict(MgLevelDev F, MgLevelDev C, int smooth0) {
  for (int64_t R = C.r0 + int64_t(blockIdx.x) * PCG_BLOCK + threadIdx.x; R < C.r1; R += int64_t(gridDim.x) * PCG_BLOCK) {
    const int64_t I = R / 3;
    const int a = int(R - 3 * I);
    double s = 0.0;
    for (int32_t m = F.mem_ptr[I]; m < F.mem_ptr[I + 1]; m++) {
      const int32_t i = F.mem[m];
      const double pa = a == 2 ? F.f_w[i] : (a == F.f_kind[i] ? 1.0 : 0.0);
      s += pa * F.r[i];
    }
    C.b[R] = s;
    if (smooth0) {
      const double d = C.c0 * C.minv[R] * s;
      C.d0[R] = d;
      C.x[R] = d;
    }
  }
}

// x_c = A_c^-1 b_c, one warp per row of the explicit inverse
__global__ void __launch_bounds__(PCG_BLOCK) k_mg_dense_apply(MgLevelDev C, const double* __restrict__ inv) {
  const int nc = C.r1 - C.r0;
  const int lane = threadIdx.x & 31;
  for (int row = blockIdx.x * (PCG_BLOCK / 32) + (threadIdx.x >> 5); row < nc; row += gridDim.x * (PCG_BLOCK / 32)) {
    const double* __restrict__ a = inv + int64_t(row) * nc;
    double s = 0.0;
    for (int j = lane; j < nc; j += 32) s += a[j] * C.b[C.r0 + j];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) C.x[C.r0 + row] = s;
  }
}

__global__ void __launch_bounds__(PCG_BLOCK) k_mg_dot(MgLevelDev L, double* dot_out, double* partial, unsigned int* counter) {
  __shared__ double sh[PCG_BLOCK / 32];
  double v[1] = {0.0};
  for (int64_t i = L.r0 + int64_t(blockIdx.x) * PCG_BLOCK + threadIdx.x; i < L.r1; i += int64_t(gridDim.x) * PCG_BLOCK)
    v[0] += L.x[i] * L.b[i];
  double* out[1] = {dot_out};
  finish_dots<1>(v, partial, counter, sh, out);
}

// e = omega P x_c;  d0 = e (the post-smoother starts with r -= A e);  x += e
__global__ void __launch_bounds__(PCG_BLOCK) k_mg_prolong(MgLevelDev F, const double* __restrict__ xc, double omega) {
  for (int64_t i = F.r0 + int64_t(blockIdx.x) * PCG_BLOCK + threadIdx.x; i < F.r1; i += int64_t(gridDim.x) * PCG_BLOCK) {
    const int32_t jb = F.f_agg[i];
    double e = 0.0;
    if (jb >= 0) {
      const int k = F.f_kind[i];
      e = F.f_w[i] * xc[jb + 2];
      if (k < 2) e += xc[jb + k];
      e *= omega;
    }
    F.d0[i] = e;
    F.x[i] += e;
  }
}

static unsigned mg_grid(int64_t items, int sm_count, int cps) {
  int64_t g = (items + PCG_BLOCK - 1) / PCG_BLOCK;
  const int64_t cap = int64_t(sm_count) * (cps > 0 ? cps : 8);
  if (g > cap) g = cap;
  return unsigned(g < 1 ? 1 : g);
}

// one launch of the block kernel, FP32 or FP64 matrix copy as the level has it
template <int MODE, bool DOT>
static void launch_blk(const MgLevelDev& L, int64_t slices, int sm_count, int cps, cudaStream_t st, const double* din,
                       const double* rin, double* dout, double c1, double c2, const double* dotv, double* dot_out,
                       double* partial, unsigned int* counter) {
  if (L.bval32)
    k_mg_spmv_blk<MODE, DOT, float><<<balanced_grid(k_mg_spmv_blk<MODE, DOT, float>, slices, sm_count, cps), PCG_BLOCK, 0, st>>>(
        L, din, rin, dout, c1, c2, dotv, dot_out, partial, counter);
  else
    k_mg_spmv_blk<MODE, DOT, double><<<balanced_grid(k_mg_spmv_blk<MODE, DOT, double>, slices, sm_count, cps), PCG_BLOCK, 0, st>>>(
        L, din, rin, dout, c1, c2, dotv, dot_out, partial, counter);
}

cudaError_t launch_mg_cheb_only(const MgLevelDev& L, int sm_count, cudaStream_t st) {
  if (L.bval) {
    const int64_t bs = ((int64_t(L.r1) / 3 + 31) >> 5) - ((L.r0 / 3) >> 5);
    launch_blk<MG_CHEB, false>(L, bs, sm_count, 0, st, L.d0, L.r, L.d1, L.c1, L.c2, nullptr, nullptr, nullptr, nullptr);
    return cudaGetLastError();
  }
  const int64_t threads = (((int64_t(L.r1) + 31) >> 5) - (L.r0 >> 5)) * 32;
  k_mg_spmv<MG_CHEB, false><<<mg_grid(threads, sm_count, 0), PCG_BLOCK, 0, st>>>(L, L.d0, L.r, L.d1, L.c1, L.c2, nullptr, nullptr,
                                                                                nullptr, nullptr);
  return cudaGetLastError();
}

cudaError_t launch_mg_vcycle(const MgJob& J, double* dot_out, double* partial, unsigned int* counter, int sm_count,
                             int cps, cudaStream_t st) {
  const int nl = J.nlev;
  auto rows = [](const MgLevelDev& L) { return int64_t(L.r1 - L.r0); };
  auto slice_threads = [](const MgLevelDev& L) { return (((int64_t(L.r1) + 31) >> 5) - (L.r0 >> 5)) * 32; };
  auto blk_slices = [](const MgLevelDev& L) { return ((int64_t(L.r1) / 3 + 31) >> 5) - ((L.r0 / 3) >> 5); };
  auto cheb = [&](const MgLevelDev& L, const double* din, const double* rin, double* dout, double c1, double c2, bool dot) {
    if (L.bval) {
      if (dot) launch_blk<MG_CHEB, true>(L, blk_slices(L), sm_count, cps, st, din, rin, dout, c1, c2, L.b, dot_out, partial, counter);
      else launch_blk<MG_CHEB, false>(L, blk_slices(L), sm_count, cps, st, din, rin, dout, c1, c2, nullptr, nullptr, nullptr, nullptr);
      return;
    }
    const unsigned g = mg_grid(slice_threads(L), sm_count, cps);
    if (dot) k_mg_spmv<MG_CHEB, true><<<g, PCG_BLOCK, 0, st>>>(L, din, rin, dout, c1, c2, L.b, dot_out, partial, counter);
    else k_mg_spmv<MG_CHEB, false><<<g, PCG_BLOCK, 0, st>>>(L, din, rin, dout, c1, c2, nullptr, nullptr, nullptr, nullptr);
  };
  // ---- down ----
  k_mg_smooth0<<<mg_grid(rows(J.lv[0]), sm_count, cps), PCG_BLOCK, 0, st>>>(J.lv[0]);
  for (int l = 0; l + 1 < nl; l++) {
    const MgLevelDev& L = J.lv[l];
    cheb(L, L.d0, L.b, L.d1, L.c1, L.c2, false);  // x = d0 + d1, r = b - A d0
    if (L.bval)
      launch_blk<MG_RESID, false>(L, blk_slices(L), sm_count, cps, st, L.d1, L.r, nullptr, 0.0, 0.0, nullptr, nullptr, nullptr, nullptr);
    else
      k_mg_spmv<MG_RESID, false><<<mg_grid(slice_threads(L), sm_count, cps), PCG_BLOCK, 0, st>>>(
          L, L.d1, L.r, nullptr, 0.0, 0.0, nullptr, nullptr, nullptr, nullptr);  // r = b - A x
    const MgLevelDev& C = J.lv[l + 1];
    const bool coarsest = l + 2 == nl;
    k_mg_restrict<<<mg_grid(rows(C), sm_count, cps), PCG_BLOCK, 0, st>>>(L, C, !(coarsest && J.dense));
  }
  // ---- coarsest ----
  {
    const MgLevelDev& C = J.lv[nl - 1];
    if (J.dense) {
      const int64_t nc = rows(C);
      k_mg_dense_apply<<<mg_grid(nc * 32, sm_count, cps), PCG_BLOCK, 0, st>>>(C, J.cinv);
      if (nl == 1)  // the whole structure is one dense block: z = A^-1 r, and r.z is all that is left to do
        k_mg_dot<<<mg_grid(nc, sm_count, cps), PCG_BLOCK, 0, st>>>(C, dot_out, partial, counter);
    } else {
      cheb(C, C.d0, C.b, C.d1, C.c1, C.c2, nl == 1);  // smoother only
    }
  }
  // ---- up ----
  for (int l = nl - 2; l >= 0; l--) {
    const MgLevelDev& L = J.lv[l];
    k_mg_prolong<<<mg_grid(rows(L), sm_count, cps), PCG_BLOCK, 0, st>>>(L, J.lv[l + 1].x, J.omega);
    cheb(L, L.d0, L.r, L.d1, 0.0, L.c0, false);         // r -= A e;  d = c0 D^-1 r;  x += d
    cheb(L, L.d1, L.r, L.d0, L.c1, L.c2, l == 0);       // second Chebyshev step (+ r.z on the fine level)
  }
  return cudaGetLastError();
}

}  // namespace obgpu
