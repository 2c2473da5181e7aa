// pcg_common.cuh — block reduction and in-kernel dot-product completion shared by the PCG and multigrid kernels.
#pragma once
#include "kernels.h"

namespace obgpu {

constexpr int PCG_BLOCK = 256;

__device__ __forceinline__ double block_sum(double v, double* sh) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  __syncthreads();
  if (l == 0) sh[w] = v;
  __syncthreads();
  double t = 0.0;
  if (w == 0) {
    t = l < PCG_BLOCK / 32 ? sh[l] : 0.0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
  }
  return t;  // valid in thread 0
}

// Finish NV dots: every block deposits its partials; the last block to arrive sums them in index order.
template <int NV>
__device__ __forceinline__ void finish_dots(const double (&v)[NV], double* partial, unsigned int* counter, double* sh,
                                            double* (&out)[NV]) {
  __shared__ bool last;
  double t[NV];
#pragma unroll
  for (int k = 0; k < NV; k++) t[k] = block_sum(v[k], sh);
  if (threadIdx.x == 0) {
#pragma unroll
    for (int k = 0; k < NV; k++) partial[k * gridDim.x + blockIdx.x] = t[k];
    __threadfence();
    const unsigned int ticket = atomicInc(counter, gridDim.x - 1);  // wraps to 0: self-resetting
    last = ticket == gridDim.x - 1;
  }
  __syncthreads();
  if (last) {
    __threadfence();
#pragma unroll
    for (int k = 0; k < NV; k++) {
      double acc = 0.0;
      for (unsigned i = threadIdx.x; i < gridDim.x; i += PCG_BLOCK) acc += __ldcg(partial + k * gridDim.x + i);
      acc = block_sum(acc, sh);
      if (threadIdx.x == 0) *out[k] = acc;
    }
  }
}

// Grid of a warp-per-slice kernel: at most `resident` CTAs per SM (what the kernel's registers allow, so that no second
// partial wave forms), capped by the caller's request, and shrunk so that every warp serves the same number of slices
// (a 3x3 block-SELL level has a third of the scalar slices: 2.2 per warp left some warps with 3 and others with 2)
template <class K>
static unsigned balanced_grid(K kernel, int64_t slices, int sm_count, int request) {
  static int resident = 0;   // one instance per kernel type
  if (resident == 0) {
    int r = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&r, kernel, PCG_BLOCK, 0) != cudaSuccess || r < 1) r = 4;
    resident = r;
  }
  const int per_sm = request > 0 && request < resident ? request : resident;
  const int64_t warps_cap = int64_t(sm_count) * per_sm * (PCG_BLOCK / 32);
  const int64_t k = (slices + warps_cap - 1) / warps_cap;                  // slices per warp
  const int64_t g = (slices + k * (PCG_BLOCK / 32) - 1) / (k * (PCG_BLOCK / 32));
  return unsigned(g < 1 ? 1 : g);
}

// y(3I..3I+2) = sum over the blocks of block row I (lane = block row of the slice), 3x3 block-SELL (BlockSell): every
// load of the warp reads 32 consecutive words; 76 bytes per block instead of 108 for nine scalar entries
__device__ __forceinline__ void block_row_product(const int32_t* __restrict__ bslptr, const int32_t* __restrict__ bcol,
                                                  const double* __restrict__ bval, int64_t sl, int lane,
                                                  const double* __restrict__ x, double (&s)[3]) {
  const int32_t base = bslptr[sl], w = (bslptr[sl + 1] - base) >> 5;
  const int32_t* __restrict__ pc = bcol + base + lane;
  const double* __restrict__ pv = bval + size_t(9) * base + lane;
  s[0] = s[1] = s[2] = 0.0;
  int k = 0;
  for (; k + 2 <= w; k += 2) {   // two blocks per trip: 2 + 18 coalesced loads in flight ahead of the 6 gathers
    const int32_t c0 = pc[32 * k], c1 = pc[32 * (k + 1)];
    double m[18];
#pragma unroll
    for (int v = 0; v < 18; v++) m[v] = pv[(9 * k + v) * 32];
    const double x0 = x[3 * c0], x1 = x[3 * c0 + 1], x2 = x[3 * c0 + 2];
    const double y0 = x[3 * c1], y1 = x[3 * c1 + 1], y2 = x[3 * c1 + 2];
#pragma unroll
    for (int a = 0; a < 3; a++)
      s[a] += m[3 * a] * x0 + m[3 * a + 1] * x1 + m[3 * a + 2] * x2 + m[9 + 3 * a] * y0 + m[9 + 3 * a + 1] * y1 + m[9 + 3 * a + 2] * y2;
  }
  for (; k < w; k++) {
    const int32_t c0 = pc[32 * k];
    double m[9];
#pragma unroll
    for (int v = 0; v < 9; v++) m[v] = pv[(9 * k + v) * 32];
    const double x0 = x[3 * c0], x1 = x[3 * c0 + 1], x2 = x[3 * c0 + 2];
#pragma unroll
    for (int a = 0; a < 3; a++) s[a] += m[3 * a] * x0 + m[3 * a + 1] * x1 + m[3 * a + 2] * x2;
  }
}

// The same on the FP32 copy the multigrid preconditioner streams: the nine values of a block as five float2 (the tenth
// float is padding), block column B (32 consecutive blocks of a slice) at float2 index 5 * 32 * B: 8-byte loads of 32
// consecutive lanes like the FP64 layout, five per block instead of nine, 44 bytes per block instead of 76.  Products and
// sums are FP64.
constexpr int BLK32_PAIRS = 5;
__device__ __forceinline__ void block_row_product_f32(const int32_t* __restrict__ bslptr, const int32_t* __restrict__ bcol,
                                                      const float* __restrict__ bval32, int64_t sl, int lane,
                                                      const double* __restrict__ x, double (&s)[3]) {
  const int32_t base = bslptr[sl], w = (bslptr[sl + 1] - base) >> 5;
  const int32_t* __restrict__ pc = bcol + base + lane;
  const float2* __restrict__ pv = reinterpret_cast<const float2*>(bval32) + size_t(BLK32_PAIRS) * base + lane;
  s[0] = s[1] = s[2] = 0.0;
  int k = 0;
  for (; k + 2 <= w; k += 2) {
    const int32_t c0 = pc[32 * k], c1 = pc[32 * (k + 1)];
    float2 m[2 * BLK32_PAIRS];
#pragma unroll
    for (int v = 0; v < 2 * BLK32_PAIRS; v++) m[v] = pv[(BLK32_PAIRS * k + v) * 32];
    const double x0 = x[3 * c0], x1 = x[3 * c0 + 1], x2 = x[3 * c0 + 2];
    const double y0 = x[3 * c1], y1 = x[3 * c1 + 1], y2 = x[3 * c1 + 2];
    s[0] += double(m[0].x) * x0 + double(m[0].y) * x1 + double(m[1].x) * x2 + double(m[5].x) * y0 + double(m[5].y) * y1 + double(m[6].x) * y2;
    s[1] += double(m[1].y) * x0 + double(m[2].x) * x1 + double(m[2].y) * x2 + double(m[6].y) * y0 + double(m[7].x) * y1 + double(m[7].y) * y2;
    s[2] += double(m[3].x) * x0 + double(m[3].y) * x1 + double(m[4].x) * x2 + double(m[8].x) * y0 + double(m[8].y) * y1 + double(m[9].x) * y2;
  }
  for (; k < w; k++) {
    const int32_t c0 = pc[32 * k];
    float2 m[BLK32_PAIRS];
#pragma unroll
    for (int v = 0; v < BLK32_PAIRS; v++) m[v] = pv[(BLK32_PAIRS * k + v) * 32];
    const double x0 = x[3 * c0], x1 = x[3 * c0 + 1], x2 = x[3 * c0 + 2];
    s[0] += double(m[0].x) * x0 + double(m[0].y) * x1 + double(m[1].x) * x2;
    s[1] += double(m[1].y) * x0 + double(m[2].x) * x1 + double(m[2].y) * x2;
    s[2] += double(m[3].x) * x0 + double(m[3].y) * x1 + double(m[4].x) * x2;
  }
}

// block_row_product leaves row 3 l + a of the slice in lane l; the row-local work that follows (vector reads and writes)
// wants lane l on rows l, 32 + l, 64 + l so that every access of the warp is 32 consecutive doubles: t[j] = the sum of
// slice row 32 j + l, fetched from lane (32 j + l) / 3, component (32 j + l) % 3
__device__ __forceinline__ void block_sums_to_rows(const double (&s)[3], int lane, double (&t)[3]) {
#pragma unroll
  for (int j = 0; j < 3; j++) {
    const int row = 32 * j + lane, src = row / 3, comp = row - 3 * src;
    const double v0 = __shfl_sync(0xffffffffu, s[0], src), v1 = __shfl_sync(0xffffffffu, s[1], src),
                 v2 = __shfl_sync(0xffffffffu, s[2], src);
    t[j] = comp == 0 ? v0 : comp == 1 ? v1 : v2;
  }
}

}  // namespace obgpu
