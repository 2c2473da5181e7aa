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

}  // namespace obgpu
