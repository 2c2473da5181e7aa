// Store-pattern micro-benchmark 2 (scratch experiment): what makes the multi-stream record writes of the thread-per-case
// kernel slower than a plain fill?  Variants: number of arrays, grid-stride vs blocked case partition, cache hints.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
constexpr int E = 3, NF = 6, NB = 6;
template <int HINT>
__device__ __forceinline__ void st2(double2* p, double2 v) {
  if (HINT == 0) *p = v;
  if (HINT == 1) __stcs(p, v);
  if (HINT == 2) __stwt(p, v);
  if (HINT == 3) __stcg(p, v);
}
// ARR: 1 = stress only (same total bytes by writing a wider record), 4 = stress, uf, ff+fb, info
template <int ARR, int BLOCKED, int HINT>
__global__ void __launch_bounds__(128) k(double* stress, double* uf, double* fb, int* info, long long n, double v) {
  const int lane = threadIdx.x & 31, wl0 = threadIdx.x & ~31;
  const long long nblk = (n + 127) / 128;
  long long b0, b1, bs;
  if (BLOCKED) { b0 = nblk * blockIdx.x / gridDim.x; b1 = nblk * (blockIdx.x + 1) / gridDim.x; bs = 1; }
  else { b0 = blockIdx.x; b1 = nblk; bs = gridDim.x; }
  for (long long blk = b0; blk < b1; blk += bs) {
    const long long c0 = blk * 128 + wl0;
    if (ARR == 1) {
      double2* gs = reinterpret_cast<double2*>(stress + c0 * 48);  // 384 B per case in one array
#pragma unroll
      for (int r = 0; r < 24; r++) st2<HINT>(gs + r * 32 + lane, make_double2(v + r, v));
    } else {
      double2* gs = reinterpret_cast<double2*>(stress + c0 * E * 12);
#pragma unroll
      for (int r = 0; r < 18; r++) st2<HINT>(gs + r * 32 + lane, make_double2(v + r, v));
      double2* g2 = reinterpret_cast<double2*>(uf + c0 * NF);
#pragma unroll
      for (int r = 0; r < 3; r++) st2<HINT>(g2 + r * 32 + lane, make_double2(v, v + r));
      double2* g3 = reinterpret_cast<double2*>(fb + c0 * NB);
#pragma unroll
      for (int r = 0; r < 3; r++) st2<HINT>(g3 + r * 32 + lane, make_double2(v, v + r));
      info[c0 + lane] = 0;
    }
  }
}
template <int ARR, int BLOCKED, int HINT>
void run(const char* name, int per_sm, double* s, double* u, double* f, int* info, long long n) {
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  const double bytes = ARR == 1 ? double(n) * 384 : double(n) * (E * 12 * 8 + NF * 8 + NB * 8 + 4);
  float tot = 0, best = 1e9f;
  for (int it = 0; it < 13; it++) {
    cudaEventRecord(a);
    k<ARR, BLOCKED, HINT><<<148 * per_sm, 128>>>(s, u, f, info, n, 1.0 + it);
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    if (it >= 3) { tot += ms; if (ms < best) best = ms; }
  }
  printf("%-34s per_sm %d: avg %.1f us best %.1f us  %.0f GB/s\n", name, per_sm, tot * 100, best * 1e3, bytes / (tot / 10 * 1e-3) / 1e9);
}
int main() {
  const long long n = 1000000 / 128 * 128;
  double *s, *u, *f; int* info;
  cudaMalloc(&s, n * 48 * 8); cudaMalloc(&u, n * NF * 8); cudaMalloc(&f, n * NB * 8); cudaMalloc(&info, n * 4);
  for (int per_sm = 3; per_sm <= 16; per_sm *= 2) {
    run<1, 0, 0>("1 array, grid-stride", per_sm, s, u, f, info, n);
    run<1, 1, 0>("1 array, blocked", per_sm, s, u, f, info, n);
    run<4, 0, 0>("4 arrays, grid-stride", per_sm, s, u, f, info, n);
    run<4, 1, 0>("4 arrays, blocked", per_sm, s, u, f, info, n);
    run<4, 0, 1>("4 arrays, grid-stride, st.cs", per_sm, s, u, f, info, n);
    run<4, 0, 2>("4 arrays, grid-stride, st.wt", per_sm, s, u, f, info, n);
    run<4, 0, 3>("4 arrays, grid-stride, st.cg", per_sm, s, u, f, info, n);
    run<1, 0, 1>("1 array, grid-stride, st.cs", per_sm, s, u, f, info, n);
  }
  cudaMemset(s, 0, n * 48 * 8);
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  for (int it = 0; it < 3; it++) {
    cudaEventRecord(a); cudaMemsetAsync(s, it, n * 48 * 8); cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    printf("cudaMemset 384 MB: %.1f us %.0f GB/s\n", ms * 1e3, n * 384.0 / ms / 1e6);
  }
  printf("err %s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
