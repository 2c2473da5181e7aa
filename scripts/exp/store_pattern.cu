// Store-pattern micro-benchmark for the thread-per-case kernel's result records (scratch experiment, not shipped):
// 1M cases x (E=3 elements x 12 doubles stress + 6 U_f + 3 F_b + 1 int info), persistent grid of 128-thread CTAs.
//   mode 0: current pattern  — per element 12 STG.64 per warp, lanes cover 96-B pieces of 3 records (stride 288 B)
//   mode 1: contiguous STG.64 — the warp's 32 stress records (9216 B) written as 36 fully contiguous 256-B stores
//   mode 2: contiguous STG.128 — 18 stores of 512 B
//   mode 3: mode 2 + U_f/F_b with 128-bit stores where aligned
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
constexpr int E = 3, NF = 6, NB = 3;
template <int MODE>
__global__ void __launch_bounds__(128) k(double* stress, double* uf, double* fb, int* info, long long n, double v) {
  const int lane = threadIdx.x & 31, wl0 = threadIdx.x & ~31;
  const long long stride = (long long)gridDim.x * 128;
  for (long long base = (long long)blockIdx.x * 128; base < n; base += stride) {
    const long long c0 = base + wl0;
    if (MODE == 0) {
      for (int e = 0; e < E; e++) {
        double* gs = stress + (c0 * E + e) * 12;
#pragma unroll
        for (int r = 0; r < 12; r++) {
          const int j = r * 32 + lane, t = j / 12, i = j - t * 12;
          gs[t * E * 12 + i] = v + r;
        }
      }
    } else if (MODE == 1) {
      double* gs = stress + c0 * E * 12;
#pragma unroll
      for (int r = 0; r < 36; r++) gs[r * 32 + lane] = v + r;
    } else {
      double2* gs = reinterpret_cast<double2*>(stress + c0 * E * 12);
#pragma unroll
      for (int r = 0; r < 18; r++) gs[r * 32 + lane] = make_double2(v + r, v);
    }
    if (MODE == 3) {
      double2* g2 = reinterpret_cast<double2*>(uf + c0 * NF);
#pragma unroll
      for (int r = 0; r < 3; r++) g2[r * 32 + lane] = make_double2(v, v + r);
      double2* g3 = reinterpret_cast<double2*>(fb + c0 * NB);
      g3[lane] = make_double2(v, v);
      if (lane < 16) g3[32 + lane] = make_double2(v, v);
    } else {
#pragma unroll
      for (int r = 0; r < NF; r++) uf[c0 * NF + r * 32 + lane] = v + r;
#pragma unroll
      for (int r = 0; r < NB; r++) fb[c0 * NB + r * 32 + lane] = v + r;
    }
    info[base + threadIdx.x] = 0;
  }
}
int main(int argc, char** argv) {
  const long long n = 1 << 20;
  double *s, *u, *f; int* info;
  cudaMalloc(&s, n * E * 12 * 8); cudaMalloc(&u, n * NF * 8); cudaMalloc(&f, n * NB * 8); cudaMalloc(&info, n * 4);
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  const double bytes = double(n) * (E * 12 * 8 + NF * 8 + NB * 8 + 4);
  for (int per_sm = 2; per_sm <= 8; per_sm += 3)
    for (int mode = 0; mode < 4; mode++) {
      const int grid = 148 * per_sm;
      float best = 1e9f, tot = 0;
      for (int it = 0; it < 13; it++) {
        cudaEventRecord(a);
        if (mode == 0) k<0><<<grid, 128>>>(s, u, f, info, n, 1.0 + it);
        if (mode == 1) k<1><<<grid, 128>>>(s, u, f, info, n, 1.0 + it);
        if (mode == 2) k<2><<<grid, 128>>>(s, u, f, info, n, 1.0 + it);
        if (mode == 3) k<3><<<grid, 128>>>(s, u, f, info, n, 1.0 + it);
        cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b);
        if (it >= 3) { tot += ms; if (ms < best) best = ms; }
      }
      printf("per_sm %d mode %d: avg %.1f us  best %.1f us  %.0f GB/s\n", per_sm, mode, tot * 100, best * 1e3, bytes / (tot / 10 * 1e-3) / 1e9);
    }
  printf("err %s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
