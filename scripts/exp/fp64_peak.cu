// fp64_peak.cu — measured DFMA peak of the device ("the builder must measure a DFMA peak ... before quoting any
// % of FP64 roofline").  16 independent FMA chains per thread, no memory traffic inside the timed loop.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o fp64_peak fp64_peak.cu && ./fp64_peak
#include <cstdio>
#include <cuda_runtime.h>

constexpr int CHAINS = 16, ITERS = 4096;

__global__ void __launch_bounds__(256) k_dfma(double* out, double a, double b) {
  double x[CHAINS];
#pragma unroll
  for (int i = 0; i < CHAINS; i++) x[i] = double(threadIdx.x + i) * 1e-3;
  for (int it = 0; it < ITERS; it++) {
#pragma unroll
    for (int i = 0; i < CHAINS; i++) x[i] = fma(x[i], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < CHAINS; i++) s += x[i];
  if (s == 123.456) out[0] = s;  // keeps the chains alive
}

int main() {
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, 0);
  double* out;
  cudaMalloc(&out, 8);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  double best = 0;
  for (int per_sm : {2, 4, 8}) {
    const int grid = p.multiProcessorCount * per_sm;
    k_dfma<<<grid, 256>>>(out, 0.999999, 1e-7);
    cudaDeviceSynchronize();
    for (int rep = 0; rep < 5; rep++) {
      cudaEventRecord(e0);
      k_dfma<<<grid, 256>>>(out, 0.999999, 1e-7);
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
      float ms;
      cudaEventElapsedTime(&ms, e0, e1);
      const double flops = 2.0 * double(grid) * 256 * CHAINS * ITERS;
      const double tf = flops / (ms * 1e-3) / 1e12;
      if (tf > best) best = tf;
    }
    printf("{\"ctas_per_sm\": %d, \"best_tflops_so_far\": %.3f}\n", per_sm, best);
  }
  printf("{\"device\": \"%s\", \"sms\": %d, \"fp64_dfma_peak_tflops\": %.3f, \"how\": \"16 independent DFMA chains/thread, 256 threads/CTA, best of 5\"}\n",
         p.name, p.multiProcessorCount, best);
  return 0;
}
