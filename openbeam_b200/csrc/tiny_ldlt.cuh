// tiny_ldlt.cuh — fully unrolled in-register LDL^T factor + solve of one small dense SPD system (thread-local).
// Shared by k_tiny_fused (kernels_tiny.cu) and the run-time specialised k_tiny_jit (tiny_jit_kernel.cuh).
#pragma once

namespace obgpu {

template <int NF>
__device__ __forceinline__ int ldlt_solve_regs(double (&A)[NF + 1][NF + 1], double (&b)[NF + 1]) {
  // A: lower triangle A[i][j], j <= i.  In place: L (unit lower) below the diagonal, 1/D on it.
  // Rows below the current column stay unnormalised (V = L D) until their own column is reached, so the only
  // divisions are the NF reciprocals of the pivots (an FP64 division is ~15 instructions).
  int info = 0;
#pragma unroll
  for (int j = 0; j < NF; j++) {
    double w[NF + 1];
    double dj = A[j][j];
#pragma unroll
    for (int k = 0; k < NF; k++)
      if (k < j) {
        w[k] = A[j][k] * A[k][k];  // L(j,k) = V(j,k) / D(k)
        dj -= A[j][k] * w[k];
      }
    if (!(dj > 0.0) && info == 0) info = j + 1;  // Eigen LLT fails on x <= 0 (solver_static.cpp:97-105)
    A[j][j] = 1.0 / dj;
#pragma unroll
    for (int i = 0; i < NF; i++) {
      if (i > j) {
        double v = A[i][j];
#pragma unroll
        for (int k = 0; k < NF; k++)
          if (k < j) v -= A[i][k] * w[k];
        A[i][j] = v;
      }
    }
#pragma unroll
    for (int k = 0; k < NF; k++)
      if (k < j) A[j][k] = w[k];
  }
#pragma unroll
  for (int i = 0; i < NF; i++) {  // L y = b
#pragma unroll
    for (int k = 0; k < NF; k++)
      if (k < i) b[i] -= A[i][k] * b[k];
  }
#pragma unroll
  for (int i = 0; i < NF; i++) b[i] *= A[i][i];
#pragma unroll
  for (int ii = 0; ii < NF; ii++) {  // L^T x = z
    const int i = NF - 1 - ii;
#pragma unroll
    for (int k = 0; k < NF; k++)
      if (k > i) b[i] -= A[k][i] * b[k];
  }
  return info;
}

}  // namespace obgpu
