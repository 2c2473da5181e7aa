// kernels_tiny.cu — K1+K2+K4+K5+K6+K10 fused for batches of load cases over ONE small shared topology
// (BASELINE config 2: demo3_frame x 1M cases).  One THREAD per system: at n_f <= 12 a CTA per
// system would leave >90% of the lanes idle, and the whole pipeline is HBM-bound (440 B/case for
// demo3_frame) — so each lane carries one case end to end:
//   element stiffness (closed form, beam_math.cuh) -> scatter into its private packed K_ff in shared
//   memory through the topology program (uniform indices: every lane of a warp follows the same
//   program, so smem accesses are conflict-free [idx][lane] columns) -> equivalent nodal loads ->
//   K_ff pulled into registers -> fully unrolled LDL^T + solves -> element end forces, reactions.
// Sums follow the reference's element order (assemble.cpp:123-223), no atomics anywhere.
// Replaces, per case, one full solveStatic()+postProcCalcStress() call of the reference.
#include "beam_math.cuh"
#include "kernels.h"
#include "tiny_program.h"

namespace obgpu {

template <int NF>
__device__ __forceinline__ int ldlt_solve_regs(double (&A)[NF * (NF + 1) / 2 + 1], double (&b)[NF + 1]) {
  // A: packed lower, A[i*(i+1)/2+j].  In place: L (unit lower) below the diagonal, D on it.
  int info = 0;
#pragma unroll
  for (int j = 0; j < NF; j++) {
    double dj = A[j * (j + 1) / 2 + j];
#pragma unroll
    for (int k = 0; k < j; k++) {
      const double l = A[j * (j + 1) / 2 + k];
      dj -= l * l * A[k * (k + 1) / 2 + k];
    }
    if (!(dj > 0.0) && info == 0) info = j + 1;  // Eigen LLT fails on x <= 0 (solver_static.cpp:97-105)
    A[j * (j + 1) / 2 + j] = dj;
    const double inv = 1.0 / dj;
#pragma unroll
    for (int i = j + 1; i < NF; i++) {
      double v = A[i * (i + 1) / 2 + j];
#pragma unroll
      for (int k = 0; k < j; k++) v -= A[i * (i + 1) / 2 + k] * (A[j * (j + 1) / 2 + k] * A[k * (k + 1) / 2 + k]);
      A[i * (i + 1) / 2 + j] = v * inv;
    }
  }
#pragma unroll
  for (int i = 0; i < NF; i++) {  // L y = b
#pragma unroll
    for (int k = 0; k < i; k++) b[i] -= A[i * (i + 1) / 2 + k] * b[k];
  }
#pragma unroll
  for (int i = 0; i < NF; i++) b[i] /= A[i * (i + 1) / 2 + i];
#pragma unroll
  for (int ii = 0; ii < NF; ii++) {  // L^T x = z  (ascending counter so the loop unrolls)
    const int i = NF - 1 - ii;
#pragma unroll
    for (int kk = 0; kk < ii; kk++) {
      const int k = NF - 1 - kk;
      b[i] -= A[k * (k + 1) / 2 + i] * b[k];
    }
  }
  return info;
}

template <int NF, int BLOCK>
__global__ void __launch_bounds__(BLOCK) k_tiny_fused(TinyArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  int32_t* PI = reinterpret_cast<int32_t*>(smem_raw);
  double* PD = reinterpret_cast<double*>(smem_raw + size_t(a.n_ints) * 4);
  double* scratch = PD + a.n_dbls;
  for (int i = threadIdx.x; i < a.n_ints; i += BLOCK) PI[i] = a.prog_i[i];
  for (int i = threadIdx.x; i < a.n_dbls; i += BLOCK) PD[i] = a.prog_d[i];
  __syncthreads();
  const int N = PI[0], E = PI[1], n = PI[2], nb = PI[4], NL = PI[5], EL = PI[6];
  constexpr int NFP = NF * (NF + 1) / 2;
  const TinyLayout lay(N, E, n, NF, nb, NL, EL);
  // private columns: Kp[NFP] | F[n] | U[n], element [idx] of lane t at scratch[idx*BLOCK + t]
  double* Kp = scratch + threadIdx.x;
  double* F = Kp + size_t(NFP) * BLOCK;
  double* U = F + size_t(n) * BLOCK;

  for (int64_t cs = int64_t(blockIdx.x) * BLOCK + threadIdx.x; cs < a.n_cases; cs += int64_t(gridDim.x) * BLOCK) {
#pragma unroll
    for (int i = 0; i < NFP; i++) Kp[i * BLOCK] = 0.0;
    for (int i = 0; i < n; i++) {
      F[i * BLOCK] = 0.0;
      U[i * BLOCK] = 0.0;
    }
    // ---- nodal loads (assemble.cpp:326-341, free DOFs only) ----
    for (int i = 0; i < NL; i++) {
      const int k = PI[lay.o_nldst + i];
      const double v = a.case_nl ? __ldg(a.case_nl + cs * NL + i) : PD[lay.o_nlval + i];
      if (k >= 0) F[k * BLOCK] += v;
    }
    // ---- stage A: element stiffness + scatter, equivalent loads ----
    for (int e = 0; e < E; e++) {
      const int na = PI[lay.o_enode + 2 * e], nbn = PI[lay.o_enode + 2 * e + 1], type = PI[lay.o_etype + e];
      const ElemGeom g = elem_geom(PD[lay.o_xy + 2 * na], PD[lay.o_xy + 2 * na + 1], PD[lay.o_xy + 2 * nbn],
                                   PD[lay.o_xy + 2 * nbn + 1]);
      double Ee, Ae, Ie;
      if (a.case_eai) {
        const double* p = a.case_eai + (cs * E + e) * 3;
        Ee = __ldg(p); Ae = __ldg(p + 1); Ie = __ldg(p + 2);
      } else {
        Ee = PD[lay.o_eai + 3 * e]; Ae = PD[lay.o_eai + 3 * e + 1]; Ie = PD[lay.o_eai + 3 * e + 2];
      }
      const ElemLocalK lk = elem_local_k(type, Ee, Ae, Ie, g);
      if (NF > 0) {
        double ke[ELEM_K];
        elem_global_k(lk, g, ke);
        const int32_t* dst = PI + lay.o_edst + ELEM_K * e;
#pragma unroll
        for (int k = 0; k < ELEM_K; k++) {
          const int d = dst[k];
          if (d >= 0) Kp[d * BLOCK] += ke[k];
        }
      }
      for (int j = PI[lay.o_elptr + e]; j < PI[lay.o_elptr + e + 1]; j++) {
        const int slot = PI[lay.o_elslot + j];
        const double q = a.case_q ? __ldg(a.case_q + cs * EL + slot) : PD[lay.o_elq + slot];
        double gl[6], ex[6];
        uniform_load(type, g, q, PD[lay.o_eldir + 3 * slot], PD[lay.o_eldir + 3 * slot + 1], gl, ex);
        const int32_t* ud = PI + lay.o_udof + 6 * e;
#pragma unroll
        for (int k = 0; k < 6; k++) {
          const int d = ud[k];
          if (d >= 0) F[d * BLOCK] += gl[k];
        }
      }
    }
    // ---- K5/K6: gather K_ff, F_f into registers; LDL^T; solve ----
    double A[NFP + 1], b[NF + 1];
#pragma unroll
    for (int i = 0; i < NFP; i++) A[i] = Kp[i * BLOCK];
#pragma unroll
    for (int i = 0; i < NF; i++) b[i] = F[PI[lay.o_freek + i] * BLOCK];
    if (a.F_f) {
#pragma unroll
      for (int i = 0; i < NF; i++) a.F_f[cs * NF + i] = b[i];
    }
    const int info = ldlt_solve_regs<NF>(A, b);
    a.info[cs] = info;
    if (info != 0) atomicOr(a.fail, 1);
#pragma unroll
    for (int i = 0; i < NF; i++) {
      U[PI[lay.o_freek + i] * BLOCK] = b[i];
      a.U_f[cs * NF + i] = b[i];
    }
    // reactions accumulate in F: bounded entries currently hold the equivalent loads (solver_static.cpp:208-218)
    for (int i = 0; i < nb; i++) {
      const int k = PI[lay.o_bndk + i];
      F[k * BLOCK] = -F[k * BLOCK];
    }
    // ---- stage B: element end forces (K10) and reactions (K_bf*U_f == sum of element end forces) ----
    for (int e = 0; e < E; e++) {
      const int na = PI[lay.o_enode + 2 * e], nbn = PI[lay.o_enode + 2 * e + 1], type = PI[lay.o_etype + e];
      const ElemGeom g = elem_geom(PD[lay.o_xy + 2 * na], PD[lay.o_xy + 2 * na + 1], PD[lay.o_xy + 2 * nbn],
                                   PD[lay.o_xy + 2 * nbn + 1]);
      double Ee, Ae, Ie;
      if (a.case_eai) {
        const double* p = a.case_eai + (cs * E + e) * 3;
        Ee = __ldg(p); Ae = __ldg(p + 1); Ie = __ldg(p + 2);
      } else {
        Ee = PD[lay.o_eai + 3 * e]; Ae = PD[lay.o_eai + 3 * e + 1]; Ie = PD[lay.o_eai + 3 * e + 2];
      }
      const ElemLocalK lk = elem_local_k(type, Ee, Ae, Ie, g);
      const int32_t* ud = PI + lay.o_udof + 6 * e;
      double ug[6], es[6];
#pragma unroll
      for (int k = 0; k < 6; k++) ug[k] = ud[k] >= 0 ? U[ud[k] * BLOCK] : 0.0;
      elem_end_forces(lk, g, ug, es);
      const double gf[6] = {g.c * es[0] - g.s * es[1], g.s * es[0] + g.c * es[1], es[2],
                            g.c * es[3] - g.s * es[4], g.s * es[3] + g.c * es[4], es[5]};
#pragma unroll
      for (int k = 0; k < 6; k++)
        if (ud[k] >= 0) F[ud[k] * BLOCK] += gf[k];
      if (a.stress) {
        double ext[6] = {0, 0, 0, 0, 0, 0};
        for (int j = PI[lay.o_elptr + e]; j < PI[lay.o_elptr + e + 1]; j++) {
          const int slot = PI[lay.o_elslot + j];
          const double q = a.case_q ? __ldg(a.case_q + cs * EL + slot) : PD[lay.o_elq + slot];
          double gl[6], ex[6];
          uniform_load(type, g, q, PD[lay.o_eldir + 3 * slot], PD[lay.o_eldir + 3 * slot + 1], gl, ex);
#pragma unroll
          for (int k = 0; k < 6; k++) ext[k] += ex[k];
        }
        // StressInfo::element_stress[e][face] = {N,Vy,Vz,Mx,My,Mz}; face 0 negated (postproc_stress.cpp:152-165)
        double2* o = reinterpret_cast<double2*>(a.stress + (cs * E + e) * 12);
        o[0] = make_double2(-es[0] + ext[0], -es[1] + ext[1]);
        o[1] = make_double2(0.0, 0.0);
        o[2] = make_double2(0.0, -es[2] + ext[2]);
        o[3] = make_double2(es[3] + ext[3], es[4] + ext[4]);
        o[4] = make_double2(0.0, 0.0);
        o[5] = make_double2(0.0, es[5] + ext[5]);
      }
    }
    for (int i = 0; i < nb; i++) a.F_b[cs * nb + i] = F[PI[lay.o_bndk + i] * BLOCK];
  }
}

constexpr int TINY_BLOCK = 128;

template <int NF>
static cudaError_t launch_nf(const TinyArgs& a, int n, int sm_count, cudaStream_t st) {
  constexpr int NFP = NF * (NF + 1) / 2;
  const size_t smem = size_t(a.n_ints) * 4 + size_t(a.n_dbls) * 8 + size_t(NFP + 2 * n) * TINY_BLOCK * 8;
  auto kern = k_tiny_fused<NF, TINY_BLOCK>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
  if (e != cudaSuccess) return e;
  int per_sm = 1;
  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, TINY_BLOCK, smem);
  if (e != cudaSuccess) return e;
  if (per_sm < 1) per_sm = 1;
  const int64_t need = (a.n_cases + TINY_BLOCK - 1) / TINY_BLOCK;
  // persistent-ish grid: a whole number of waves over the SMs, grid-stride over cases
  int64_t grid = int64_t(sm_count) * per_sm;
  if (need < grid) grid = need;
  if (grid < 1) grid = 1;
  kern<<<unsigned(grid), TINY_BLOCK, smem, st>>>(a);
  return cudaGetLastError();
}

cudaError_t launch_tiny_fused(int nf, int n, const TinyArgs& a, int sm_count, cudaStream_t st) {
  switch (nf) {
#define OB_CASE(K) case K: return launch_nf<K>(a, n, sm_count, st);
    OB_CASE(0) OB_CASE(1) OB_CASE(2) OB_CASE(3) OB_CASE(4) OB_CASE(5) OB_CASE(6)
    OB_CASE(7) OB_CASE(8) OB_CASE(9) OB_CASE(10) OB_CASE(11) OB_CASE(12)
#undef OB_CASE
  }
  return cudaErrorInvalidValue;
}

}  // namespace obgpu
