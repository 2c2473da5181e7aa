// kernels_tiny.cu — K1+K2+K4+K5+K6+K10 fused for batches of load cases over ONE small shared topology
// (BASELINE config 2: demo3_frame x 1M cases).  One THREAD per system: at n_f <= 12 a CTA per system would
// leave >90% of the lanes idle, and the pipeline moves only 440 B per case — so each lane
// carries one case end to end:
//   per CTA, once:  case-invariant geometry (orientation, L, 1/L: internals.cpp:30-123) of the E elements and
//                   the projections of the element-load directions -> shared memory
//   per case:       local stiffness from (E,A,Iz) -> rotation to global (closed form, beam_math.cuh) -> scatter
//                   into the lane's private packed K_ff column in shared memory through the topology program
//                   -> equivalent nodal loads -> K_ff pulled into registers -> fully unrolled LDL^T + solves
//                   -> element end forces, reactions.
// The topology "program" (integer indices produced by the host symbolic phase) rides in the kernel parameter,
// i.e. the constant bank: every lane follows the same program, so index reads are uniform constant loads and
// the skip-tests are uniform branches; shared memory holds only the per-lane columns [idx][lane] (conflict-free).
// Sums follow the reference's element order (assemble.cpp:123-223); no atomics on the data path.
#include <cstdlib>
#include <cstring>

#include "beam_math.cuh"
#include "kernels.h"
#include "tiny_ldlt.cuh"
#include "tiny_program.h"

namespace obgpu {

#ifndef TINY_BLOCK_SIZE
#define TINY_BLOCK_SIZE 128
#endif
constexpr int TINY_BLOCK = TINY_BLOCK_SIZE;

// fixed-capacity copy of the program, passed by value (constant bank)
struct TinyParam {
  int N, E, n, nb, NL, EL;
  int8_t enode[2 * TINY_MAX_E];
  int8_t etype[TINY_MAX_E];
  int8_t udof[6 * TINY_MAX_E];       // -1 absent, k free, k|64 bounded
  int8_t edst[ELEM_K * TINY_MAX_E];  // packed K_ff index or -1
  int8_t elptr[TINY_MAX_E + 1];
  int8_t elslot[TINY_MAX_EL];
  int8_t nldst[TINY_MAX_NL];
  int8_t freek[TINY_MAX_NF];
  int8_t bndk[TINY_MAX_N];
  int8_t eblk[2 * TINY_MAX_E];       // node-blocked path: free node block of each element end, -1 = no free DOF
  double xy[2 * TINY_MAX_NODES];
  double eai[3 * TINY_MAX_E];
  double nlval[TINY_MAX_NL];
  double elq[TINY_MAX_EL];
  double eldir[2 * TINY_MAX_EL];
};

// ---- node-blocked fast path -------------------------------------------------------------------------------------
// When every element end is either entirely free with its three DOFs numbered consecutively from a multiple of 3
// (free "node block" m: free indices 3m..3m+2) or has no free DOF at all (clamped supports: every frame of the
// BASELINE configs), K_ff is a grid of 3x3 node blocks and the scatter of an element's K11/K22/K12 is a warp-uniform
// choice of block position.  K_ff then lives in REGISTERS from its first touch to the end of the factorisation and
// the displacement gather of stress recovery is a register select: no shared-memory round trips (the LSU data pipe
// was the bound of the column-in-shared-memory version, profiles/README.md r01c).  Same adds in the same order.
template <int NF>
__device__ __forceinline__ void blk_add_diag(double (&A)[NF + 1][NF + 1], int m, const double* k6) {
#pragma unroll
  for (int I = 0; I < NF / 3; I++)
    if (m == I) {  // k6 = upper triangle (0,0),(0,1),(0,2),(1,1),(1,2),(2,2) of a symmetric block
      A[3 * I][3 * I] += k6[0];
      A[3 * I + 1][3 * I] += k6[1];
      A[3 * I + 1][3 * I + 1] += k6[3];
      A[3 * I + 2][3 * I] += k6[2];
      A[3 * I + 2][3 * I + 1] += k6[4];
      A[3 * I + 2][3 * I + 2] += k6[5];
    }
}
// k9 = K12 row-major (rows: end 0, columns: end 1); ma/mb = free node blocks of end 0 / end 1 (both >= 0, distinct)
template <int NF>
__device__ __forceinline__ void blk_add_off(double (&A)[NF + 1][NF + 1], int ma, int mb, const double* k9) {
#pragma unroll
  for (int I = 1; I < NF / 3; I++)
#pragma unroll
    for (int J = 0; J < I; J++) {
      if (ma == I && mb == J) {
#pragma unroll
        for (int r = 0; r < 3; r++)
#pragma unroll
          for (int c = 0; c < 3; c++) A[3 * I + r][3 * J + c] += k9[3 * r + c];
      } else if (mb == I && ma == J) {
#pragma unroll
        for (int r = 0; r < 3; r++)
#pragma unroll
          for (int c = 0; c < 3; c++) A[3 * I + r][3 * J + c] += k9[3 * c + r];
      }
    }
}
template <int NF>
__device__ __forceinline__ void blk_get3(const double (&b)[NF + 1], int m, double* u) {
  u[0] = u[1] = u[2] = 0.0;  // no free DOF at this end: U_b = 0
#pragma unroll
  for (int I = 0; I < NF / 3; I++)
    if (m == I) { u[0] = b[3 * I]; u[1] = b[3 * I + 1]; u[2] = b[3 * I + 2]; }
}

__device__ __forceinline__ void cp_async8(double* smem_dst, const double* gsrc) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(uint32_t(__cvta_generic_to_shared(smem_dst))), "l"(gsrc));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// Warp-cooperative store of a [32 cases][C] record block whose values sit in staging columns W[i*CS + t]
// (component i, case-in-warp t): 32 consecutive doubles per instruction.  Element j = r*32 + lane of the block is
// component i = j % C of case t = j / C; both repeat with period PER = C/gcd(C,32) in r (t advancing by
// TSTEP = 32/gcd(C,32)), so the PER lane-dependent staging offsets are computed ONCE per thread, outside the case
// loop, and every other round is an immediate offset away.  (The divisions cost 26 instructions per store in v4:
// 29 % of all instructions issued, profiles/README.md r01d.)
__host__ __device__ constexpr int gcd32(int c) { return (c % 32 == 0) ? 32 : (c % 16 == 0) ? 16 : (c % 8 == 0) ? 8 : (c % 4 == 0) ? 4 : (c % 2 == 0) ? 2 : 1; }
template <int C, int CS>
struct StoreMap {
  static constexpr int G = gcd32(C), PER = C / G, TSTEP = 32 / G;
  int so[PER];
  __device__ __forceinline__ void init(int lane) {
#pragma unroll
    for (int r0 = 0; r0 < PER; r0++) {
      const int j = r0 * 32 + lane, t = j / C, i = j - t * C;
      so[r0] = i * CS + t;
    }
  }
  // contiguous destination g[0 .. cnt*C)
  __device__ __forceinline__ void store(double* __restrict__ g, const double* W, int lane, int cnt) const {
    if (cnt == 32) {
#pragma unroll
      for (int r = 0; r < C; r++) g[r * 32 + lane] = W[so[r % PER] + (r / PER) * TSTEP];
    } else {
      for (int j = lane; j < cnt * C; j += 32) {
        const int t = j / C, i = j - t * C;
        g[j] = W[i * CS + t];
      }
    }
  }
};

template <int NF, int MINB, bool BLK>
__global__ void __launch_bounds__(TINY_BLOCK, MINB) k_tiny_fused(const __grid_constant__ TinyParam P, TinyArgs a) {
  constexpr int BLOCK = TINY_BLOCK;
  constexpr int CS = TINY_BLOCK + 1;  // column stride (+1: conflict-light transposed reads in the coalesced stores)
  constexpr int NFP = NF * (NF + 1) / 2;
  extern __shared__ __align__(16) double smem[];
  // shared: G[E][4] = c,s,L,1/L | LP[EL][2] | per-lane columns Kp[NFP][BLOCK], FU[n][BLOCK], SN[NL][BLOCK], SQ[2][EL][BLOCK]
  const int E = P.E, n = P.n, nb = P.nb, NL = P.NL, EL = P.EL;
  double* G = smem;
  double* LP = G + 4 * E;
  double* Kp = LP + 2 * EL + threadIdx.x;
  // Staging columns of the coalesced result stores: SS = the 12-wide element_stress record, whose Vz,Mx,My
  // components (2,3,4,8,9,10: zero in 2D) are plain zero columns; SX = U_f / F_f / F_b records, XW wide.  The
  // node-blocked path gives SS its own columns (zeroed once per CTA); the generic path overlays both on its K columns.
  constexpr int XW = BLK ? (NF > 6 ? NF : 6) : (NFP > 12 ? NFP : 12);
  constexpr int KC = BLK ? 12 + XW : XW;
  double* SX = BLK ? Kp + 12 * CS : Kp;
  double* FU = Kp + size_t(KC) * CS;
  const int lane = threadIdx.x & 31, wl0 = threadIdx.x & ~31;
  const double* SSW = Kp - threadIdx.x + wl0;  // warp views of the staging columns: [col*CS + case-in-warp]
  const double* SXW = SX - threadIdx.x + wl0;
  StoreMap<12, CS> map12;
  StoreMap<(NF ? NF : 1), CS> mapNF;
  map12.init(lane);
  mapNF.init(lane);
  int go12[3];  // global offsets of the stress record rounds: case t, component i -> t*E*12 + i
#pragma unroll
  for (int r0 = 0; r0 < 3; r0++) {
    const int j = r0 * 32 + lane, t = j / 12, i = j - t * 12;
    go12[r0] = t * P.E * 12 + i;
  }
  if (BLK) {
    Kp[2 * CS] = 0.0; Kp[3 * CS] = 0.0; Kp[4 * CS] = 0.0;
    Kp[8 * CS] = 0.0; Kp[9 * CS] = 0.0; Kp[10 * CS] = 0.0;
  }
  if (threadIdx.x < E) {
    const int e = threadIdx.x, na = P.enode[2 * e], nbn = P.enode[2 * e + 1];
    const ElemGeom g = elem_geom(P.xy[2 * na], P.xy[2 * na + 1], P.xy[2 * nbn], P.xy[2 * nbn + 1]);
    G[4 * e] = g.c; G[4 * e + 1] = g.s; G[4 * e + 2] = g.L; G[4 * e + 3] = g.invL;
    for (int j = P.elptr[e]; j < P.elptr[e + 1]; j++) {
      const int slot = P.elslot[j];
      const LoadProj pr = load_proj(g, P.eldir[2 * slot], P.eldir[2 * slot + 1]);
      LP[2 * slot] = pr.plong; LP[2 * slot + 1] = pr.pperp;
    }
  }
  __syncthreads();

  // Per-case inputs arrive by cp.async into the lane's staging columns one block of cases ahead of their use:
  // SN[NL] (nodal-load values, consumed at the top of a case) and SQ[2][EL] (element-load q, double-buffered
  // because it is still read at the end of the case), so global-load latency hides behind a whole case of compute.
  double* SN = FU + size_t(n) * CS;
  double* SQ = SN + (a.case_nl ? size_t(NL) * CS : 0);
  const int64_t stride = int64_t(gridDim.x) * BLOCK;
  const int64_t last = a.n_cases - 1;
  auto prefetch = [&](int64_t b, int buf) {
    const int64_t c = min(b + threadIdx.x, last);
    if (a.case_nl)
      for (int i = 0; i < NL; i++) cp_async8(SN + i * CS, a.case_nl + c * NL + i);
    if (a.case_q)
      for (int i = 0; i < EL; i++) cp_async8(SQ + (buf * EL + i) * CS, a.case_q + c * EL + i);
    cp_async_commit();
  };
  int buf = 0;
  int64_t base = int64_t(blockIdx.x) * BLOCK;
  if (base < a.n_cases) prefetch(base, 0);
  // CTA-uniform loop: lanes past the end recompute the last case and skip their stores, so the whole body is
  // convergent (uniform program reads, uniform branches).
  for (; base < a.n_cases; base += stride, buf ^= 1) {
    const bool valid = base + threadIdx.x <= last;
    const int64_t cs = min(base + threadIdx.x, last);
    cp_async_wait_all();
    double A[NF + 1][NF + 1], b[NF + 1];
    if (BLK) {
#pragma unroll
      for (int i = 0; i < NF; i++)
#pragma unroll
        for (int j = 0; j < NF; j++)
          if (j <= i) A[i][j] = 0.0;
    } else {
#pragma unroll
      for (int i = 0; i < NFP; i++) Kp[i * CS] = 0.0;
    }
    for (int i = 0; i < n; i++) FU[i * CS] = 0.0;
    // ---- nodal loads (assemble.cpp:326-341, free DOFs only) ----
    for (int i = 0; i < NL; i++) {
      const int k = P.nldst[i];
      if (k >= 0) FU[k * CS] += a.case_nl ? SN[i * CS] : P.nlval[i];
    }
    if (base + stride < a.n_cases) prefetch(base + stride, buf ^ 1);
    const double* Q = SQ + size_t(buf) * EL * CS;
    // ---- stage A: element stiffness + scatter, equivalent loads ----
    for (int e = 0; e < E; e++) {
      const int type = P.etype[e];
      ElemGeom g;
      g.c = G[4 * e]; g.s = G[4 * e + 1]; g.L = G[4 * e + 2]; g.invL = G[4 * e + 3];
      double Ee, Ae, Ie;
      if (a.case_eai) {
        const double* p = a.case_eai + (cs * E + e) * 3;
        Ee = __ldg(p); Ae = __ldg(p + 1); Ie = __ldg(p + 2);
      } else {
        Ee = P.eai[3 * e]; Ae = P.eai[3 * e + 1]; Ie = P.eai[3 * e + 2];
      }
      const ElemLocalK lk = elem_local_k(type, Ee, Ae, Ie, g);
      if (NF > 0) {
        double ke[ELEM_K];
        elem_global_k(lk, g, ke);
        if (BLK) {
          const int ma = P.eblk[2 * e], mb = P.eblk[2 * e + 1];
          blk_add_diag<NF>(A, ma, ke);
          blk_add_diag<NF>(A, mb, ke + 6);
          blk_add_off<NF>(A, ma, mb, ke + 12);
        } else {
#pragma unroll
          for (int k = 0; k < ELEM_K; k++) {
            const int d = P.edst[ELEM_K * e + k];
            if (d >= 0) Kp[d * CS] += ke[k];
          }
        }
      }
      for (int j = P.elptr[e]; j < P.elptr[e + 1]; j++) {
        const int slot = P.elslot[j];
        const double q = a.case_q ? Q[slot * CS] : P.elq[slot];
        LoadProj pr;
        pr.plong = LP[2 * slot]; pr.pperp = LP[2 * slot + 1];
        double gl[6], ex[6];
        uniform_load_p(type, g, q, pr, gl, ex);
#pragma unroll
        for (int k = 0; k < 6; k++) {
          const int d = P.udof[6 * e + k];
          if (d >= 0) FU[(d & 63) * CS] += gl[k];
        }
      }
    }
    // ---- K5/K6: gather K_ff, F_f into registers; LDL^T; solve ----
    if (!BLK) {
#pragma unroll
      for (int i = 0; i < NF; i++)
#pragma unroll
        for (int j = 0; j < NF; j++)
          if (j <= i) A[i][j] = Kp[(i * (i + 1) / 2 + j) * CS];
    }
#pragma unroll
    for (int i = 0; i < NF; i++) b[i] = FU[P.freek[i] * CS];
    // Results leave through the warp's (now dead) K columns: each lane parks its values in its column, then the
    // warp writes the 32 cases' contiguous output range with fully coalesced stores (a per-lane store of a
    // 48..288-byte record would touch 32 different 128-byte lines per instruction and bind the kernel on L1 tags).
    const int64_t c0 = base + wl0;                                   // first case of this warp
    const int cnt = int(min(int64_t(32), a.n_cases - c0));           // valid cases in this warp (<= 0: none)
    if (a.F_f) {
      __syncwarp();
#pragma unroll
      for (int i = 0; i < NF; i++) SX[i * CS] = b[i];
      __syncwarp();
      if (NF > 0) mapNF.store(a.F_f + c0 * NF, SXW, lane, cnt);
      __syncwarp();
    }
    const int info = ldlt_solve_regs<NF>(A, b);
    if (valid) {
      a.info[cs] = info;
      if (info != 0) atomicOr(a.fail, 1);
    }
#pragma unroll
    for (int i = 0; i < NF; i++) {
      if (!BLK) FU[P.freek[i] * CS] = b[i];
      SX[i * CS] = b[i];
    }
    __syncwarp();
    if (NF > 0) mapNF.store(a.U_f + c0 * NF, SXW, lane, cnt);
    if (!BLK) {  // the K columns become the stress record: its Vz,Mx,My columns must read zero
      __syncwarp();
      Kp[2 * CS] = 0.0; Kp[3 * CS] = 0.0; Kp[4 * CS] = 0.0;
      Kp[8 * CS] = 0.0; Kp[9 * CS] = 0.0; Kp[10 * CS] = 0.0;
    }
    // bounded entries hold the equivalent loads: reactions start from their negative (solver_static.cpp:208-218)
    for (int i = 0; i < nb; i++) {
      const int k = P.bndk[i];
      FU[k * CS] = -FU[k * CS];
    }
    // ---- stage B: element end forces (K10) and reactions (K_bf*U_f == sum of element end forces) ----
    for (int e = 0; e < E; e++) {
      const int type = P.etype[e];
      ElemGeom g;
      g.c = G[4 * e]; g.s = G[4 * e + 1]; g.L = G[4 * e + 2]; g.invL = G[4 * e + 3];
      double Ee, Ae, Ie;
      if (a.case_eai) {
        const double* p = a.case_eai + (cs * E + e) * 3;
        Ee = __ldg(p); Ae = __ldg(p + 1); Ie = __ldg(p + 2);
      } else {
        Ee = P.eai[3 * e]; Ae = P.eai[3 * e + 1]; Ie = P.eai[3 * e + 2];
      }
      const ElemLocalK lk = elem_local_k(type, Ee, Ae, Ie, g);
      double ug[6], es[6];
      if (BLK) {
        blk_get3<NF>(b, P.eblk[2 * e], ug);
        blk_get3<NF>(b, P.eblk[2 * e + 1], ug + 3);
      } else {
#pragma unroll
        for (int k = 0; k < 6; k++) {
          const int d = P.udof[6 * e + k];
          ug[k] = (d >= 0 && d < 64) ? FU[d * CS] : 0.0;   // bounded DOFs: U_b = 0
        }
      }
      elem_end_forces(lk, g, ug, es);
      const double gf[6] = {g.c * es[0] - g.s * es[1], g.s * es[0] + g.c * es[1], es[2],
                            g.c * es[3] - g.s * es[4], g.s * es[3] + g.c * es[4], es[5]};
#pragma unroll
      for (int k = 0; k < 6; k++) {
        const int d = P.udof[6 * e + k];
        if (d >= 64) FU[(d & 63) * CS] += gf[k];
      }
      if (a.stress) {
        double ext[6] = {0, 0, 0, 0, 0, 0};
        for (int j = P.elptr[e]; j < P.elptr[e + 1]; j++) {
          const int slot = P.elslot[j];
          const double q = a.case_q ? Q[slot * CS] : P.elq[slot];
          LoadProj pr;
          pr.plong = LP[2 * slot]; pr.pperp = LP[2 * slot + 1];
          double gl[6], ex[6];
          uniform_load_p(type, g, q, pr, gl, ex);
#pragma unroll
          for (int k = 0; k < 6; k++) ext[k] += ex[k];
        }
        // StressInfo::element_stress[e][face] = {N,Vy,Vz,Mx,My,Mz}; face 0 negated (postproc_stress.cpp:152-165)
        __syncwarp();  // previous element's record has been read out
        Kp[0 * CS] = -es[0] + ext[0]; Kp[1 * CS] = -es[1] + ext[1]; Kp[5 * CS] = -es[2] + ext[2];
        Kp[6 * CS] = es[3] + ext[3];  Kp[7 * CS] = es[4] + ext[4];  Kp[11 * CS] = es[5] + ext[5];
        __syncwarp();
        double* gs = a.stress + (c0 * E + e) * 12;
        if (cnt == 32) {
#pragma unroll
          for (int r = 0; r < 12; r++)  // round r = r0 + 3k: case t0 + 8k, same component (96 = 8 * 12)
            gs[go12[r % 3] + (r / 3) * 96 * E] = SSW[map12.so[r % 3] + (r / 3) * 8];
        } else {
          for (int j = lane; j < cnt * 12; j += 32) {
            const int t = j / 12, i = j - t * 12;
            gs[t * E * 12 + i] = SSW[i * CS + t];
          }
        }
      }
    }
    __syncwarp();
    if (NF > 0 && nb == NF) {  // same record width as U_f: reuse its hoisted store map
#pragma unroll
      for (int i = 0; i < NF; i++) SX[i * CS] = FU[P.bndk[i] * CS];
      __syncwarp();
      mapNF.store(a.F_b + c0 * NF, SXW, lane, cnt);
      __syncwarp();
    } else {
      for (int i0 = 0; i0 < nb; i0 += XW) {  // reactions leave through the staging columns too, XW components at a time
        const int w = min(XW, nb - i0);
        for (int i = 0; i < w; i++) SX[i * CS] = FU[P.bndk[i0 + i] * CS];
        __syncwarp();
        const float rw = 1.0f / float(w);
        for (int j = lane; j < cnt * w; j += 32) {
          const int t = int((float(j) + 0.5f) * rw), i = j - t * w;  // exact for j < 2^20
          a.F_b[(c0 + t) * nb + i0 + i] = SXW[i * CS + t];
        }
        __syncwarp();  // columns are rewritten at the top of the next block of cases
      }
    }
  }
}

static bool fill_param(TinyParam& P, const int32_t* I, const double* D) {
  const int N = I[0], E = I[1], n = I[2], nf = I[3], nb = I[4], NL = I[5], EL = I[6];
  const TinyLayout lay(N, E, n, nf, nb, NL, EL);
  P.N = N; P.E = E; P.n = n; P.nb = nb; P.NL = NL; P.EL = EL;
  for (int i = 0; i < 2 * E; i++) P.enode[i] = int8_t(I[lay.o_enode + i]);
  for (int i = 0; i < E; i++) P.etype[i] = int8_t(I[lay.o_etype + i]);
  for (int i = 0; i < 6 * E; i++) P.udof[i] = int8_t(I[lay.o_udof + i]);
  for (int i = 0; i < ELEM_K * E; i++) P.edst[i] = int8_t(I[lay.o_edst + i]);
  for (int i = 0; i <= E; i++) P.elptr[i] = int8_t(I[lay.o_elptr + i]);
  for (int i = 0; i < EL; i++) P.elslot[i] = int8_t(I[lay.o_elslot + i]);
  for (int i = 0; i < NL; i++) P.nldst[i] = int8_t(I[lay.o_nldst + i]);
  for (int i = 0; i < nf; i++) P.freek[i] = int8_t(I[lay.o_freek + i]);
  for (int i = 0; i < nb; i++) P.bndk[i] = int8_t(I[lay.o_bndk + i]);
  for (int i = 0; i < 2 * N; i++) P.xy[i] = D[lay.o_xy + i];
  for (int i = 0; i < 3 * E; i++) P.eai[i] = D[lay.o_eai + i];
  for (int i = 0; i < NL; i++) P.nlval[i] = D[lay.o_nlval + i];
  for (int i = 0; i < EL; i++) {
    P.elq[i] = D[lay.o_elq + i];
    P.eldir[2 * i] = D[lay.o_eldir + 3 * i];
    P.eldir[2 * i + 1] = D[lay.o_eldir + 3 * i + 1];
  }
  // node-blocked? every element end: three free DOFs at free indices 3m,3m+1,3m+2, or no free DOF at all
  int finv[64];
  for (int i = 0; i < 64; i++) finv[i] = -1;
  for (int i = 0; i < nf; i++) finv[P.freek[i]] = i;
  bool blocked = nf % 3 == 0 && nf > 0;
  for (int e = 0; e < E && blocked; e++)
    for (int end = 0; end < 2; end++) {
      int f[3];
      for (int k = 0; k < 3; k++) {
        const int d = P.udof[6 * e + 3 * end + k];
        f[k] = (d >= 0 && d < 64) ? finv[d] : -1;
      }
      if (f[0] < 0 && f[1] < 0 && f[2] < 0) P.eblk[2 * e + end] = -1;
      else if (f[0] >= 0 && f[0] % 3 == 0 && f[1] == f[0] + 1 && f[2] == f[0] + 2) P.eblk[2 * e + end] = int8_t(f[0] / 3);
      else blocked = false;
    }
  for (int e = 0; e < E && blocked; e++)  // an element between two ends of the same node block would alias K11/K22/K12
    if (P.eblk[2 * e] >= 0 && P.eblk[2 * e] == P.eblk[2 * e + 1]) blocked = false;
  return blocked;
}

template <int NF, bool BLK>
static cudaError_t launch_nf(const TinyParam& P, const TinyArgs& a, int n, int sm_count, cudaStream_t st) {
  constexpr int NFP = NF * (NF + 1) / 2;
#ifndef TINY_MINB_SMALL
#define TINY_MINB_SMALL 6
#endif
#ifndef TINY_MINB_BLK
#define TINY_MINB_BLK 5
#endif
  // the node-blocked path keeps K_ff, F_f and the element blocks live together: measured (v5, 1M cases) 115 us at
  // 5 CTAs/SM (96 registers), 123 us at 4 (118 registers, no spills), 140 us at 6 (80 registers, spills)
  constexpr int MINB = NF <= 6 ? (BLK ? TINY_MINB_BLK : TINY_MINB_SMALL) : NF <= 9 ? 3 : 2;  // register budget: K_ff (NFP doubles) lives in registers
  const size_t smem = size_t(4 * P.E + 2 * P.EL) * 8 +
                      size_t((BLK ? 12 + (NF > 6 ? NF : 6) : (NFP > 12 ? NFP : 12)) + n + (a.case_nl ? P.NL : 0) + (a.case_q ? 2 * P.EL : 0)) * (TINY_BLOCK + 1) * 8;
  auto kern = k_tiny_fused<NF, MINB, BLK>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);  // per function, not per launch: other host threads launch it with other sizes
  if (e != cudaSuccess) return e;
  int per_sm = 1;
  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, TINY_BLOCK, smem);
  if (e != cudaSuccess) return e;
  if (per_sm < 1) per_sm = 1;
  const int64_t need = (a.n_cases + TINY_BLOCK - 1) / TINY_BLOCK;
  // persistent-ish grid: a whole number of CTAs per SM, grid-stride over cases
  int64_t grid = int64_t(sm_count) * per_sm;
  if (need < grid) grid = need;
  if (grid < 1) grid = 1;
  kern<<<unsigned(grid), TINY_BLOCK, smem, st>>>(P, a);
  return cudaGetLastError();
}

static cudaError_t launch_tiny_precompiled(int nf, int n, const TinyArgs& a, int sm_count, cudaStream_t st);

cudaError_t launch_tiny_fused(int nf, int n, const TinyArgs& a, int sm_count, cudaStream_t st, const char** which) {
  // first choice: the kernel specialised to this topology at run time (tiny_jit.cu); the precompiled interpreter
  // below serves when NVRTC or the driver library is not available
  {
    int dev = 0;
    cudaGetDevice(&dev);
    const cudaError_t je = launch_tiny_jit(dev, a, sm_count, st, which);
    if (je != cudaErrorNotSupported) return je;
  }
  if (a.stress_compact) return cudaErrorNotYetImplemented;  // the precompiled kernels write the full record
  {  // the precompiled interpreter knows uniform element loads only: the caller takes the warp-per-structure route
    const int32_t* I = a.prog_i;
    const TinyLayout lay(I[0], I[1], I[2], I[3], I[4], I[5], I[6]);
    for (int i = 0; i < I[6]; i++)
      if (I[lay.o_elkind + i] != OBGPU_LOAD_UNIFORM) return cudaErrorNotSupported;
  }
  if (which) *which = "k_tiny_fused (precompiled, thread per case, 128 threads/CTA)";
  if (a.summ && a.summ_fallback) {  // the precompiled kernels do not write summaries: separate pass over U_f / F_b
    const cudaError_t e = launch_tiny_precompiled(nf, n, a, sm_count, st);
    if (e != cudaSuccess) return e;
    const cudaError_t e2 = launch_summaries(*a.summ_fallback, sm_count, st);
    if (e2 != cudaSuccess) return e2;
    for (int p = 0; p < a.n_summ_peer; p++) {  // peer copies: plain device-to-device copies behind the kernel
      const cudaError_t e3 = cudaMemcpyAsync(a.summ_peer[p], a.summ, size_t(a.n_cases) * 32, cudaMemcpyDeviceToDevice, st);
      if (e3 != cudaSuccess) return e3;
    }
    return cudaSuccess;
  }
  return launch_tiny_precompiled(nf, n, a, sm_count, st);
}

cudaError_t launch_tiny_precompiled(int nf, int n, const TinyArgs& a, int sm_count, cudaStream_t st) {
  TinyParam P;
  std::memset(&P, 0, sizeof(P));
  const bool blocked = fill_param(P, a.prog_i, a.prog_d) && !getenv("OBGPU_TINY_GENERIC");
  if (blocked) switch (nf) {
#define OB_CASE(K) case K: return launch_nf<K, true>(P, a, n, sm_count, st);
    OB_CASE(3) OB_CASE(6) OB_CASE(9) OB_CASE(12)
#undef OB_CASE
  }
  switch (nf) {
#define OB_CASE(K) case K: return launch_nf<K, false>(P, a, n, sm_count, st);
    OB_CASE(0) OB_CASE(1) OB_CASE(2) OB_CASE(3) OB_CASE(4) OB_CASE(5) OB_CASE(6)
    OB_CASE(7) OB_CASE(8) OB_CASE(9) OB_CASE(10) OB_CASE(11) OB_CASE(12)
#undef OB_CASE
  }
  return cudaErrorInvalidValue;
}

}  // namespace obgpu
