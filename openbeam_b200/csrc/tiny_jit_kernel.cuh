// tiny_jit_kernel.cuh — the thread-per-case kernel SPECIALISED TO ONE TOPOLOGY (compiled at run time by NVRTC, see
// tiny_jit.cpp).  Same pipeline and same arithmetic as k_tiny_fused (kernels_tiny.cu): per case the element
// stiffnesses are formed, rotated, assembled, factored (LDL^T) and solved, then end forces and reactions — nothing is
// hoisted out of the case loop.  What changes is that the integer topology program (DOF maps, element->K_ff slots,
// load slots) is a set of compile-time constants (the generated prelude `TJ_*`), so
//   * every loop over elements / loads / DOFs unrolls and every index is an immediate,
//   * K_ff, F and U are plain register arrays from first touch to last use: no shared-memory read-modify-write, no
//     dispatch on block positions, no per-entry skip tests (entries that leave K_ff simply generate no code),
//   * it serves ANY topology within the tiny limits, not only the node-blocked ones.
// Shared memory only stages the per-case inputs (cp.async, one block of cases ahead) and the coalesced result stores.
// The floating-point inputs (coordinates, sections, base load values) stay run-time kernel parameters.
//
// The prelude must define:  TJ_N, TJ_E, TJ_n, TJ_NF, TJ_NB, TJ_NL, TJ_EL, TJ_MINB and the constexpr arrays
//   TJ_enode[2E], TJ_etype[E], TJ_udof[6E] (-1 absent, k free, k|64 bounded), TJ_edst[21E] (packed K_ff index or -1),
//   TJ_elkind[EL] (OBGPU_LOAD_* of every element load),
//   TJ_SUMM + TJ_bnddof[NB] (per-case summaries written by this kernel; DoF_index of each bounded DOF),
//   TJ_BULK (records through the bulk-copy engine), TJ_FF (F_f is kept), TJ_EVICT_FIRST,
//   TJ_SC (compact stress records [E][2][3] = N, Vy, Mz per face instead of the reference's [E][2][6]; TJ_BULK only),
//   TJ_elptr[E+1], TJ_elslot[EL], TJ_nldst[NL], TJ_freek[NF], TJ_bndk[NB]
// (arrays padded to at least one entry), then include this file after beam_math.cuh.
#pragma once

namespace obgpu {

struct TinyJitVals {  // run-time values of the shared topology (kernel parameter -> constant bank)
  double xy[2 * TJ_N];
  double eai[3 * TJ_E];
  double nlval[TJ_NL > 0 ? TJ_NL : 1];
  double elq[TJ_EL > 0 ? TJ_EL : 1];
  double eldir[TJ_EL > 0 ? 2 * TJ_EL : 1];
  double elq2[TJ_EL > 0 ? TJ_EL : 1];  // second load parameter (q_end / dist); fixed over the cases
};

__host__ __device__ constexpr int tj_row(int d) {
  int i = 0;
  while ((i + 1) * (i + 2) / 2 <= d) i++;
  return i;
}
__host__ __device__ constexpr int tj_gcd32(int c) {
  return (c % 32 == 0) ? 32 : (c % 16 == 0) ? 16 : (c % 8 == 0) ? 8 : (c % 4 == 0) ? 4 : (c % 2 == 0) ? 2 : 1;
}

__device__ __forceinline__ void tj_cp_async8(double* smem_dst, const double* gsrc) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(unsigned(__cvta_generic_to_shared(smem_dst))), "l"(gsrc));
}

// Staging columns of a [cases][C] record: column i starts at i*TJ_RS + sigma_C(i) doubles, one entry per thread after
// that.  The thread-per-case writes are trivially conflict-free; the transposed reads of the coalesced store (lane ->
// flat index j = C*t + i of the 32-case block) touch bank pair (sigma_C(i) + t) mod 16, and sigma is chosen so that this
// is a BIJECTION on every aligned run of 16 flat indices (one 64-bit half-warp access):
//   C = 2^a q (q odd), i = 2^a u + v:  sigma = (16 / 2^a) v + q^{-1} u  (mod 16)   [2^a >= 16: sigma = i]
// since with J = q t + u (j = 2^a J + v) the bank is (16/2^a) v + q^{-1} J, distinct over 16/2^a consecutive J and all v.
__host__ __device__ constexpr int tj_sigma(int C, int i) {
  int p2 = 1, q = C;
  while (q % 2 == 0 && p2 < 16) { q /= 2; p2 *= 2; }
  if (p2 >= 16) return i % 16;
  int qi = 1;
  while ((qi * (q % 16)) % 16 != 1) qi += 2;
  return ((16 / p2) * (i % p2) + qi * (i / p2)) % 16;
}
constexpr int TJ_RS = 128 + 16;  // column pitch of the record staging (room for the bank offset)
__host__ __device__ constexpr int tj_col(int C, int i) { return i * TJ_RS + tj_sigma(C, i); }

// Coalesced store of a [32 cases][C] record block staged in tj_col columns (see StoreMap in kernels_tiny.cu).
template <int C>
struct TjStoreMap {
  static constexpr int G = tj_gcd32(C), PER = C / G, TSTEP = 32 / G;
  int so[PER];
  __device__ __forceinline__ void init(int lane) {
#pragma unroll
    for (int r0 = 0; r0 < PER; r0++) {
      const int j = r0 * 32 + lane, t = j / C, i = j - t * C;
      so[r0] = tj_col(C, i) + t;
    }
  }
  __device__ __forceinline__ void store(double* __restrict__ g, const double* W, int lane, int cnt) const {
    if (cnt == 32) {
#pragma unroll
      for (int r = 0; r < C; r++) g[r * 32 + lane] = W[so[r % PER] + (r / PER) * TSTEP];
    } else {
      for (int j = lane; j < cnt * C; j += 32) {
        const int t = j / C, i = j - t * C;
        g[j] = W[tj_col(C, i) + t];
      }
    }
  }
};

// ---- TJ_BULK: result records leave through the TMA (cp.async.bulk shared -> global) ----
// Each warp keeps its 32 cases' records in shared memory in the FINAL global layout ([case][record], every thread writes
// its own row), then one lane hands the whole contiguous block to the bulk-copy engine and the warp goes on computing:
// the stores no longer sit in the warps' instruction streams, so a saturated memory system cannot stall the arithmetic
// of the other warps behind a queue of STGs (measured: compute 45 us + stores 65 us ran back to back as 98 us).
// The staging rows are reused one iteration later, after cp.async.bulk.wait_group.read.
__device__ __forceinline__ void tj_fence_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
// waits until at most N of the most recently committed bulk groups of the issuing lane still have their source rows
// unread: the records leave in a fixed order per iteration (F_f | info, U_f | stress, F_b, summaries), so each staging
// array is waited for only where it is written again, one iteration later, with the younger groups left in flight
template <int N>
__device__ __forceinline__ void tj_bulk_wait_read(int lane) {
  if (lane == 0) asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
  __syncwarp();
}
// All lanes call this after their row writes + tj_fence_async() + __syncwarp().  rec_bytes is a multiple of 4.
__device__ __forceinline__ void tj_bulk_store(void* g, const void* s, int rec_bytes, int cnt, int lane) {
  if (cnt <= 0) return;
  const int bytes = cnt * rec_bytes;
  if (((unsigned long long)(g) & 15ULL) == 0 && (bytes & 15) == 0) {
    if (lane == 0) {
#if TJ_EVICT_FIRST  // results are written once and not read again by this kernel: do not let them displace the inputs in L2
      unsigned long long pol;
      asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group.L2::cache_hint [%0], [%1], %2, %3;" ::"l"(g),
                   "r"(unsigned(__cvta_generic_to_shared(s))), "r"(bytes), "l"(pol)
                   : "memory");
#else
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(g),
                   "r"(unsigned(__cvta_generic_to_shared(s))), "r"(bytes)
                   : "memory");
#endif
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
  } else {  // ragged tail / unaligned destination: plain stores
    for (int j = lane; j < bytes / 4; j += 32) static_cast<int*>(g)[j] = static_cast<const int*>(s)[j];
  }
}

constexpr int TJ_BLOCK = TJ_BLOCK_SIZE;  // threads per CTA (128 unless OBGPU_TINY_JIT_BLOCK says otherwise)
constexpr int TJ_CS = TJ_BLOCK;  // pitch of the per-case input columns (read back by the owning thread only)
constexpr int TJ_XW = TJ_NF > TJ_NB ? (TJ_NF > 0 ? TJ_NF : 1) : (TJ_NB > 0 ? TJ_NB : 1);  // widest U_f / F_f / F_b record

extern "C" __global__ void __launch_bounds__(TJ_BLOCK, TJ_MINB) k_tiny_jit(const __grid_constant__ TinyJitVals V, TinyArgs a) {
  constexpr int E = TJ_E, n = TJ_n, NF = TJ_NF, NB = TJ_NB, NL = TJ_NL, EL = TJ_EL, CS = TJ_CS;
  extern __shared__ __align__(16) double smem[];
  // shared: G[E][4] = c,s,L,1/L | LP[EL][2] | staging columns SS[12] (stress record, zero columns 2,3,4,8,9,10) |
  // SX[XW] (U_f/F_f/F_b records) | SN[NL] | SQ[2][EL]   (record columns: tj_col; input columns: [CS] wide)
  double* G = smem;
  double* LP = G + 4 * E;
  const int lane = threadIdx.x & 31, wl0 = threadIdx.x & ~31;
#if TJ_BULK
  // per warp: RS[32][12E] stress | RU[32][NF] U_f | RF[32][NF] F_f | RB[32][NB] F_b | RM[32][4] summaries | RI[32] info
  // (final layouts)
  constexpr int SR = TJ_SC ? 6 : 12;  // doubles per element of the stress record
  constexpr int WR = 32 * (SR * E + (TJ_FF ? 2 : 1) * NF + NB + (TJ_SUMM ? 4 : 0)) + 16;  // doubles per warp
  double* RS = LP + 2 * (EL > 0 ? EL : 1) + (threadIdx.x >> 5) * WR;
  double* RU = RS + 32 * SR * E;
  double* RF = RU + 32 * NF;
  double* RB = RF + (TJ_FF ? 32 * NF : 0);
  double* RM = RB + 32 * NB;  // [32][4] summaries
  int* RI = reinterpret_cast<int*>(RM + (TJ_SUMM ? 32 * 4 : 0));
  double* SN = LP + 2 * (EL > 0 ? EL : 1) + (TJ_BLOCK / 32) * WR + threadIdx.x;
  double* SQ = SN + (a.case_nl ? NL * CS : 0);
  double* rs = RS + lane * SR * E;
#pragma unroll
  for (int i = 0; i < SR * E; i++) rs[i] = 0.0;  // the structural zeros of the full stress records stay in place
  // bulk groups a warp commits per iteration, in order: [F_f] | info, U_f | [stress], [F_b], [summaries (+ peers)]
  const bool late_wait0 = TJ_SUMM && a.n_summ_peer > 0;  // peer copies add groups: fall back to waiting for everything
#else
  double* SS = LP + 2 * (EL > 0 ? EL : 1) + threadIdx.x;
  double* SX = SS + 12 * TJ_RS;
  double* SN = SX + TJ_XW * TJ_RS;
  double* SQ = SN + (a.case_nl ? NL * CS : 0);
#endif
#ifdef TJ_ABL_NOSTORE  // ablation experiment (OBGPU_TINY_JIT_ABL=1): everything computed, no global result stores
  const bool st_on = a.fail == nullptr;
#else
  constexpr bool st_on = true;
#endif
#if !TJ_BULK
  const double* SSW = SS - threadIdx.x + wl0;
  const double* SXW = SX - threadIdx.x + wl0;
#endif
  if (threadIdx.x < E) {
    const int e = threadIdx.x, na = TJ_enode[2 * e], nbn = TJ_enode[2 * e + 1];
    const ElemGeom g = elem_geom(V.xy[2 * na], V.xy[2 * na + 1], V.xy[2 * nbn], V.xy[2 * nbn + 1]);
    G[4 * e] = g.c; G[4 * e + 1] = g.s; G[4 * e + 2] = g.L; G[4 * e + 3] = g.invL;
    for (int j = TJ_elptr[e]; j < TJ_elptr[e + 1]; j++) {
      const int slot = TJ_elslot[j];
      const LoadProj pr = load_proj(g, V.eldir[2 * slot], V.eldir[2 * slot + 1]);
      LP[2 * slot] = pr.plong; LP[2 * slot + 1] = pr.pperp;
    }
  }
#if !TJ_BULK
  constexpr int NFC = NF > 0 ? NF : 1, NBC = NB > 0 ? NB : 1;
  SS[tj_col(12, 2)] = 0.0; SS[tj_col(12, 3)] = 0.0; SS[tj_col(12, 4)] = 0.0;
  SS[tj_col(12, 8)] = 0.0; SS[tj_col(12, 9)] = 0.0; SS[tj_col(12, 10)] = 0.0;
  TjStoreMap<12> map12;
  TjStoreMap<NFC> mapNF;
  TjStoreMap<NBC> mapNB;
  map12.init(lane);
  mapNF.init(lane);
  mapNB.init(lane);
  int go12[3];
#pragma unroll
  for (int r0 = 0; r0 < 3; r0++) {
    const int j = r0 * 32 + lane, t = j / 12, i = j - t * 12;
    go12[r0] = t * E * 12 + i;
  }
#endif
  __syncthreads();

  const long long stride = (long long)gridDim.x * TJ_BLOCK;
  const long long last = a.n_cases - 1;
  auto prefetch = [&](long long b, int buf) {
    const long long c = min(b + (long long)threadIdx.x, last);
    if (a.case_nl)
#pragma unroll
      for (int i = 0; i < NL; i++) tj_cp_async8(SN + i * CS, a.case_nl + c * NL + i);
    if (a.case_q)
#pragma unroll
      for (int i = 0; i < EL; i++) tj_cp_async8(SQ + (buf * EL + i) * CS, a.case_q + c * EL + i);
    asm volatile("cp.async.commit_group;");
  };
  int buf = 0;
  long long base = (long long)blockIdx.x * TJ_BLOCK;
  if (base < a.n_cases) prefetch(base, 0);
  for (; base < a.n_cases; base += stride, buf ^= 1) {
    const bool valid = base + threadIdx.x <= last;
    const long long cs = min(base + (long long)threadIdx.x, last);
    asm volatile("cp.async.wait_all;" ::: "memory");
    // ---- registers: lower triangle of K_ff, F/U over all problem DOFs ----
    double A[NF + 1][NF + 1], fu[n + 1], b[NF + 1];
#pragma unroll
    for (int i = 0; i < NF; i++)
#pragma unroll
      for (int j = 0; j <= i; j++) A[i][j] = 0.0;
#pragma unroll
    for (int i = 0; i < n; i++) fu[i] = 0.0;
    // nodal loads (assemble.cpp:326-341, free DOFs only)
#pragma unroll
    for (int i = 0; i < NL; i++)
      if (TJ_nldst[i] >= 0) fu[TJ_nldst[i] >= 0 ? TJ_nldst[i] : 0] += a.case_nl ? SN[i * CS] : V.nlval[i];
    if (base + stride < a.n_cases) prefetch(base + stride, buf ^ 1);
    if (a.case_eai && base + stride < a.n_cases) {  // next block's section rows: pull them into L2 ahead of the __ldg reads
      const char* p = reinterpret_cast<const char*>(a.case_eai + min(base + stride + (long long)threadIdx.x, last) * (3 * E));
#pragma unroll
      for (int o = 0; o < 3 * E * 8; o += 128) asm volatile("prefetch.global.L2 [%0];" ::"l"(p + o));
    }
    const double* Q = SQ + buf * EL * CS;
    // ---- stage A: element stiffness -> K_ff registers, equivalent loads -> F registers ----
#pragma unroll
    for (int e = 0; e < E; e++) {
      constexpr int dummy = 0;
      (void)dummy;
      const int type = TJ_etype[e];
      ElemGeom g;
      g.c = G[4 * e]; g.s = G[4 * e + 1]; g.L = G[4 * e + 2]; g.invL = G[4 * e + 3];
      double Ee, Ae, Ie;
      if (a.case_eai) {
        const double* p = a.case_eai + (cs * E + e) * 3;
        Ee = __ldg(p); Ae = __ldg(p + 1); Ie = __ldg(p + 2);
      } else {
        Ee = V.eai[3 * e]; Ae = V.eai[3 * e + 1]; Ie = V.eai[3 * e + 2];
      }
      const ElemLocalK lk = elem_local_k(type, Ee, Ae, Ie, g);
      double ke[ELEM_K];
      elem_global_k(lk, g, ke);
#pragma unroll
      for (int k = 0; k < ELEM_K; k++) {
        const int d = TJ_edst[ELEM_K * e + k];
        if (d >= 0) A[tj_row(d >= 0 ? d : 0)][(d >= 0 ? d : 0) - tj_row(d >= 0 ? d : 0) * (tj_row(d >= 0 ? d : 0) + 1) / 2] += ke[k];
      }
#pragma unroll
      for (int j = TJ_elptr[e]; j < TJ_elptr[e + 1]; j++) {
        const int slot = TJ_elslot[j];
        const double q = a.case_q ? Q[slot * CS] : V.elq[slot];
        LoadProj pr;
        pr.plong = LP[2 * slot]; pr.pperp = LP[2 * slot + 1];
        double gl[6], ex[6];
        elem_load_p(TJ_elkind[slot], type, g, q, V.elq2[slot], pr, Ee * Ae, gl, ex);
#pragma unroll
        for (int k = 0; k < 6; k++) {
          const int d = TJ_udof[6 * e + k];
          if (d >= 0) fu[(d >= 0 ? d : 0) & 63] += gl[k];
        }
      }
    }
#pragma unroll
    for (int i = 0; i < NF; i++) b[i] = fu[TJ_freek[i]];
    const long long c0 = base + wl0;
    const int cnt = int(min(32LL, a.n_cases - c0));
#if TJ_BULK
    // last iteration's F_f / info / U_f rows have been read out (its stress, F_b and summary groups, committed after
    // them, may still be in flight: their rows are not touched before stage B)
    constexpr int G_LATE = 1 + (NB > 0 ? 1 : 0) + (TJ_SUMM ? 1 : 0);
    if (late_wait0 || !a.stress) tj_bulk_wait_read<0>(lane);
    else tj_bulk_wait_read<G_LATE>(lane);
    if (TJ_FF && a.F_f && NF > 0) {
#pragma unroll
      for (int i = 0; i < NF; i++) RF[lane * NF + i] = b[i];
      tj_fence_async();
      __syncwarp();
      if (st_on) tj_bulk_store(a.F_f + c0 * NF, RF, NF * 8, cnt, lane);
    }
#else
    if (a.F_f) {
      __syncwarp();
#pragma unroll
      for (int i = 0; i < NF; i++) SX[tj_col(NFC, i)] = b[i];
      __syncwarp();
      if (NF > 0 && st_on) mapNF.store(a.F_f + c0 * NF, SXW, lane, cnt);
      __syncwarp();
    }
#endif
#ifdef TJ_ABL_NOSOLVE  // ablation experiment (OBGPU_TINY_JIT_ABL=2): K_ff assembled but not factored
    int info = 0;
#pragma unroll
    for (int i = 0; i < NF; i++)
#pragma unroll
      for (int j = 0; j <= i; j++) b[i] += A[i][j];
#else
    const int info = ldlt_solve_regs<NF>(A, b);
#endif
#if TJ_BULK
    if (valid && info != 0) atomicOr(a.fail, 1);
    RI[lane] = info;
#pragma unroll
    for (int i = 0; i < NF; i++) {
      fu[TJ_freek[i]] = b[i];
      if (NF % 2 == 0) {
        if (i % 2 == 1) reinterpret_cast<double2*>(RU + lane * NF)[i / 2] = make_double2(b[i - 1], b[i]);
      } else {
        RU[lane * NF + i] = b[i];
      }
    }
    tj_fence_async();
    __syncwarp();
    if (st_on) {
      tj_bulk_store(a.info + c0, RI, 4, cnt, lane);
      if (NF > 0) tj_bulk_store(a.U_f + c0 * NF, RU, NF * 8, cnt, lane);
    }
#else
    if (valid && st_on) {
      a.info[cs] = info;
      if (info != 0) atomicOr(a.fail, 1);
    }
#pragma unroll
    for (int i = 0; i < NF; i++) {
      fu[TJ_freek[i]] = b[i];
      SX[tj_col(NFC, i)] = b[i];
    }
    __syncwarp();
    if (NF > 0 && st_on) mapNF.store(a.U_f + c0 * NF, SXW, lane, cnt);
#endif
#if TJ_SUMM
    double umax = 0.0;  // k_summaries' max |U_f| (kernels_misc.cu), same order of operations
#pragma unroll
    for (int i = 0; i < NF; i++) umax = fmax(umax, fabs(b[i]));
#endif
#if TJ_BULK
    // stage B rewrites the stress / F_b / summary rows: only this iteration's info and U_f groups (and F_f) may be pending
    if (!late_wait0 && a.stress) tj_bulk_wait_read<2 + ((TJ_FF && NF > 0) ? 1 : 0)>(lane);
#endif
    // bounded entries hold the equivalent loads: reactions start from their negative (solver_static.cpp:208-218)
#pragma unroll
    for (int i = 0; i < NB; i++) fu[TJ_bndk[i]] = -fu[TJ_bndk[i]];
    // ---- stage B: element end forces (K10), reactions (K_bf U_f = sum of element end forces, U_b = 0) ----
#pragma unroll
    for (int e = 0; e < E; e++) {
      const int type = TJ_etype[e];
      ElemGeom g;
      g.c = G[4 * e]; g.s = G[4 * e + 1]; g.L = G[4 * e + 2]; g.invL = G[4 * e + 3];
      double Ee, Ae, Ie;
      if (a.case_eai) {
        const double* p = a.case_eai + (cs * E + e) * 3;
        Ee = __ldg(p); Ae = __ldg(p + 1); Ie = __ldg(p + 2);
      } else {
        Ee = V.eai[3 * e]; Ae = V.eai[3 * e + 1]; Ie = V.eai[3 * e + 2];
      }
      const ElemLocalK lk = elem_local_k(type, Ee, Ae, Ie, g);
      double ug[6], es[6];
#pragma unroll
      for (int k = 0; k < 6; k++) {
        const int d = TJ_udof[6 * e + k];
        ug[k] = (d >= 0 && d < 64) ? fu[(d >= 0 && d < 64) ? d : 0] : 0.0;
      }
      elem_end_forces(lk, g, ug, es);
      const double gf[6] = {g.c * es[0] - g.s * es[1], g.s * es[0] + g.c * es[1], es[2],
                            g.c * es[3] - g.s * es[4], g.s * es[3] + g.c * es[4], es[5]};
#pragma unroll
      for (int k = 0; k < 6; k++) {
        const int d = TJ_udof[6 * e + k];
        if (d >= 64) fu[(d >= 64 ? d : 64) & 63] += gf[k];
      }
      if (a.stress) {
        double ext[6] = {0, 0, 0, 0, 0, 0};
#pragma unroll
        for (int j = TJ_elptr[e]; j < TJ_elptr[e + 1]; j++) {
          const int slot = TJ_elslot[j];
          const double q = a.case_q ? Q[slot * CS] : V.elq[slot];
          LoadProj pr;
          pr.plong = LP[2 * slot]; pr.pperp = LP[2 * slot + 1];
          double gl[6], ex[6];
          elem_load_p(TJ_elkind[slot], type, g, q, V.elq2[slot], pr, Ee * Ae, gl, ex);
#pragma unroll
          for (int k = 0; k < 6; k++) ext[k] += ex[k];
        }
#if TJ_BULK && TJ_SC
        reinterpret_cast<double2*>(rs + 6 * e)[0] = make_double2(-es[0] + ext[0], -es[1] + ext[1]);
        reinterpret_cast<double2*>(rs + 6 * e)[1] = make_double2(-es[2] + ext[2], es[3] + ext[3]);
        reinterpret_cast<double2*>(rs + 6 * e)[2] = make_double2(es[4] + ext[4], es[5] + ext[5]);
#elif TJ_BULK
        reinterpret_cast<double2*>(rs + 12 * e)[0] = make_double2(-es[0] + ext[0], -es[1] + ext[1]);
        rs[12 * e + 5] = -es[2] + ext[2];
        reinterpret_cast<double2*>(rs + 12 * e)[3] = make_double2(es[3] + ext[3], es[4] + ext[4]);
        rs[12 * e + 11] = es[5] + ext[5];
#else
        __syncwarp();
        SS[tj_col(12, 0)] = -es[0] + ext[0]; SS[tj_col(12, 1)] = -es[1] + ext[1]; SS[tj_col(12, 5)] = -es[2] + ext[2];
        SS[tj_col(12, 6)] = es[3] + ext[3];  SS[tj_col(12, 7)] = es[4] + ext[4];  SS[tj_col(12, 11)] = es[5] + ext[5];
        __syncwarp();
        double* gs = a.stress + (c0 * E + e) * 12;
        if (!st_on) {
        } else if (cnt == 32) {
#pragma unroll
          for (int r = 0; r < 12; r++) gs[go12[r % 3] + (r / 3) * 96 * E] = SSW[map12.so[r % 3] + (r / 3) * 8];
        } else {
          for (int j = lane; j < cnt * 12; j += 32) {
            const int t = j / 12, i = j - t * 12;
            gs[t * E * 12 + i] = SSW[tj_col(12, i) + t];
          }
        }
#endif
      }
    }
#if TJ_SUMM
    {
      double fx = 0.0, fy = 0.0;
#pragma unroll
      for (int i = 0; i < NB; i++) {
        if (TJ_bnddof[i] == 0) fx += fu[TJ_bndk[i]];  // OBGPU_DX
        if (TJ_bnddof[i] == 1) fy += fu[TJ_bndk[i]];  // OBGPU_DY
      }
#if TJ_BULK
      reinterpret_cast<double2*>(RM + lane * 4)[0] = make_double2(umax, fx);
      reinterpret_cast<double2*>(RM + lane * 4)[1] = make_double2(fy, double(info));
#else
      if (valid && st_on) {
        reinterpret_cast<double4*>(a.summ)[cs] = make_double4(umax, fx, fy, double(info));
        for (int p = 0; p < a.n_summ_peer; p++)
          reinterpret_cast<double4*>(a.summ_peer[p])[cs] = make_double4(umax, fx, fy, double(info));
      }
#endif
    }
#endif
#if TJ_BULK
#pragma unroll
    for (int i = 0; i < NB; i++) RB[lane * NB + i] = fu[TJ_bndk[i]];
    tj_fence_async();
    __syncwarp();
    if (st_on) {
      if (a.stress) tj_bulk_store(a.stress + c0 * E * SR, RS, E * SR * 8, cnt, lane);
      if (NB > 0) tj_bulk_store(a.F_b + c0 * NB, RB, NB * 8, cnt, lane);
      if (TJ_SUMM) {
        tj_bulk_store(a.summ + c0 * 4, RM, 32, cnt, lane);
        for (int p = 0; p < a.n_summ_peer; p++) tj_bulk_store(a.summ_peer[p] + c0 * 4, RM, 32, cnt, lane);
      }
    }
#else
    __syncwarp();
#pragma unroll
    for (int i = 0; i < NB; i++) SX[tj_col(NBC, i)] = fu[TJ_bndk[i]];
    __syncwarp();
    if (NB > 0 && st_on) mapNB.store(a.F_b + c0 * NB, SXW, lane, cnt);
    __syncwarp();
#endif
  }
#if TJ_BULK
  if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");  // staging rows stay valid until read
  __syncwarp();
#endif
}

}  // namespace obgpu
