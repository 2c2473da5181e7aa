// kernels_frame.cu — K1+K2+K4+K5+K6+K10 fused for batches of DIFFERENT small structures (BASELINE config 3:
// random 2D frames, 10-60 elements, n_f up to ~180).  One WARP per system (structure s, load case c):
//   stage A  lane = element (chunks of 16): orientation, local stiffness, rotation to global (beam_math.cuh),
//            summed equivalent nodal loads of the element's DISTRIB_UNIFORM loads -> shared-memory staging;
//            the staged elements are then scattered ONE ELEMENT AT A TIME (lanes = the 21 compact entries + 6
//            load entries) into the banded K_ff / F_f in shared memory.  Entries of one element never collide,
//            elements are applied in the reference's order (assemble.cpp:123-223), so the sums are
//            deterministic and need no atomics.  Bounded rows/columns are dropped on the fly (BC elimination as
//            a gather through the per-node slots, assemble.cpp:249-341).
//   stage B  banded LDL^T in shared memory on the unscaled factor V = L D, by PANELS of 4 columns: the panel's pivots are
//            eliminated in registers (lane = window row), the trailing triangle gets one rank-4 update per panel with
//            a 2 x 2 tile per lane; L y = b rides along in the panel; L^T x = D^-1 y as a column sweep.
//            Non-positive pivot -> info = row+1 (solver_static.cpp:97-105).
//   stage C  lane = element: end forces (postproc_stress.cpp:86-165) staged and written coalesced; global end
//            forces of the elements that touch a support scattered into the bounded DOFs = K_bf*U_f
//            (solver_static.cpp:196-218).
// Structures are served in SIZE CLASSES (host: capi_general.cu): shared memory per warp is sized by the class
// maxima, so small frames run at high occupancy, and the kernel is instantiated per number of update rounds.
#include "beam_math.cuh"
#include "kernels.h"

namespace obgpu {

#ifndef OBGPU_FRAME_MINB
#define OBGPU_FRAME_MINB(MAXR) 4   // CTAs per SM the register allocation aims at: 128 registers, no spills (5 and 6 CTAs
                                  // were measured: the spill code costs more than the extra warps hide)
#endif
#ifdef OBGPU_FRAME_CLOCKS  // experiment build only: per-stage cycle sums (lane 0 of every warp), read by scripts/exp
__device__ unsigned long long g_frame_clk[8];
#define OB_CLK(i)                                               \
  do {                                                          \
    const long long t_now = clock64();                          \
    if (lane == 0) atomicAdd(&g_frame_clk[i], (unsigned long long)(t_now - t_prev)); \
    t_prev = t_now;                                             \
  } while (0)
extern "C" void obgpu_debug_frame_clocks(unsigned long long* out, int reset) {
  cudaMemcpyFromSymbol(out, g_frame_clk, sizeof(g_frame_clk));
  if (reset) {
    unsigned long long z[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    cudaMemcpyToSymbol(g_frame_clk, z, sizeof(z));
  }
}
#else
#define OB_CLK(i)
#endif
#ifndef OBGPU_FRAME_SPLIT_MINB
#define OBGPU_FRAME_SPLIT_MINB 6
#endif
#ifndef OBGPU_FRAME_PRE_MINB
#define OBGPU_FRAME_PRE_MINB 6    // k_frame_pre / k_frame_post: 80 registers, 24 warps per SM (64 registers spill)
#endif
constexpr int STAGE_D = 27;  // doubles staged per element: 21 stiffness + 6 equivalent loads
constexpr int CHUNK = 16;    // elements computed and staged at a time.  Measured and rejected: 32 (the larger staging buffer costs
                             // resident warps), computing 32 in registers + staging in two halves (no fewer instructions), and
                             // 272-byte records that carry the 27 scatter offsets precomputed (5 % fewer instructions in the
                             // factorisation kernel, but the larger staging area costs the 56-KB class a resident CTA: slower)

// per-warp shared memory: band[max_band] | Ff[max_nf] | R[max_nb] | Eq[max_nb] | Dinv[max_nf] | stage[CHUNK*STAGE_D]
// | slots[CHUNK*6] (int).  Stage C reuses stage[] as [CHUNK][12] stress + [CHUNK][6] global end forces.
size_t frame_smem_per_warp(int max_band, int max_nf, int max_nb) {
  size_t d = size_t(max_band) + 2 * size_t(max_nf) + 2 * size_t(max_nb) + CHUNK * STAGE_D;
  return d * 8 + CHUNK * 6 * 4;
}

// compact entry cid (0..20) -> the two compact DOF positions (0..5 = end0 {x,y,rz}, end1 {x,y,rz})
__device__ __forceinline__ void cid_pair(int cid, int& a, int& b) {
  if (cid < 6) {
    a = cid < 3 ? 0 : cid < 5 ? 1 : 2;
    b = cid < 3 ? cid : cid < 5 ? cid - 2 : 2;
  } else if (cid < 12) {
    const int u = cid - 6;
    a = 3 + (u < 3 ? 0 : u < 5 ? 1 : 2);
    b = 3 + (u < 3 ? u : u < 5 ? u - 2 : 2);
  } else {
    const int u = cid - 12;
    a = u / 3;
    b = 3 + u % 3;
  }
}

// Stage B works on PANELS of 4 columns (see the kernel).  The trailing triangle of a panel is cut into 2 x 2 tiles, tile
// e = r*32 + lane is (P, Q), 0 <= Q <= P, rows P in order, i.e. the tiles of tile rows P < m are exactly e < m(m+1)/2.
// Window coordinates (row a, column c relative to the panel's pivot (j, j)): tile rows a0 = 4 + 2P, a0 + 1, tile columns
// c0 = 4 + 2Q, c0 + 1; V(j+a, j+c) sits (a*(LD-1) + c) doubles after the pivot.  (P, Q) depend on (round, lane) only, so
// the map is built once per structure.
template <int MAXR>
struct TileMap {
  int v[MAXR > 0 ? MAXR : 1], a0[MAXR > 0 ? MAXR : 1], c0[MAXR > 0 ? MAXR : 1];
  __device__ __forceinline__ void init(int lane, int LD) {
#pragma unroll
    for (int r = 0; r < MAXR; r++) {
      const int e = r * 32 + lane;
      int pp = int((sqrtf(8.0f * float(e) + 1.0f) + 1.0f) * 0.5f);
      while (pp * (pp - 1) / 2 > e) pp--;
      while (pp * (pp + 1) / 2 <= e) pp++;
      const int qq = e - pp * (pp - 1) / 2 + 1;  // 1 <= qq <= pp
      a0[r] = 4 + 2 * (pp - 1);
      c0[r] = 4 + 2 * (qq - 1);
      v[r] = a0[r] * (LD - 1) + c0[r];
    }
  }
};

// Rank-4 update of one 2 x 2 tile: V(a, c) -= sum_k V(a, k) * (V(c, k) / D(k)), k = the panel's 4 columns in order (the
// same operations, in the same order, as four successive column eliminations).  pu = unscaled panel rows, ps = scaled.
__device__ __forceinline__ void tile_update(double* bj, int voff, int a0, int c0, int LDm1, int nrows, const double2* pu01,
                                            const double2* pu23, const double2* ps01, const double2* ps23) {
  const bool row1 = a0 + 1 < nrows, offd = c0 < a0;  // second row inside the window; (a0, c0+1) below the diagonal
  double* v0 = bj + voff;
  double* v1 = v0 + LDm1;
  const double2 pa = pu01[a0], pb = pu23[a0], pc = pu01[a0 + 1], pd = pu23[a0 + 1];
  const double2 qa = ps01[c0], qb = ps23[c0], qc = ps01[c0 + 1], qd = ps23[c0 + 1];
  double v00 = v0[0], v01 = offd ? v0[1] : 0.0, v10 = row1 ? v1[0] : 0.0, v11 = row1 ? v1[1] : 0.0;
  v00 -= pa.x * qa.x; v01 -= pa.x * qc.x; v10 -= pc.x * qa.x; v11 -= pc.x * qc.x;
  v00 -= pa.y * qa.y; v01 -= pa.y * qc.y; v10 -= pc.y * qa.y; v11 -= pc.y * qc.y;
  v00 -= pb.x * qb.x; v01 -= pb.x * qd.x; v10 -= pd.x * qb.x; v11 -= pd.x * qd.x;
  v00 -= pb.y * qb.y; v01 -= pb.y * qd.y; v10 -= pd.y * qb.y; v11 -= pd.y * qd.y;
  v0[0] = v00;
  if (offd) v0[1] = v01;
  if (row1) {
    v1[0] = v10;
    v1[1] = v11;
  }
}

// Reciprocal of a pivot: hardware seed (about 20 bits) + two Newton steps, 5 instructions against the 9 + a range check of
// the correctly rounded 1.0 / d; the result is within an ulp of it (pivots of a stiffness matrix are far from the
// subnormal range the seed flushes).  OBGPU_FRAME_EXACT_RCP restores 1.0 / d.
__device__ __forceinline__ double pivot_rcp(double d) {
#ifdef OBGPU_FRAME_EXACT_RCP
  return 1.0 / d;
#else
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(d));
  double e = fma(-d, r, 1.0);
  r = fma(r, e, r);
  e = fma(-d, r, 1.0);
  return fma(r, e, r);
#endif
}

// GEN: the general path (nodal rotated frames and/or non-zero prescribed displacements) is a separate
// instantiation so that its extra state never costs the common kernel a register.
// SPLIT: the middle kernel of the three-kernel path (k_frame_pre -> this -> k_frame_post, see below): the element records
// arrive precomputed from global memory instead of being formed here, and stage C is left to k_frame_post.  (Without the
// element arithmetic it needs 80 registers instead of 128; 6, 7 or 8 CTAs per SM measured the same: the stage is bound by
// its instruction stream and shared-memory traffic, not by occupancy.)
template <int MAXR, bool GEN, bool SPLIT>
__global__ void __launch_bounds__(FRAME_WARPS * 32, SPLIT ? OBGPU_FRAME_SPLIT_MINB : OBGPU_FRAME_MINB(MAXR))
    k_frame_fused(FrameArgs a, int doubles_per_warp) {
  extern __shared__ __align__(16) double smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  double* band = smem + size_t(warp) * doubles_per_warp;
  double* Ff = band + a.max_band;
  double* R = Ff + a.max_nf;
  double* Eq = R + a.max_nb;
  double* Dinv = Eq + a.max_nb;                                        // reciprocal pivots
  double* stage = Dinv + a.max_nf;                                     // [CHUNK][STAGE_D] (A) / [CHUNK][12]+[CHUNK][6] (C)
  int32_t* slots = reinterpret_cast<int32_t*>(stage + CHUNK * STAGE_D);  // [CHUNK][6]
  const int64_t n_cls = a.o1 - a.o0;
  const int64_t n_sys = n_cls * a.L;
  const int64_t wstride = int64_t(gridDim.x) * FRAME_WARPS;
  int pa = 0, pb = 0;  // scatter role of this lane: compact entry (pa, pb) or load component pa
  if (lane < ELEM_K) cid_pair(lane, pa, pb);
  else if (lane < STAGE_D) pa = lane - ELEM_K;
  const int offF = int(Ff - band), offEq = int(Eq - band);
  for (int64_t it = int64_t(blockIdx.x) * FRAME_WARPS + warp; it < n_sys; it += wstride) {
    const int64_t c = it / n_cls;
    const int64_t s = a.order[a.o0 + (it - c * n_cls)];
    const int64_t sys = c * a.S + s;
    const FrameMeta m = a.meta[s];
    const int nf = m.nf, nb = m.nb, hb = m.hb, W = frame_ld(hb);  // row stride (>= hb + 1)
#ifdef OBGPU_FRAME_CLOCKS
    long long t_prev = clock64();
#endif
    for (int i = lane; i < W * nf; i += 32) band[i] = 0.0;
    for (int i = lane; i < nf; i += 32) Ff[i] = 0.0;
    for (int i = lane; i < nb; i += 32) { R[i] = 0.0; Eq[i] = 0.0; }
    __syncwarp();
    // ---- nodal loads (assemble.cpp:326-341): free DOFs only; sequential so repeated DOFs add in input order ----
    if (lane == 0) {
      const double* v = a.case_nl ? a.case_nl + c * a.NLtot + m.nl0 : a.nload_val + m.nl0;
      for (int l = 0; l < m.NL; l++) {
        const int32_t t = a.nload_slot[m.nl0 + l];
        if (t >= 0) Ff[t] += v[l];
      }
    }
    __syncwarp();
    const double* eai_base = a.case_eai ? a.case_eai + c * 3 * a.Etot : a.elem_EAI;
    const double* q_base = a.case_q ? a.case_q + c * a.ELtot : a.eload_q;
    OB_CLK(0);
    // ---- stage A ----  (pass 1 only with prescribed displacements: F_f -= K_fb U_b, solver_static.cpp:37-43)
    const double* Ub = GEN && a.U_b ? a.U_b + m.b0 : nullptr;
    const double* node_rot = GEN ? a.node_rot : nullptr;
    for (int pass = 0; pass < (Ub ? 2 : 1); pass++) {
    for (int eb = 0; eb < m.E; eb += CHUNK) {
      const int cnt = min(CHUNK, m.E - eb);
      if (SPLIT) {  // records of k_frame_pre: [FRAME_EREC] doubles per element = 21 stiffness + 6 loads + 6 int32 slots
        const double* src = a.erec + (m.e0 + eb) * FRAME_EREC;
        for (int i = lane; i < cnt * FRAME_EREC; i += 32) stage[i] = src[i];
      } else
      if (lane < cnt) {
        const int64_t e = m.e0 + eb + lane;
        const uint2 nd = a.elem_nodes[e];
        const int type = a.elem_type[e];
        const double2 p0 = a.node_xy[m.n0 + nd.x], p1 = a.node_xy[m.n0 + nd.y];
        const ElemGeom g = elem_geom(p0.x, p0.y, p1.x, p1.y);
        const ElemLocalK lk = elem_local_k(type, eai_base[3 * e], eai_base[3 * e + 1], eai_base[3 * e + 2], g);
        double ke[ELEM_K];
        elem_global_k(lk, g, ke);
        if (node_rot) {
          const double r0 = node_rot[m.n0 + nd.x], r1 = node_rot[m.n0 + nd.y];
          if (r0 != 0.0 || r1 != 0.0) rotate_nodal(ke, cos(r0), sin(r0), cos(r1), sin(r1));
        }
        double* st = stage + lane * STAGE_D;
        int32_t sl6[6];
#pragma unroll
        for (int k = 0; k < 3; k++) {
          sl6[k] = a.node_slot[3 * (m.n0 + nd.x) + k];
          sl6[3 + k] = a.node_slot[3 * (m.n0 + nd.y) + k];
        }
        double gsum[6] = {0, 0, 0, 0, 0, 0};
        if (pass == 0) {
#pragma unroll
          for (int k = 0; k < ELEM_K; k++) st[k] = ke[k];
          for (int j = a.el_ptr[e]; j < a.el_ptr[e + 1]; j++) {
            const int sl = a.el_order[j];
            double gl[6], ex[6];
            if (a.eload_type)
              elem_load(a.eload_type[sl], type, g, q_base[sl], a.eload_q2[sl], a.eload_dir[3 * sl], a.eload_dir[3 * sl + 1],
                        eai_base[3 * e] * eai_base[3 * e + 1], gl, ex);
            else
              uniform_load(type, g, q_base[sl], a.eload_dir[3 * sl], a.eload_dir[3 * sl + 1], gl, ex);
#pragma unroll
            for (int k = 0; k < 6; k++) gsum[k] += gl[k];
          }
        } else {  // free rows of the element: -sum over its bounded DOFs of K(a,b) U_b(b)
#pragma unroll
          for (int b = 0; b < 6; b++) {
            const int32_t tb = sl6[b];
            if (tb < 0 && tb != SLOT_NONE) {
              const double ub = Ub[-tb - 1];
#pragma unroll
              for (int r = 0; r < 6; r++)
                if (sl6[r] >= 0) gsum[r] -= ke_at(ke, r, b) * ub;
            }
          }
        }
#pragma unroll
        for (int k = 0; k < 6; k++) {
          st[ELEM_K + k] = gsum[k];
          slots[lane * 6 + k] = sl6[k];
        }
      }
      __syncwarp();
      OB_CLK(1);
      // Scatter, one element per step in the reference's order.  band | Ff | R | Eq are one contiguous block, so a
      // lane's target is a single offset from `band`; the next element's offset and value are fetched while the
      // current read-modify-write is in flight (the RMW chain is the only serial dependency left).
      auto fetch = [&](int ee, int& off, double& v) {
        const int32_t* sl = SPLIT ? reinterpret_cast<const int32_t*>(stage + ee * FRAME_EREC + STAGE_D) : slots + ee * 6;
        off = -1;
        v = 0.0;
        if (lane < STAGE_D && (pass == 0 || lane >= ELEM_K)) {
          v = stage[ee * (SPLIT ? FRAME_EREC : STAGE_D) + lane];
          const int32_t fi = sl[pa];
          if (lane < ELEM_K) {
            const int32_t fj = sl[pb];
            if (fi >= 0 && fj >= 0) {
              const int i = max(fi, fj), j = min(fi, fj);
              off = i * W + (j - i + hb);
            }
          } else if (fi >= 0) off = offF + fi;
          else if (!SPLIT && fi != SLOT_NONE) off = offEq + (-fi - 1);   // SPLIT: k_frame_post sums the bounded loads
        }
      };
      int off_n;
      double v_n;
      fetch(0, off_n, v_n);
      for (int ee = 0; ee < cnt; ee++) {
        const int off = off_n;
        const double v = v_n;
        if (ee + 1 < cnt) fetch(ee + 1, off_n, v_n);
        if (off >= 0) band[off] += v;
        __syncwarp();
      }
      OB_CLK(2);
    }
    if (pass == 0) {
      if (node_rot) {  // assemble.cpp:349-435: (Fx,Fy) of a rotated node's FREE DOFs -> nodal frame (bounded ones count as 0)
        for (int n = lane; n < m.N; n += 32) {
          const double r = node_rot[m.n0 + n];
          if (r != 0.0) {
            const int32_t sx = a.node_slot[3 * (m.n0 + n)], sy = a.node_slot[3 * (m.n0 + n) + 1];
            const double fx = sx >= 0 ? Ff[sx] : 0.0, fy = sy >= 0 ? Ff[sy] : 0.0;
            const double cr = cos(r), sr = sin(r);
            if (sx >= 0) Ff[sx] = cr * fx + sr * fy;
            if (sy >= 0) Ff[sy] = -sr * fx + cr * fy;
          }
        }
        __syncwarp();
      }
      if (a.F_f)
        for (int i = lane; i < nf; i += 32) a.F_f[c * a.nf_tot + m.f0 + (a.free_inv ? a.free_inv[m.f0 + i] : i)] = Ff[i];
    }
    }  // pass
    if (SPLIT) {
      // the element records of the structure this warp serves NEXT are asked into L2 now: their DRAM latency then passes
      // behind this structure's factorisation instead of in front of the next one's copy loop
      const int64_t it_n = it + wstride;
      if (it_n < n_sys) {
        const int64_t s_n = a.order[a.o0 + (it_n - (it_n / n_cls) * n_cls)];
        const int64_t e_n = a.meta[s_n].e0;
        const int bytes_n = a.meta[s_n].E * FRAME_EREC * 8;
        const char* pn = reinterpret_cast<const char*>(a.erec + e_n * FRAME_EREC);
        for (int o = lane * 128; o < bytes_n; o += 32 * 128) asm volatile("prefetch.global.L2 [%0];" ::"l"(pn + o));
      }
    }
    // ---- stage B: banded LDL^T (row i holds columns i-hb..i at band[i*W + (j-i+hb)]) ----
    // Right-looking, one column j per step, on the UNSCALED factor V = L D (V(i,j) = L(i,j) D(j)): the column is
    // never rewritten, the only division is the reciprocal pivot (kept in Dinv), and L y = b rides along
    // (F(j+q) -= V(j+q,j) * y(j)/D(j)).  The trailing update V(j+p,j+q) -= V(j+p,j) V(j+q,j) / D(j), 1 <= q <= p <= m,
    // is a triangle of m(m+1)/2 independent entries, flattened over the 32 lanes in rounds (TriMap): a round costs
    // 3 loads + 2 FP64 ops + 1 store and all lanes work; all loads of a column are issued ahead of its stores.
    int info = 0;
    const int q = lane + 1;
    if (MAXR > 0) {
      // PANELS of 4 columns (hb + 4 <= 32).  Lane p owns window row j+p: it holds V(j+p, j..j+3) and F(j+p) in registers, the
      // four pivots are eliminated there (values of the pivot rows travel by shuffle), and the finished panel goes to a
      // scratch copy [row][k] (unscaled and scaled by 1/D) that the trailing update reads with 16-byte loads.  The trailing
      // triangle is then updated ONCE per panel, every lane a 2 x 2 tile: 16 FMAs per 12 shared-memory accesses, where a
      // column-at-a-time update pays 4 accesses per FMA (the shared-memory pipe was the limiter of this stage).
      TileMap<MAXR> tm;
      tm.init(lane, W);
      const int LDm1 = W - 1;
      double2* pu01 = reinterpret_cast<double2*>((reinterpret_cast<uintptr_t>(stage) + 15) & ~uintptr_t(15));  // scratch over stage[]
      double2* pu23 = pu01 + 32;
      double2* ps01 = pu23 + 32;
      double2* ps23 = ps01 + 32;
      for (int j = 0; j < nf; j += 4) {
        double* bj = band + j * W + hb;  // (j, j)
        const int nrows = min(hb + 4, nf - j);
        double* rp = bj + lane * LDm1;   // V(j+lane, j+k) = rp[k]
        double x0 = 0.0, x1 = 0.0, x2 = 0.0, x3 = 0.0, f = 0.0;
        const bool in = lane < nrows;
        const bool e0 = in && lane <= hb, e1 = in && lane >= 1 && lane <= hb + 1, e2 = in && lane >= 2 && lane <= hb + 2,
                   e3 = in && lane >= 3;
        if (e0) x0 = rp[0];
        if (e1) x1 = rp[1];
        if (e2) x2 = rp[2];
        if (e3) x3 = rp[3];
        if (in) f = Ff[j + lane];
        // The four pivots of the panel, in order.  Straight-line code: a non-positive pivot (Eigen's LLT stops there,
        // solver_static.cpp:97-105) latches info = its row + 1 and is replaced by 1, so that no branch separates the
        // warp shuffles; what is computed after it is never used (info != 0 skips the back substitution).
        double i0, i1, i2, i3;
        {  // column j
          double d = __shfl_sync(0xffffffffu, x0, 0);
          if (!(d > 0.0)) {
            if (info == 0) info = j + 1;
            d = 1.0;
          }
          i0 = pivot_rcp(d);
          const double t = __shfl_sync(0xffffffffu, f, 0) * i0;
          const double c1 = __shfl_sync(0xffffffffu, x0, 1) * i0, c2 = __shfl_sync(0xffffffffu, x0, 2) * i0,
                       c3 = __shfl_sync(0xffffffffu, x0, 3) * i0;
          if (lane >= 1) f -= x0 * t;
          if (lane >= 1) x1 -= x0 * c1;
          if (lane >= 2) x2 -= x0 * c2;
          if (lane >= 3) x3 -= x0 * c3;
        }
        {  // column j+1 (rows past the end of the matrix hold zeros: their "pivot" is replaced by 1 without an error)
          double d = __shfl_sync(0xffffffffu, x1, 1);
          if (!(d > 0.0)) {
            if (info == 0 && nrows > 1) info = j + 2;
            d = 1.0;
          }
          i1 = pivot_rcp(d);
          const double t = __shfl_sync(0xffffffffu, f, 1) * i1;
          const double c2 = __shfl_sync(0xffffffffu, x1, 2) * i1, c3 = __shfl_sync(0xffffffffu, x1, 3) * i1;
          if (lane >= 2) f -= x1 * t;
          if (lane >= 2) x2 -= x1 * c2;
          if (lane >= 3) x3 -= x1 * c3;
        }
        {  // column j+2
          double d = __shfl_sync(0xffffffffu, x2, 2);
          if (!(d > 0.0)) {
            if (info == 0 && nrows > 2) info = j + 3;
            d = 1.0;
          }
          i2 = pivot_rcp(d);
          const double t = __shfl_sync(0xffffffffu, f, 2) * i2;
          const double c3 = __shfl_sync(0xffffffffu, x2, 3) * i2;
          if (lane >= 3) f -= x2 * t;
          if (lane >= 3) x3 -= x2 * c3;
        }
        {  // column j+3
          double d = __shfl_sync(0xffffffffu, x3, 3);
          if (!(d > 0.0)) {
            if (info == 0 && nrows > 3) info = j + 4;
            d = 1.0;
          }
          i3 = pivot_rcp(d);
          const double t = __shfl_sync(0xffffffffu, f, 3) * i3;
          if (lane >= 4) f -= x3 * t;
        }
        if (in) Ff[j + lane] = f;
        if (e1) rp[1] = x1;  // column j was final already
        if (e2) rp[2] = x2;
        if (e3) rp[3] = x3;
        if (lane == 0) {
          Dinv[j] = i0;
          if (nrows > 1) Dinv[j + 1] = i1;
          if (nrows > 2) Dinv[j + 2] = i2;
          if (nrows > 3) Dinv[j + 3] = i3;
        }
        pu01[lane] = make_double2(x0, x1);
        pu23[lane] = make_double2(x2, x3);
        ps01[lane] = make_double2(x0 * i0, x1 * i1);
        ps23[lane] = make_double2(x2 * i2, x3 * i3);
        __syncwarp();
        const int mt = (nrows - 4 + 1) >> 1;       // tile rows of the trailing triangle (<= 0: none)
        const int rem = mt * (mt + 1) / 2 - lane;  // tile (r, lane) exists iff rem > 32 r
#pragma unroll
        for (int r = 0; r < MAXR; r++)
          if (rem > 32 * r) tile_update(bj, tm.v[r], tm.a0[r], tm.c0[r], LDm1, nrows, pu01, pu23, ps01, ps23);
        __syncwarp();
      }
    } else {  // wide bands (hb >= 32: more than 10 nodes per storey): lanes stride over the columns of each row
      for (int j = 0; j < nf; j++) {
        const double d = band[j * W + hb];
        if (!(d > 0.0)) {
          info = j + 1;
          break;
        }
        const double inv = 1.0 / d;
        const double tj = Ff[j] * inv;
        const int mrows = min(hb, nf - 1 - j);
        for (int i = j + 1; i <= j + mrows; i++) {
          const double lij = band[i * W + (j - i + hb)] * inv;
          for (int k = j + 1 + lane; k <= i; k += 32) band[i * W + (k - i + hb)] -= lij * band[k * W + (j - k + hb)];
        }
        for (int i = j + 1 + lane; i <= j + mrows; i += 32) Ff[i] -= band[i * W + (j - i + hb)] * tj;
        if (lane == 0) Dinv[j] = inv;
        __syncwarp();
      }
    }
    OB_CLK(3);
    if (info == 0 && MAXR > 0) {
      // L^T x = D^-1 y as a column sweep from the last row up: x(j) = Dinv(j) * (y(j) - sum_k V(k,j) x(k)).  The
      // running sums live in REGISTERS: row i is carried by lane i%32 from the step it first receives a term
      // (j = i + hb) to its own step (hb < 32, so the rows in flight sit in distinct lanes); x(j) reaches the
      // other lanes by shuffle, so the serial chain of a step is shuffle + 2 FP64 ops with no shared-memory
      // round trip or barrier.
      // Loop state is carried incrementally (dist = j - myrow, po = offset of V(j, myrow), the source lane of x(j)):
      // the step is a predicated load, a shuffle, a multiply and an FMA plus four integer updates.
      constexpr int FAR = 1 << 20;  // "no row": never within the band
      int myrow = (nf - 1) - ((nf - 1 - lane) & 31);
      int dist = myrow >= 0 ? (nf - 1) - myrow : FAR;
      int po = (nf - 1) * W + hb - dist;
      int jl = (nf - 1) & 31;
      double acc = myrow >= 0 ? Ff[myrow] : 0.0;
      for (int j = nf - 1; j >= 0; j--) {
        const double vj = unsigned(dist - 1) < unsigned(hb) ? band[po] : 0.0;
        const double xj = __shfl_sync(0xffffffffu, acc, jl) * Dinv[j];
        if (dist == 0) {
          Ff[j] = xj;
          myrow -= 32;
          dist = myrow >= 0 ? 32 : FAR;
          po -= 32;
          acc = myrow >= 0 ? Ff[myrow] : 0.0;
        } else {
          acc -= vj * xj;
        }
        dist--;
        po -= W - 1;
        jl = (jl - 1) & 31;
      }
      __syncwarp();
    } else if (info == 0) {
      for (int j = nf - 1; j >= 0; j--) {
        const double xj = Ff[j] * Dinv[j];
        for (int qq = q; qq <= min(hb, j); qq += 32) Ff[j - qq] -= band[j * W + hb - qq] * xj;
        __syncwarp();
        if (lane == 0) Ff[j] = xj;
      }
      __syncwarp();
    }
    if (lane == 0) {
      a.info[sys] = info;
      if (info) atomicOr(a.fail, 1);
    }
    for (int i = lane; i < nf; i += 32) a.U_f[c * a.nf_tot + m.f0 + (a.free_inv ? a.free_inv[m.f0 + i] : i)] = Ff[i];
    OB_CLK(4);
    if (SPLIT) {
      __syncwarp();
      continue;
    }
    // ---- stage C: element end forces, reactions ----
    double* sst = stage;                  // [CHUNK][12] stress staging
    double* sgf = stage + CHUNK * 12;     // [CHUNK][6]  global end forces
    for (int eb = 0; eb < m.E; eb += CHUNK) {
      const int cnt = min(CHUNK, m.E - eb);
      bool hasb = false;
      if (lane < cnt) {
        const int64_t e = m.e0 + eb + lane;
        const uint2 nd = a.elem_nodes[e];
        const int type = a.elem_type[e];
        const double2 p0 = a.node_xy[m.n0 + nd.x], p1 = a.node_xy[m.n0 + nd.y];
        const ElemGeom g = elem_geom(p0.x, p0.y, p1.x, p1.y);
        const ElemLocalK lk = elem_local_k(type, eai_base[3 * e], eai_base[3 * e + 1], eai_base[3 * e + 2], g);
        // postproc_stress.cpp:124-125 reads the lengths from the CURRENT node poses and the orientation from the element's
        // cache: after nonLinearIterative these differ (solver_static.cpp:190-193 restores the poses)
        ElemLocalK lks = lk;
        if (GEN && a.node_xy_len) {
          const double2 q0 = a.node_xy_len[m.n0 + nd.x], q1 = a.node_xy_len[m.n0 + nd.y];
          lks = elem_local_k(type, eai_base[3 * e], eai_base[3 * e + 1], eai_base[3 * e + 2], elem_geom(q0.x, q0.y, q1.x, q1.y));
        }
        double ug[6], es[6];
        int32_t sl6[6];
#pragma unroll
        for (int k = 0; k < 3; k++) {
          sl6[k] = a.node_slot[3 * (m.n0 + nd.x) + k];
          sl6[3 + k] = a.node_slot[3 * (m.n0 + nd.y) + k];
        }
#pragma unroll
        for (int k = 0; k < 6; k++) {
          const int32_t t = sl6[k];
          slots[lane * 6 + k] = t;
          hasb |= t < 0 && t != SLOT_NONE;
          // postproc_stress.cpp:86-119: U as solved (nodal frames are NOT undone there, quirk Q6), U_b on bounded DOFs
          ug[k] = t >= 0 ? Ff[t] : (Ub && t != SLOT_NONE) ? Ub[-t - 1] : 0.0;
        }
        elem_end_forces(lks, g, ug, es);
        if (!Ub && !node_rot && !(GEN && a.node_xy_len)) {  // K_bf U_f of this element = its global end forces (U_b = 0, identity node frames)
          sgf[lane * 6 + 0] = g.c * es[0] - g.s * es[1];
          sgf[lane * 6 + 1] = g.s * es[0] + g.c * es[1];
          sgf[lane * 6 + 2] = es[2];
          sgf[lane * 6 + 3] = g.c * es[3] - g.s * es[4];
          sgf[lane * 6 + 4] = g.s * es[3] + g.c * es[4];
          sgf[lane * 6 + 5] = es[5];
        } else {
          // general reactions, solver_static.cpp:196-200: bounded row r of K_bf U_f + K_bb U_b with the element's
          // block in nodal frames; K_bb holds only its upper triangle in the reference and is multiplied as a
          // general matrix, so a bounded column c contributes to row r only when c >= r
          double ke[ELEM_K];
          elem_global_k(lk, g, ke);
          if (node_rot) {
            const double r0 = node_rot[m.n0 + nd.x], r1 = node_rot[m.n0 + nd.y];
            if (r0 != 0.0 || r1 != 0.0) rotate_nodal(ke, cos(r0), sin(r0), cos(r1), sin(r1));
          }
#pragma unroll
          for (int r = 0; r < 6; r++) {
            double acc = 0.0;
            const int32_t tr = sl6[r];
            if (tr < 0 && tr != SLOT_NONE) {
#pragma unroll
              for (int b = 0; b < 6; b++) {
                const int32_t tb = sl6[b];
                if (tb >= 0) acc += ke_at(ke, r, b) * Ff[tb];
                else if (Ub && tb != SLOT_NONE && -tb >= -tr) acc += ke_at(ke, r, b) * Ub[-tb - 1];
              }
            }
            sgf[lane * 6 + r] = acc;
          }
        }
        if (a.stress) {
          double ext[6] = {0, 0, 0, 0, 0, 0};
          for (int j = a.el_ptr[e]; j < a.el_ptr[e + 1]; j++) {
            const int sl = a.el_order[j];
            double gl[6], ex[6];
            if (a.eload_type)
              elem_load(a.eload_type[sl], type, g, q_base[sl], a.eload_q2[sl], a.eload_dir[3 * sl], a.eload_dir[3 * sl + 1],
                        eai_base[3 * e] * eai_base[3 * e + 1], gl, ex);
            else
              uniform_load(type, g, q_base[sl], a.eload_dir[3 * sl], a.eload_dir[3 * sl + 1], gl, ex);
#pragma unroll
            for (int k = 0; k < 6; k++) ext[k] += ex[k];
          }
          double* o = sst + lane * 12;  // {N,Vy,Vz,Mx,My,Mz} x 2 faces; face 0 negated (postproc_stress.cpp:152-165)
          o[0] = -es[0] + ext[0]; o[1] = -es[1] + ext[1]; o[2] = 0.0; o[3] = 0.0; o[4] = 0.0; o[5] = -es[2] + ext[2];
          o[6] = es[3] + ext[3];  o[7] = es[4] + ext[4];  o[8] = 0.0; o[9] = 0.0; o[10] = 0.0; o[11] = es[5] + ext[5];
        }
      }
      __syncwarp();
      if (a.stress) {
        double* o = a.stress + (c * a.Etot + m.e0 + eb) * 12;
        for (int i = lane; i < cnt * 12; i += 32) o[i] = sst[i];
      }
      unsigned todo = __ballot_sync(0xffffffffu, hasb);  // elements touching a support, still in element order
      while (todo) {
        const int ee = __ffs(todo) - 1;
        todo &= todo - 1;
        if (lane < 6) {
          const int32_t t = slots[ee * 6 + lane];
          if (t < 0 && t != SLOT_NONE) R[-t - 1] += sgf[ee * 6 + lane];
        }
        __syncwarp();
      }
    }
    // F_b = K_bf*U_f - equivalent loads on bounded DOFs (solver_static.cpp:196-218)
    for (int i = lane; i < nb; i += 32) a.F_b[c * a.nb_tot + m.b0 + i] = R[i] - Eq[i];
    __syncwarp();
    OB_CLK(5);
  }
}

// ---- the three-kernel path (FrameArgs::erec != null): k_frame_pre -> k_frame_fused<., false, true> per class -> k_frame_post
// For one load case, identity node frames and U_b = 0.  The element arithmetic of stages A and C runs in two kernels of
// their own — warp per structure, lane per element, almost no shared memory, so 48-64 resident warps per SM instead of 16 —
// and the factorisation kernel in between keeps only what it needs.  The price is one round trip of the element records
// through HBM (240 B per element written and read once).  Sums are formed in the same order as in the fused kernel.

// k_frame_pre: stage A's arithmetic.  Record of element e at erec[e * FRAME_EREC]: the 21 compact global stiffness entries,
// the 6 summed equivalent nodal loads, the 6 node slots (int32) of its two ends.
constexpr int PRE_LD = FRAME_EREC + 1;  // odd row stride of the staging rows: the per-lane record writes hit 16 banks
__global__ void __launch_bounds__(128, OBGPU_FRAME_PRE_MINB) k_frame_pre(FrameArgs a) {
  // The records leave through a shared-memory transpose: a lane forms its element's record, the warp then writes the
  // chunk's records — contiguous in erec — as whole 128-byte lines (per-lane stores at a 240-byte stride cost one L2
  // write transaction per lane and instruction, and that, not HBM, bounded the first version of this kernel).
  __shared__ double rows_all[4][32 * PRE_LD];
  const int lane = threadIdx.x & 31;
  double* rows = rows_all[threadIdx.x >> 5];
  const int64_t wstride = int64_t(gridDim.x) * (blockDim.x >> 5);
  for (int64_t s = int64_t(blockIdx.x) * (blockDim.x >> 5) + (threadIdx.x >> 5); s < a.S; s += wstride) {
    const FrameMeta m = a.meta[s];
    for (int eb = 0; eb < m.E; eb += 32) {
      const int cnt = min(32, m.E - eb);
      if (lane < cnt) {
        const int64_t e = m.e0 + eb + lane;
        const uint2 nd = a.elem_nodes[e];
        const int type = a.elem_type[e];
        const double2 p0 = a.node_xy[m.n0 + nd.x], p1 = a.node_xy[m.n0 + nd.y];
        const ElemGeom g = elem_geom(p0.x, p0.y, p1.x, p1.y);
        const ElemLocalK lk = elem_local_k(type, a.elem_EAI[3 * e], a.elem_EAI[3 * e + 1], a.elem_EAI[3 * e + 2], g);
        double* o = rows + lane * PRE_LD;
        {
          double ke[ELEM_K];
          elem_global_k(lk, g, ke);
#pragma unroll
          for (int k = 0; k < ELEM_K; k++) o[k] = ke[k];
        }
        double gsum[6] = {0, 0, 0, 0, 0, 0};
        for (int j = a.el_ptr[e]; j < a.el_ptr[e + 1]; j++) {
          const int sl = a.el_order[j];
          double gl[6], ex[6];
          if (a.eload_type)
            elem_load(a.eload_type[sl], type, g, a.eload_q[sl], a.eload_q2[sl], a.eload_dir[3 * sl], a.eload_dir[3 * sl + 1],
                      a.elem_EAI[3 * e] * a.elem_EAI[3 * e + 1], gl, ex);
          else
            uniform_load(type, g, a.eload_q[sl], a.eload_dir[3 * sl], a.eload_dir[3 * sl + 1], gl, ex);
#pragma unroll
          for (int k = 0; k < 6; k++) gsum[k] += gl[k];
        }
#pragma unroll
        for (int k = 0; k < 6; k++) o[ELEM_K + k] = gsum[k];
        int32_t* os = reinterpret_cast<int32_t*>(o + STAGE_D);
#pragma unroll
        for (int k = 0; k < 3; k++) {
          os[k] = a.node_slot[3 * (m.n0 + nd.x) + k];
          os[3 + k] = a.node_slot[3 * (m.n0 + nd.y) + k];
        }
      }
      __syncwarp();
      double* dst = a.erec + (m.e0 + eb) * FRAME_EREC;
      for (int i = lane; i < cnt * FRAME_EREC; i += 32) {
        const int r = i / FRAME_EREC;
        dst[i] = rows[r * PRE_LD + (i - r * FRAME_EREC)];
      }
      __syncwarp();
    }
  }
}

// k_frame_post: stage C (postproc_stress.cpp:86-165, solver_static.cpp:196-218).  Per-warp shared memory:
// sg[32][12] (global end forces | equivalent loads of the chunk's elements) | slots[32][6] | R[max_nb] | Eq[max_nb]
constexpr int POST_WARPS = 4;
__global__ void __launch_bounds__(POST_WARPS * 32, OBGPU_FRAME_PRE_MINB) k_frame_post(FrameArgs a, int max_nb) {
  extern __shared__ __align__(16) double smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int per_warp = 32 * 12 + 32 * 3 + 2 * max_nb;
  double* sg = smem + size_t(warp) * per_warp;
  int32_t* slots = reinterpret_cast<int32_t*>(sg + 32 * 12);
  double* R = sg + 32 * 12 + 32 * 3;
  double* Eq = R + max_nb;
  const int64_t wstride = int64_t(gridDim.x) * POST_WARPS;
  for (int64_t s = int64_t(blockIdx.x) * POST_WARPS + warp; s < a.S; s += wstride) {
    const FrameMeta m = a.meta[s];
    const int nb = m.nb;
    for (int i = lane; i < nb; i += 32) { R[i] = 0.0; Eq[i] = 0.0; }
    const double* U = a.U_f + m.f0;
    const int32_t* finv = a.free_inv ? a.free_inv + m.f0 : nullptr;
    __syncwarp();
    for (int eb = 0; eb < m.E; eb += 32) {
      const int cnt = min(32, m.E - eb);
      bool hasb = false;
      if (lane < cnt) {
        const int64_t e = m.e0 + eb + lane;
        const uint2 nd = a.elem_nodes[e];
        const int type = a.elem_type[e];
        const double2 p0 = a.node_xy[m.n0 + nd.x], p1 = a.node_xy[m.n0 + nd.y];
        const ElemGeom g = elem_geom(p0.x, p0.y, p1.x, p1.y);
        const ElemLocalK lk = elem_local_k(type, a.elem_EAI[3 * e], a.elem_EAI[3 * e + 1], a.elem_EAI[3 * e + 2], g);
        double ug[6], es[6];
#pragma unroll
        for (int k = 0; k < 6; k++) {
          const int32_t t = a.node_slot[3 * (m.n0 + (k < 3 ? nd.x : nd.y)) + (k < 3 ? k : k - 3)];
          slots[lane * 6 + k] = t;
          hasb |= t < 0 && t != SLOT_NONE;
          ug[k] = t >= 0 ? U[finv ? finv[t] : t] : 0.0;   // postproc_stress.cpp:86-119 (U_b = 0 on this path)
        }
        if (hasb) {  // the element's summed equivalent loads (k_frame_pre's record): only supports need them here
          const double* rec = a.erec + e * FRAME_EREC + ELEM_K;
#pragma unroll
          for (int k = 0; k < 6; k++) sg[lane * 12 + 6 + k] = rec[k];
        }
        elem_end_forces(lk, g, ug, es);
        // K_bf U_f of this element = its global end forces (U_b = 0, identity node frames)
        sg[lane * 12 + 0] = g.c * es[0] - g.s * es[1];
        sg[lane * 12 + 1] = g.s * es[0] + g.c * es[1];
        sg[lane * 12 + 2] = es[2];
        sg[lane * 12 + 3] = g.c * es[3] - g.s * es[4];
        sg[lane * 12 + 4] = g.s * es[3] + g.c * es[4];
        sg[lane * 12 + 5] = es[5];
        if (a.stress) {
          double ext[6] = {0, 0, 0, 0, 0, 0};
          for (int j = a.el_ptr[e]; j < a.el_ptr[e + 1]; j++) {
            const int sl = a.el_order[j];
            double gl[6], ex[6];
            if (a.eload_type)
              elem_load(a.eload_type[sl], type, g, a.eload_q[sl], a.eload_q2[sl], a.eload_dir[3 * sl], a.eload_dir[3 * sl + 1],
                        a.elem_EAI[3 * e] * a.elem_EAI[3 * e + 1], gl, ex);
            else
              uniform_load(type, g, a.eload_q[sl], a.eload_dir[3 * sl], a.eload_dir[3 * sl + 1], gl, ex);
#pragma unroll
            for (int k = 0; k < 6; k++) ext[k] += ex[k];
          }
          // {N,Vy,Vz,Mx,My,Mz} x 2 faces; face 0 negated (postproc_stress.cpp:152-165).  96 B = three whole sectors per
          // lane, written directly: a shared-memory transpose like k_frame_pre's measured slower here (785 against 720 us)
          double2* o = reinterpret_cast<double2*>(a.stress + e * 12);
          o[0] = make_double2(-es[0] + ext[0], -es[1] + ext[1]);
          o[1] = make_double2(0.0, 0.0);
          o[2] = make_double2(0.0, -es[2] + ext[2]);
          o[3] = make_double2(es[3] + ext[3], es[4] + ext[4]);
          o[4] = make_double2(0.0, 0.0);
          o[5] = make_double2(0.0, es[5] + ext[5]);
        }
      }
      __syncwarp();
      unsigned todo = __ballot_sync(0xffffffffu, hasb);  // elements touching a support, in element order
      while (todo) {
        const int ee = __ffs(todo) - 1;
        todo &= todo - 1;
        if (lane < 6) {
          const int32_t t = slots[ee * 6 + lane];
          if (t < 0 && t != SLOT_NONE) {
            R[-t - 1] += sg[ee * 12 + lane];
            Eq[-t - 1] += sg[ee * 12 + 6 + lane];
          }
        }
        __syncwarp();
      }
    }
    // F_b = K_bf*U_f - equivalent loads on bounded DOFs (solver_static.cpp:196-218)
    for (int i = lane; i < nb; i += 32) a.F_b[m.b0 + i] = R[i] - Eq[i];
    __syncwarp();
  }
}

cudaError_t launch_frame_pre(const FrameArgs& a, int sm_count, cudaStream_t st) {
  int64_t grid = (a.S + 3) / 4;
  const int64_t cap = int64_t(sm_count) * OBGPU_FRAME_PRE_MINB;
  if (grid > cap) grid = cap;
  k_frame_pre<<<unsigned(grid < 1 ? 1 : grid), 128, 0, st>>>(a);
  return cudaGetLastError();
}

cudaError_t launch_frame_post(const FrameArgs& a, int max_nb, int sm_count, cudaStream_t st) {
  const size_t smem = size_t(32 * 12 + 32 * 3 + 2 * max_nb) * 8 * POST_WARPS;
  if (smem > 227 * 1024) return cudaErrorInvalidConfiguration;
  cudaError_t e = cudaFuncSetAttribute(k_frame_post, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
  if (e != cudaSuccess) return e;
  int per_sm = 1;
  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_frame_post, POST_WARPS * 32, smem);
  if (e != cudaSuccess) return e;
  int64_t grid = int64_t(sm_count) * (per_sm < 1 ? 1 : per_sm);
  const int64_t need = (a.S + POST_WARPS - 1) / POST_WARPS;
  if (grid > need) grid = need;
  k_frame_post<<<unsigned(grid < 1 ? 1 : grid), POST_WARPS * 32, smem, st>>>(a, max_nb);
  return cudaGetLastError();
}

template <int MAXR, bool GEN, bool SPLIT>
static cudaError_t launch_frame_rg(const FrameArgs& a, int sm_count, cudaStream_t st) {
  const size_t per_warp = frame_smem_per_warp(a.max_band, a.max_nf, a.max_nb);
  const size_t smem = per_warp * FRAME_WARPS;
  if (smem > 227 * 1024) return cudaErrorInvalidConfiguration;
  auto kern = k_frame_fused<MAXR, GEN, SPLIT>;
  // always the device maximum: the attribute is per function, not per launch, and contexts on other host threads launch
  // the same instantiation with other class sizes (a smaller value set there would fail this launch)
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
  if (e != cudaSuccess) return e;
  int per_sm = 1;
  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, FRAME_WARPS * 32, smem);
  if (e != cudaSuccess) return e;
  if (per_sm < 1) per_sm = 1;
  const int64_t n_sys = (a.o1 - a.o0) * a.L;
  int64_t grid = int64_t(sm_count) * per_sm;  // persistent: a whole number of CTAs per SM, warps stride over systems
  const int64_t need = (n_sys + FRAME_WARPS - 1) / FRAME_WARPS;
  if (grid > need) grid = need;
  if (grid < 1) grid = 1;
  kern<<<unsigned(grid), FRAME_WARPS * 32, smem, st>>>(a, int(per_warp / 8));
  return cudaGetLastError();
}

template <int MAXR>
static cudaError_t launch_frame_r(const FrameArgs& a, int sm_count, cudaStream_t st) {
  if (a.U_b || a.node_rot || a.node_xy_len) return launch_frame_rg<MAXR, true, false>(a, sm_count, st);
  if (a.erec) return launch_frame_rg<MAXR, false, true>(a, sm_count, st);
  return launch_frame_rg<MAXR, false, false>(a, sm_count, st);
}

cudaError_t launch_frame_fused(const FrameArgs& a, int sm_count, cudaStream_t st) {
  if (a.o1 <= a.o0) return cudaSuccess;
  if (a.max_hb + 4 > 32) return launch_frame_r<0>(a, sm_count, st);  // a panel's window (hb + 4 rows) must fit the 32 lanes
  const int ht = (a.max_hb + 1) / 2;                 // tile rows of the widest band in the class
  const int rounds = (ht * (ht + 1) / 2 + 31) / 32;  // 2 x 2 tiles per lane
  if (rounds <= 1) return launch_frame_r<1>(a, sm_count, st);
  if (rounds <= 2) return launch_frame_r<2>(a, sm_count, st);
  return launch_frame_r<4>(a, sm_count, st);
}

}  // namespace obgpu
