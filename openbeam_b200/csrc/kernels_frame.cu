// kernels_frame.cu — K1+K2+K4+K5+K6+K10 fused for batches of DIFFERENT small structures (BASELINE config 3:
// random 2D frames, 10-60 elements, n_f up to ~180).  One WARP per system (structure s, load case c):
//   stage A  lane = element: orientation, local stiffness, rotation to global (beam_math.cuh), summed
//            equivalent nodal loads of the element's DISTRIB_UNIFORM loads -> shared-memory staging;
//            then the 32 staged elements are scattered ONE ELEMENT AT A TIME (lanes = the 21 compact
//            entries + 6 load entries) into the banded K_ff / F_f in shared memory.  Entries of one element
//            never collide, elements are applied in the reference's order (assemble.cpp:123-223), so the sums
//            are deterministic and need no atomics.  Bounded rows/columns are dropped on the fly
//            (BC elimination as a gather through the per-node slots, assemble.cpp:249-341).
//   stage B  banded LDL^T in shared memory (lanes over the trailing band block), forward/back substitution
//            with warp-shuffle dot products.  Non-positive pivot -> info = row+1 (solver_static.cpp:97-105).
//   stage C  lane = element: end forces (postproc_stress.cpp:86-165) staged and written coalesced; global end
//            forces scattered element by element into the bounded DOFs = K_bf*U_f (solver_static.cpp:196-218).
#include "beam_math.cuh"
#include "kernels.h"

namespace obgpu {

constexpr int FRAME_WARPS = 4;       // warps per CTA
constexpr int STAGE_D = 27;          // doubles staged per element: 21 stiffness + 6 equivalent loads

// per-warp shared memory: band[max_band] | Ff[max_nf] | R[max_nb] | Eq[max_nb] | stage[32*STAGE_D] | slots[32*6] (int)
// (stage C reuses stage[] as [32][12] stress + [32][6] global end forces)
size_t frame_smem_per_warp(int max_band, int max_nf, int max_nb) {
  size_t d = size_t(max_band) + max_nf + 2 * size_t(max_nb) + 32 * STAGE_D;
  return d * 8 + 32 * 6 * 4;
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

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__global__ void __launch_bounds__(FRAME_WARPS * 32) k_frame_fused(FrameArgs a, int doubles_per_warp) {
  extern __shared__ __align__(16) double smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  double* band = smem + size_t(warp) * doubles_per_warp;
  double* Ff = band + a.max_band;
  double* R = Ff + a.max_nf;
  double* Eq = R + a.max_nb;
  double* stage = Eq + a.max_nb;                                   // [32][STAGE_D] (stage A) / [32][12]+[32][6] (stage C)
  int32_t* slots = reinterpret_cast<int32_t*>(stage + 32 * STAGE_D);  // [32][6]
  const int64_t n_sys = a.S * a.L;
  const int64_t wstride = int64_t(gridDim.x) * FRAME_WARPS;
  for (int64_t sys = int64_t(blockIdx.x) * FRAME_WARPS + warp; sys < n_sys; sys += wstride) {
    const int64_t c = sys / a.S, s = sys - c * a.S;
    const FrameMeta m = a.meta[s];
    const int nf = m.nf, nb = m.nb, hb = m.hb, W = hb + 1;
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
    // ---- stage A ----
    for (int eb = 0; eb < m.E; eb += 32) {
      const int cnt = min(32, m.E - eb);
      if (lane < cnt) {
        const int64_t e = m.e0 + eb + lane;
        const uint2 nd = a.elem_nodes[e];
        const int type = a.elem_type[e];
        const double2 p0 = a.node_xy[m.n0 + nd.x], p1 = a.node_xy[m.n0 + nd.y];
        const ElemGeom g = elem_geom(p0.x, p0.y, p1.x, p1.y);
        const ElemLocalK lk = elem_local_k(type, eai_base[3 * e], eai_base[3 * e + 1], eai_base[3 * e + 2], g);
        double ke[ELEM_K];
        elem_global_k(lk, g, ke);
        double* st = stage + lane * STAGE_D;
#pragma unroll
        for (int k = 0; k < ELEM_K; k++) st[k] = ke[k];
        double gsum[6] = {0, 0, 0, 0, 0, 0};
        for (int j = a.el_ptr[e]; j < a.el_ptr[e + 1]; j++) {
          const int sl = a.el_order[j];
          double gl[6], ex[6];
          uniform_load(type, g, q_base[sl], a.eload_dir[3 * sl], a.eload_dir[3 * sl + 1], gl, ex);
#pragma unroll
          for (int k = 0; k < 6; k++) gsum[k] += gl[k];
        }
#pragma unroll
        for (int k = 0; k < 6; k++) st[ELEM_K + k] = gsum[k];
#pragma unroll
        for (int k = 0; k < 3; k++) {
          slots[lane * 6 + k] = a.node_slot[3 * (m.n0 + nd.x) + k];
          slots[lane * 6 + 3 + k] = a.node_slot[3 * (m.n0 + nd.y) + k];
        }
      }
      __syncwarp();
      int pa = 0, pb = 0;
      if (lane < ELEM_K) cid_pair(lane, pa, pb);
      for (int ee = 0; ee < cnt; ee++) {
        const int32_t* sl = slots + ee * 6;
        if (lane < ELEM_K) {
          const int32_t fi = sl[pa], fj = sl[pb];
          if (fi >= 0 && fj >= 0) {
            const int i = max(fi, fj), j = min(fi, fj);
            band[i * W + (j - i + hb)] += stage[ee * STAGE_D + lane];
          }
        } else if (lane < STAGE_D) {
          const int32_t t = sl[lane - ELEM_K];
          const double v = stage[ee * STAGE_D + lane];
          if (t >= 0) Ff[t] += v;
          else if (t != SLOT_NONE) Eq[-t - 1] += v;
        }
        __syncwarp();
      }
    }
    if (a.F_f)
      for (int i = lane; i < nf; i += 32) a.F_f[c * a.nf_tot + m.f0 + i] = Ff[i];
    // ---- stage B: banded LDL^T (row i holds columns i-hb..i at band[i*W + (j-i+hb)]) ----
    int info = 0;
    for (int j = 0; j < nf; j++) {
      const double d = band[j * W + hb];
      if (!(d > 0.0)) {  // Eigen LLT stops at the first x <= 0 (solver_static.cpp:97-105)
        info = j + 1;
        break;
      }
      const double inv = 1.0 / d;
      const int imax = min(j + hb, nf - 1);
      // trailing update with the still-unscaled column j: B(i,k) -= B(i,j)*B(k,j)/d, j < k <= i
      for (int i = j + 1; i <= imax; i++) {
        const double lij = band[i * W + (j - i + hb)] * inv;
        for (int k = j + 1 + lane; k <= i; k += 32) band[i * W + (k - i + hb)] -= lij * band[k * W + (j - k + hb)];
      }
      __syncwarp();
      for (int i = j + 1 + lane; i <= imax; i += 32) band[i * W + (j - i + hb)] *= inv;  // store L
      __syncwarp();
    }
    if (info == 0) {
      // L y = b
      for (int i = 1; i < nf; i++) {
        const int k0 = max(0, i - hb);
        double acc = 0.0;
        for (int k = k0 + lane; k < i; k += 32) acc += band[i * W + (k - i + hb)] * Ff[k];
        acc = warp_sum(acc);
        if (lane == 0) Ff[i] -= acc;
        __syncwarp();
      }
      for (int i = lane; i < nf; i += 32) Ff[i] /= band[i * W + hb];
      __syncwarp();
      // L^T x = z
      for (int i = nf - 2; i >= 0; i--) {
        const int k1 = min(nf - 1, i + hb);
        double acc = 0.0;
        for (int k = i + 1 + lane; k <= k1; k += 32) acc += band[k * W + (i - k + hb)] * Ff[k];
        acc = warp_sum(acc);
        if (lane == 0) Ff[i] -= acc;
        __syncwarp();
      }
    }
    if (lane == 0) {
      a.info[sys] = info;
      if (info) atomicOr(a.fail, 1);
    }
    for (int i = lane; i < nf; i += 32) a.U_f[c * a.nf_tot + m.f0 + i] = Ff[i];
    // ---- stage C: element end forces, reactions ----
    double* sst = stage;               // [32][12] stress staging
    double* sgf = stage + 32 * 12;     // [32][6]  global end forces
    for (int eb = 0; eb < m.E; eb += 32) {
      const int cnt = min(32, m.E - eb);
      if (lane < cnt) {
        const int64_t e = m.e0 + eb + lane;
        const uint2 nd = a.elem_nodes[e];
        const int type = a.elem_type[e];
        const double2 p0 = a.node_xy[m.n0 + nd.x], p1 = a.node_xy[m.n0 + nd.y];
        const ElemGeom g = elem_geom(p0.x, p0.y, p1.x, p1.y);
        const ElemLocalK lk = elem_local_k(type, eai_base[3 * e], eai_base[3 * e + 1], eai_base[3 * e + 2], g);
        double ug[6], es[6];
#pragma unroll
        for (int k = 0; k < 3; k++) {
          const int32_t t0 = a.node_slot[3 * (m.n0 + nd.x) + k], t1 = a.node_slot[3 * (m.n0 + nd.y) + k];
          slots[lane * 6 + k] = t0;
          slots[lane * 6 + 3 + k] = t1;
          ug[k] = t0 >= 0 ? Ff[t0] : 0.0;
          ug[3 + k] = t1 >= 0 ? Ff[t1] : 0.0;
        }
        elem_end_forces(lk, g, ug, es);
        sgf[lane * 6 + 0] = g.c * es[0] - g.s * es[1];
        sgf[lane * 6 + 1] = g.s * es[0] + g.c * es[1];
        sgf[lane * 6 + 2] = es[2];
        sgf[lane * 6 + 3] = g.c * es[3] - g.s * es[4];
        sgf[lane * 6 + 4] = g.s * es[3] + g.c * es[4];
        sgf[lane * 6 + 5] = es[5];
        if (a.stress) {
          double ext[6] = {0, 0, 0, 0, 0, 0};
          for (int j = a.el_ptr[e]; j < a.el_ptr[e + 1]; j++) {
            const int sl = a.el_order[j];
            double gl[6], ex[6];
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
      for (int ee = 0; ee < cnt; ee++) {
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
  }
}

cudaError_t launch_frame_fused(const FrameArgs& a, int sm_count, cudaStream_t st) {
  const size_t per_warp = frame_smem_per_warp(a.max_band, a.max_nf, a.max_nb);
  const size_t smem = per_warp * FRAME_WARPS;
  if (smem > 227 * 1024) return cudaErrorInvalidConfiguration;
  cudaError_t e = cudaFuncSetAttribute(k_frame_fused, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
  if (e != cudaSuccess) return e;
  int per_sm = 1;
  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_frame_fused, FRAME_WARPS * 32, smem);
  if (e != cudaSuccess) return e;
  if (per_sm < 1) per_sm = 1;
  const int64_t n_sys = a.S * a.L;
  int64_t grid = int64_t(sm_count) * per_sm;
  const int64_t need = (n_sys + FRAME_WARPS - 1) / FRAME_WARPS;
  if (grid > need) grid = need;
  if (grid < 1) grid = 1;
  k_frame_fused<<<unsigned(grid), FRAME_WARPS * 32, smem, st>>>(a, int(per_warp / 8));
  return cudaGetLastError();
}

}  // namespace obgpu
