// beam_math.cuh — device-side element arithmetic shared by every kernel of libobgpu (FP64).
//
// What the reference computes with virtual calls, heap vectors and dense 6x6 Eigen products
//   orientation        src/internals.cpp:30-123            (2D: R = [[c,-s,0],[s,c,0],[0,0,1]])
//   local stiffness    src/CElementBeam_2D_{RR,RA,AR,AA}.cpp:41-127
//   rotation to global src/CElement.cpp:53-79              (B <- R B R^T per non-zero 3x3 sub-block)
//   equivalent loads   src/CLoadDistributedUniform.cpp:33-159
//   end forces         src/CFiniteElementProblem_postproc_stress.cpp:86-165
// is written here in closed form over the only DOFs a planar beam touches (DX,DY,RZ per end).
// Results agree with the reference to rounding (<= a few ulp; the contract is 1e-9 relative).
#pragma once
#include "obgpu_internal.h"

namespace obgpu {

struct ElemGeom {
  double c, s, L, invL;
};

// internals.cpp:35-52 : len = |p2-p1|, X = p12 * (1/len)
__device__ __forceinline__ ElemGeom elem_geom(double x0, double y0, double x1, double y1) {
  ElemGeom g;
  const double dx = x1 - x0, dy = y1 - y0;
  const double L2 = dx * dx + dy * dy;
  g.L = sqrt(L2);
  if (L2 == 0.0) {  // internals.cpp:100-104: coincident nodes (springs) -> identity orientation
    g.invL = 0.0; g.c = 1.0; g.s = 0.0;
    return g;
  }
  g.invL = 1.0 / g.L;
  g.c = dx * g.invL;
  g.s = dy * g.invL;
  return g;
}

// Non-zero coefficients of the local blocks (all four element types):
//   K11 = [[a,0,0],[0,b,p11],[0,p11,r11]]   K22 = [[a,0,0],[0,b,p22],[0,p22,r22]]
//   K12 = [[-a,0,0],[0,-b,q15],[0,q51,q55]]          (rows/cols = local x, y, rz)
struct ElemLocalK {
  double a, b, p11, r11, p22, r22, q15, q51, q55;
};

__device__ __forceinline__ ElemLocalK elem_local_k(int type, double E, double A, double Iz, const ElemGeom& g) {
  ElemLocalK k;
  const double EI = E * Iz;
  const double EI_L = EI * g.invL, EI_L2 = EI_L * g.invL, EI_L3 = EI_L2 * g.invL;
  k.a = E * A * g.invL;
  k.p11 = k.r11 = k.p22 = k.r22 = k.q15 = k.q51 = k.q55 = 0.0;
  if (type == OBGPU_BEAM2D_RR) {  // CElementBeam_2D_RR.cpp:104-126
    k.b = 12.0 * EI_L3;
    k.p11 = 6.0 * EI_L2;
    k.r11 = 4.0 * EI_L;
    k.p22 = -k.p11;
    k.r22 = k.r11;
    k.q15 = k.p11;
    k.q51 = -k.p11;
    k.q55 = 2.0 * EI_L;
  } else if (type == OBGPU_BEAM2D_RA) {  // CElementBeam_2D_RA.cpp:92-107
    k.b = 3.0 * EI_L3;
    k.p11 = 3.0 * EI_L2;
    k.r11 = 3.0 * EI_L;
    k.q51 = -k.p11;
  } else if (type == OBGPU_BEAM2D_AR) {  // CElementBeam_2D_AR.cpp:92-107
    k.b = 3.0 * EI_L3;
    k.p22 = -3.0 * EI_L2;
    k.r22 = 3.0 * EI_L;
    k.q15 = -k.p22;
  } else if (type == OBGPU_BEAM2D_AA) {  // CElementBeam_2D_AA.cpp:86-94
    k.b = 0.0;
  } else if (type == OBGPU_BEAM2D_RD) {  // CElementBeam_2D_RD.cpp:96-112
    k.b = 0.0;
    k.r11 = k.r22 = EI_L;
    k.q55 = -EI_L;
  } else {  // springs: (E, A, Iz) carry (K | Kx, Ky, KRz)
    k.a = k.b = 0.0;
    if (type == OBGPU_SPRING_1D) {  // CElementSpring.cpp:56-66
      k.a = E;
    } else if (type == OBGPU_SPRING_XY) {  // CElementSpringXY.cpp:47-59
      k.a = E; k.b = A;
    } else if (type == OBGPU_SPRING_DXDYRZ) {
      // CElementSpringDXDYRZ.cpp:47-62: the last line there assigns K11(5,5) = -KRz (not K12): K11(5,5) = -KRz,
      // K22(5,5) = +KRz (copied before), K12(5,5) = 0.  Reproduced as is.
      k.a = E; k.b = A; k.r11 = -Iz; k.r22 = Iz;
    } else {  // OBGPU_SPRING_TORSION, CElementTorsionSpring.cpp:57-67
      k.r11 = k.r22 = E;
      k.q55 = -E;
    }
  }
  return k;
}

// Global blocks in the compact ELEM_K order (see obgpu_internal.h): R B R^T with R the z-rotation (c,s).
__device__ __forceinline__ void elem_global_k(const ElemLocalK& k, const ElemGeom& g, double ke[ELEM_K]) {
  const double c = g.c, s = g.s;
  const double cc = c * c, ss = s * s, cs = c * s;
  const double g00 = k.a * cc + k.b * ss, g01 = (k.a - k.b) * cs, g11 = k.a * ss + k.b * cc;
  ke[0] = g00;  ke[1] = g01;  ke[2] = -s * k.p11;  ke[3] = g11;  ke[4] = c * k.p11;  ke[5] = k.r11;
  ke[6] = g00;  ke[7] = g01;  ke[8] = -s * k.p22;  ke[9] = g11;  ke[10] = c * k.p22; ke[11] = k.r22;
  ke[12] = -g00; ke[13] = -g01; ke[14] = -s * k.q15;
  ke[15] = -g01; ke[16] = -g11; ke[17] = c * k.q15;
  ke[18] = -s * k.q51; ke[19] = c * k.q51; ke[20] = k.q55;
}

// CLoadDistributedUniform.cpp:64-71: projections of the load direction on the beam axis and its normal
// (case-invariant: depend on geometry and direction only)
struct LoadProj {
  double plong, pperp;
};
__device__ __forceinline__ LoadProj load_proj(const ElemGeom& g, double dx, double dy) {
  LoadProj p;
  p.plong = g.c * dx + g.s * dy;                 // Az = 0 in plane
  const double cross_z = g.c * dy - g.s * dx;
  p.pperp = sqrt(1.0 - p.plong * p.plong) * (cross_z < 0 ? 1.0 : -1.0);
  return p;
}

// CLoadDistributedUniform.cpp:33-159.  Outputs:
//   gl[6]  equivalent nodal loads in global axes (Fx,Fy,Mz at end 0; Fx,Fy,Mz at end 1)
//   ex[6]  fixed-end stress to add in postProcCalcStress (N,Vy,Mz at face 0; N,Vy,Mz at face 1)
__device__ __forceinline__ void uniform_load_p(int type, const ElemGeom& g, double q, const LoadProj& pr, double gl[6],
                                               double ex[6]) {
  const double L = g.L;
  const double q_t = q * pr.plong, q_n = q * pr.pperp;
  const double fx = -q_t * L * 0.5;
  double f1y, f2y, m1, m2;
  if (type == OBGPU_BEAM2D_AA) {
    f1y = q_n * L * 0.5; f2y = f1y; m1 = 0.0; m2 = 0.0;
  } else if (type == OBGPU_BEAM2D_RR) {
    f1y = q_n * L * 0.5; f2y = f1y; m1 = q_n * L * L / 12.0; m2 = -m1;
  } else if (type == OBGPU_BEAM2D_RA) {
    f1y = q_n * L * (5.0 / 8.0); f2y = q_n * L * (3.0 / 8.0); m1 = q_n * L * L / 8.0; m2 = 0.0;
  } else {
    f1y = q_n * L * (3.0 / 8.0); f2y = q_n * L * (5.0 / 8.0); m1 = 0.0; m2 = -(q_n * L * L / 8.0);
  }
  ex[0] = -fx; ex[1] = -f1y; ex[2] = -m1;
  ex[3] = fx;  ex[4] = f2y;  ex[5] = m2;
  gl[0] = -(g.c * fx - g.s * f1y); gl[1] = -(g.s * fx + g.c * f1y); gl[2] = -m1;
  gl[3] = -(g.c * fx - g.s * f2y); gl[4] = -(g.s * fx + g.c * f2y); gl[5] = -m2;
}
__device__ __forceinline__ void uniform_load(int type, const ElemGeom& g, double q, double dx, double dy, double gl[6],
                                             double ex[6]) {
  uniform_load_p(type, g, q, load_proj(g, dx, dy), gl, ex);
}

// Every element load the reference has, same outputs as uniform_load_p:
//   OBGPU_LOAD_UNIFORM      p1 = q                       CLoadDistributedUniform.cpp:33-159 (all four beam types)
//   OBGPU_LOAD_TRIANGULAR   p1 = q_ini, p2 = q_end       CLoadDistributedTriangular.cpp:31-169 (RR only)
//   OBGPU_LOAD_CONCENTRATED p1 = P, p2 = distance from end 0   CLoadConcentratedForce.cpp:31-154 (RR, perpendicular only)
//   OBGPU_LOAD_TEMPERATURE  p1 = temperature increment   CLoadConstTemperature.cpp:31-93 (alpha = 12e-6 hard-wired there)
// EA = E*A of the element (temperature only).  Unsupported (load, element) pairs are refused by the host at upload,
// like the reference throws.
__device__ __forceinline__ void elem_load_p(int ltype, int etype, const ElemGeom& g, double p1, double p2, const LoadProj& pr,
                                            double EA, double gl[6], double ex[6]) {
  if (ltype == OBGPU_LOAD_UNIFORM) {
    uniform_load_p(etype, g, p1, pr, gl, ex);
    return;
  }
  const double L = g.L;
  double f1x = 0.0, f2x = 0.0, f1y = 0.0, f2y = 0.0, m1 = 0.0, m2 = 0.0;
  if (ltype == OBGPU_LOAD_TEMPERATURE) {
    const double N = EA * 12e-6 * p1;  // stress N = -N at both faces; loads -N*axis at end 0, +N*axis at end 1
    ex[0] = -N; ex[1] = 0.0; ex[2] = 0.0;
    ex[3] = -N; ex[4] = 0.0; ex[5] = 0.0;
    gl[0] = -N * g.c; gl[1] = -N * g.s; gl[2] = 0.0;
    gl[3] = N * g.c;  gl[4] = N * g.s;  gl[5] = 0.0;
    return;
  }
  if (ltype == OBGPU_LOAD_TRIANGULAR) {
    const double q0_t = p1 * pr.plong, q1_t = p2 * pr.plong, q0_n = p1 * pr.pperp, q1_n = p2 * pr.pperp;
    f1y = L * (q0_n * 0.5 + (q1_n - q0_n) * 3. / 20.);
    m1 = +L * L * (q0_n * 1. / 12. + (q1_n - q0_n) * 1. / 30.);
    f2y = L * (q0_n * 0.5 + (q1_n - q0_n) * 7. / 20.);
    m2 = -L * L * (q0_n * 1. / 12. + (q1_n - q0_n) * 1. / 20.);
    f1x = -0.5 * L * (q0_t + (q1_t - q0_t) / 3.0);
    f2x = -0.5 * L * (q0_t + 2.0 * (q1_t - q0_t) / 3.0);
  } else {  // OBGPU_LOAD_CONCENTRATED
    const double P_n = p1 * pr.pperp;
    const double a = p2, b = L - a, a2 = a * a, b2 = b * b, L2 = L * L, L3 = L * L2;
    f1y = P_n * (b2 / L3) * (3 * a + b);
    m1 = P_n * a * b2 / L2;
    f2y = P_n * (a2 / L3) * (a + 3 * b);
    m2 = -P_n * a2 * b / L2;
  }
  ex[0] = -f1x; ex[1] = -f1y; ex[2] = -m1;
  ex[3] = f2x;  ex[4] = f2y;  ex[5] = m2;
  gl[0] = -(g.c * f1x - g.s * f1y); gl[1] = -(g.s * f1x + g.c * f1y); gl[2] = -m1;
  gl[3] = -(g.c * f2x - g.s * f2y); gl[4] = -(g.s * f2x + g.c * f2y); gl[5] = -m2;
}
__device__ __forceinline__ void elem_load(int ltype, int etype, const ElemGeom& g, double p1, double p2, double dx, double dy,
                                          double EA, double gl[6], double ex[6]) {
  LoadProj pr;
  pr.plong = 0.0; pr.pperp = 0.0;
  if (ltype != OBGPU_LOAD_TEMPERATURE) pr = load_proj(g, dx, dy);
  elem_load_p(ltype, etype, g, p1, p2, pr, EA, gl, ex);
}

// postproc_stress.cpp:86-146 : raw end forces es = K_local * U_local (before the face-0 sign flip).
// ug = global (ux,uy,rz) of end 0 then end 1.  es = (N,Vy,Mz) at end 0 then end 1.
__device__ __forceinline__ void elem_end_forces(const ElemLocalK& k, const ElemGeom& g, const double ug[6], double es[6]) {
  const double u0x = g.c * ug[0] + g.s * ug[1], u0y = -g.s * ug[0] + g.c * ug[1], t0 = ug[2];
  const double u1x = g.c * ug[3] + g.s * ug[4], u1y = -g.s * ug[3] + g.c * ug[4], t1 = ug[5];
  es[0] = k.a * (u0x - u1x);
  es[1] = k.b * (u0y - u1y) + k.p11 * t0 + k.q15 * t1;
  es[2] = k.p11 * u0y + k.r11 * t0 + k.q51 * u1y + k.q55 * t1;
  es[3] = k.a * (u1x - u0x);
  es[4] = k.b * (u1y - u0y) + k.p22 * t1 + k.q51 * t0;
  es[5] = k.p22 * u1y + k.r22 * t1 + k.q15 * u0y + k.q55 * t0;
}

// ---- general path: nodal rotated frames and non-zero prescribed displacements ----------------------
// Compact element block -> full symmetric 6x6 over (end0 x,y,rz, end1 x,y,rz).
__device__ __forceinline__ double ke_at(const double* ke, int a, int b) {
  if (a > b) { const int t = a; a = b; b = t; }
  if (b < 3) return ke[a == 0 ? b : a == 1 ? 2 + b : 5];
  if (a >= 3) return ke[6 + (a == 3 ? b - 3 : a == 4 ? b - 1 : 5)];
  return ke[12 + 3 * a + (b - 3)];
}
// assemble.cpp:144-179: K_ij <- T_i^T K_ij T_j on every 3x3 sub-block, T = nodal frame (a rotation about z by the
// node's yaw: Q = [[c,-s,0],[s,c,0],[0,0,1]] on (x,y,rz)).  (ci,si)/(cj,sj) = cos/sin of end 0 / end 1.
__device__ __forceinline__ void rotate_nodal(double* ke, double ci, double si, double cj, double sj) {
  double B[3][3][3];  // [block K11,K22,K12][r][c]
#pragma unroll
  for (int r = 0; r < 3; r++)
#pragma unroll
    for (int c = 0; c < 3; c++) {
      B[0][r][c] = ke_at(ke, r, c);
      B[1][r][c] = ke_at(ke, 3 + r, 3 + c);
      B[2][r][c] = ke_at(ke, r, 3 + c);
    }
#pragma unroll
  for (int k = 0; k < 3; k++) {
    const double cl = k == 1 ? cj : ci, sl = k == 1 ? sj : si;  // left factor: node of the block's rows
    const double cr = k == 0 ? ci : cj, sr = k == 0 ? si : sj;  // right factor: node of the block's columns
#pragma unroll
    for (int c = 0; c < 3; c++) {  // Q_l^T B
      const double b0 = B[k][0][c], b1 = B[k][1][c];
      B[k][0][c] = cl * b0 + sl * b1;
      B[k][1][c] = -sl * b0 + cl * b1;
    }
#pragma unroll
    for (int r = 0; r < 3; r++) {  // B Q_r
      const double b0 = B[k][r][0], b1 = B[k][r][1];
      B[k][r][0] = cr * b0 + sr * b1;
      B[k][r][1] = -sr * b0 + cr * b1;
    }
  }
  ke[0] = B[0][0][0]; ke[1] = B[0][0][1]; ke[2] = B[0][0][2]; ke[3] = B[0][1][1]; ke[4] = B[0][1][2]; ke[5] = B[0][2][2];
  ke[6] = B[1][0][0]; ke[7] = B[1][0][1]; ke[8] = B[1][0][2]; ke[9] = B[1][1][1]; ke[10] = B[1][1][2]; ke[11] = B[1][2][2];
#pragma unroll
  for (int r = 0; r < 3; r++)
#pragma unroll
    for (int c = 0; c < 3; c++) ke[12 + 3 * r + c] = B[2][r][c];
}

}  // namespace obgpu
