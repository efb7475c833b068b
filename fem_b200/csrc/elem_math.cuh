// elem_math.cuh -- fp64 device element library (shape derivatives, Jacobians, the
// (lambda, mu) node-block form of B^T D B, stress at integration points, Xu-Needleman).
//
// What it computes is what src/m_elems.F90 / src/m_local.F90 / src/m_seplaw.F90 compute;
// how is new: no B or D matrix is materialised.  With g_a = grad N_a (physical),
//   K_ab = sum_ip w detJ [ lam g_a (x) g_b + mu g_b (x) g_a + mu (g_a . g_b) I ]
// equals B_a^T D B_b for the reference's plane-strain / 3-D isotropic D
// (src/m_local.F90:204-228: lam = c nu, mu = c (1-2nu)/2, c = E/((1+nu)(1-2nu))) and its
// Voigt order xx,yy,(zz),xy,(yz,zx) with engineering shear (src/m_local.F90:148-171).
#pragma once
#include "femx_internal.h"

namespace femx {

template <int T> struct ET;
template <> struct ET<FEMX_TRI> { static constexpr int NEN = 3, DIM = 2, NIP = 1, CP = 3; };
template <> struct ET<FEMX_QUA> { static constexpr int NEN = 4, DIM = 2, NIP = 4, CP = 3; };
template <> struct ET<FEMX_TET> { static constexpr int NEN = 4, DIM = 3, NIP = 1, CP = 6; };
template <> struct ET<FEMX_HEX> { static constexpr int NEN = 8, DIM = 3, NIP = 8, CP = 6; };
template <> struct ET<FEMX_COH> { static constexpr int NEN = 4, DIM = 2, NIP = 2, CP = 3; };

// sqrt(1/3) as the reference evaluates it in real(8) (src/m_elems.F90:311, :500)
#define FEMX_G 0.5773502691896257

__host__ __device__ inline void lame(double E, double nu, double *lam, double *mu) {
  const double c = E / ((1.0 + nu) * (1.0 - 2.0 * nu));  // DMat2d/3d, src/m_local.F90:207,217
  *lam = c * nu;
  *mu = c * ((1.0 - 2.0 * nu) / 2.0);
}

// Parent-space derivatives dN[d][a] at integration point ip, and its weight.
// ShapeFuncd* src/m_elems.F90:273-279, 357-363, 459-466, 569-580; SamPts* :212-235,
// :298-316 (ip order (-,-),(-,+),(+,-),(+,+)), :383-409, :487-509 (ip k beside node k).
template <int T>
__device__ __forceinline__ double parent_grad(int ip, double (&dN)[ET<T>::DIM][ET<T>::NEN]) {
  if constexpr (T == FEMX_TRI) {
    dN[0][0] = -1.0; dN[0][1] = 1.0; dN[0][2] = 0.0;
    dN[1][0] = -1.0; dN[1][1] = 0.0; dN[1][2] = 1.0;
    return 0.5;
  } else if constexpr (T == FEMX_TET) {
    dN[0][0] = -1.0; dN[0][1] = 1.0; dN[0][2] = 0.0; dN[0][3] = 0.0;
    dN[1][0] = -1.0; dN[1][1] = 0.0; dN[1][2] = 1.0; dN[1][3] = 0.0;
    dN[2][0] = -1.0; dN[2][1] = 0.0; dN[2][2] = 0.0; dN[2][3] = 1.0;
    return 1.0 / 6.0;
  } else if constexpr (T == FEMX_QUA) {
    const double e = (ip & 2) ? FEMX_G : -FEMX_G, n = (ip & 1) ? FEMX_G : -FEMX_G;
    dN[0][0] = -0.25 * (1.0 - n); dN[0][1] = 0.25 * (1.0 - n); dN[0][2] = 0.25 * (1.0 + n); dN[0][3] = -0.25 * (1.0 + n);
    dN[1][0] = -0.25 * (1.0 - e); dN[1][1] = -0.25 * (1.0 + e); dN[1][2] = 0.25 * (1.0 + e); dN[1][3] = 0.25 * (1.0 - e);
    return 1.0;
  } else {
    static_assert(T == FEMX_HEX, "no parent gradient for this type");
    // node / ip sign table: bottom face CCW then top face CCW
    const bool px = (ip == 1) | (ip == 2) | (ip == 5) | (ip == 6);
    const bool py = (ip == 2) | (ip == 3) | (ip == 6) | (ip == 7);
    const bool pz = ip >= 4;
    const double e = px ? FEMX_G : -FEMX_G, n = py ? FEMX_G : -FEMX_G, s = pz ? FEMX_G : -FEMX_G;
    const double em = 1.0 - e, ep = 1.0 + e, nm = 1.0 - n, np = 1.0 + n, sm = 1.0 - s, sp = 1.0 + s, c = 0.125;
    dN[0][0] = -c * nm * sm; dN[0][1] = c * nm * sm; dN[0][2] = c * np * sm; dN[0][3] = -c * np * sm;
    dN[0][4] = -c * nm * sp; dN[0][5] = c * nm * sp; dN[0][6] = c * np * sp; dN[0][7] = -c * np * sp;
    dN[1][0] = -c * em * sm; dN[1][1] = -c * ep * sm; dN[1][2] = c * ep * sm; dN[1][3] = c * em * sm;
    dN[1][4] = -c * em * sp; dN[1][5] = -c * ep * sp; dN[1][6] = c * ep * sp; dN[1][7] = c * em * sp;
    dN[2][0] = -c * em * nm; dN[2][1] = -c * ep * nm; dN[2][2] = -c * ep * np; dN[2][3] = -c * em * np;
    dN[2][4] = c * em * nm; dN[2][5] = c * ep * nm; dN[2][6] = c * ep * np; dN[2][7] = c * em * np;
    return 1.0;
  }
}

// Physical gradients g[a][d] at ip and w*detJ.  FormdNdetJ src/m_local.F90:175-189 with the
// LAPACK inverse (src/m_utils.F90:45-57) replaced by the closed-form adjugate.
template <int T>
__device__ __forceinline__ double phys_grad(const double (&X)[ET<T>::NEN][ET<T>::DIM], int ip,
                                            double (&g)[ET<T>::NEN][ET<T>::DIM]) {
  constexpr int NEN = ET<T>::NEN, DIM = ET<T>::DIM;
  double dN[DIM][NEN];
  const double w = parent_grad<T>(ip, dN);
  double J[DIM][DIM];
#pragma unroll
  for (int i = 0; i < DIM; ++i)
#pragma unroll
    for (int j = 0; j < DIM; ++j) {
      double s = 0.0;
#pragma unroll
      for (int a = 0; a < NEN; ++a) s += dN[i][a] * X[a][j];
      J[i][j] = s;
    }
  double det, inv[DIM][DIM];
  if constexpr (DIM == 2) {
    det = J[0][0] * J[1][1] - J[0][1] * J[1][0];
    const double r = 1.0 / det;
    inv[0][0] = J[1][1] * r; inv[0][1] = -J[0][1] * r;
    inv[1][0] = -J[1][0] * r; inv[1][1] = J[0][0] * r;
  } else {
    const double c00 = J[1][1] * J[2][2] - J[1][2] * J[2][1];
    const double c01 = J[1][0] * J[2][2] - J[1][2] * J[2][0];
    const double c02 = J[1][0] * J[2][1] - J[1][1] * J[2][0];
    det = J[0][0] * c00 - J[0][1] * c01 + J[0][2] * c02;  // MatDet src/m_utils.F90:63-65
    const double r = 1.0 / det;
    inv[0][0] = c00 * r;
    inv[0][1] = (J[0][2] * J[2][1] - J[0][1] * J[2][2]) * r;
    inv[0][2] = (J[0][1] * J[1][2] - J[0][2] * J[1][1]) * r;
    inv[1][0] = -c01 * r;
    inv[1][1] = (J[0][0] * J[2][2] - J[0][2] * J[2][0]) * r;
    inv[1][2] = (J[0][2] * J[1][0] - J[0][0] * J[1][2]) * r;
    inv[2][0] = c02 * r;
    inv[2][1] = (J[0][1] * J[2][0] - J[0][0] * J[2][1]) * r;
    inv[2][2] = (J[0][0] * J[1][1] - J[0][1] * J[1][0]) * r;
  }
#pragma unroll
  for (int a = 0; a < NEN; ++a)
#pragma unroll
    for (int i = 0; i < DIM; ++i) {
      double s = 0.0;
#pragma unroll
      for (int j = 0; j < DIM; ++j) s += inv[i][j] * dN[j][a];
      g[a][i] = s;
    }
  return w * det;
}

// Row-block a of Ke: acc[b][i*DIM+j] = K[(a,i),(b,j)].  FormElKE src/m_local.F90:61-65.
template <int T>
__device__ __forceinline__ void ke_row_blocks(const double (&X)[ET<T>::NEN][ET<T>::DIM], int a, double lam,
                                              double mu, double (&acc)[ET<T>::NEN][ET<T>::DIM * ET<T>::DIM]) {
  constexpr int NEN = ET<T>::NEN, DIM = ET<T>::DIM, NIP = ET<T>::NIP;
#pragma unroll
  for (int b = 0; b < NEN; ++b)
#pragma unroll
    for (int q = 0; q < DIM * DIM; ++q) acc[b][q] = 0.0;
#pragma unroll 1
  for (int ip = 0; ip < NIP; ++ip) {
    double g[NEN][DIM];
    const double wd = phys_grad<T>(X, ip, g);
    double ga[DIM];
#pragma unroll
    for (int i = 0; i < DIM; ++i) ga[i] = 0.0;
#pragma unroll
    for (int b = 0; b < NEN; ++b)
      if (b == a) {
#pragma unroll
        for (int i = 0; i < DIM; ++i) ga[i] = g[b][i];
      }
    double la[DIM], ma[DIM];
#pragma unroll
    for (int i = 0; i < DIM; ++i) { la[i] = wd * lam * ga[i]; ma[i] = wd * mu * ga[i]; }
#pragma unroll
    for (int b = 0; b < NEN; ++b) {
      double dot = 0.0;
#pragma unroll
      for (int i = 0; i < DIM; ++i) dot += ma[i] * g[b][i];
#pragma unroll
      for (int i = 0; i < DIM; ++i)
#pragma unroll
        for (int j = 0; j < DIM; ++j) {
          double v = la[i] * g[b][j] + ma[j] * g[b][i];
          if (i == j) v += dot;
          acc[b][i * DIM + j] += v;
        }
    }
  }
}

// Stress at the integration points: CalcElStress src/m_local.F90:93-123.
template <int T>
__device__ __forceinline__ void ip_stress(const double (&X)[ET<T>::NEN][ET<T>::DIM],
                                          const double (&U)[ET<T>::NEN][ET<T>::DIM], double lam, double mu,
                                          double (&sig)[ET<T>::NIP][ET<T>::CP]) {
  constexpr int NEN = ET<T>::NEN, DIM = ET<T>::DIM, NIP = ET<T>::NIP;
#pragma unroll 1
  for (int ip = 0; ip < NIP; ++ip) {
    double g[NEN][DIM];
    phys_grad<T>(X, ip, g);
    double H[DIM][DIM];  // H[i][j] = du_i/dx_j
#pragma unroll
    for (int i = 0; i < DIM; ++i)
#pragma unroll
      for (int j = 0; j < DIM; ++j) {
        double s = 0.0;
#pragma unroll
        for (int a = 0; a < NEN; ++a) s += U[a][i] * g[a][j];
        H[i][j] = s;
      }
    double tr = 0.0;
#pragma unroll
    for (int i = 0; i < DIM; ++i) tr += H[i][i];
    double out[ET<T>::CP];
#pragma unroll
    for (int i = 0; i < DIM; ++i) out[i] = lam * tr + 2.0 * mu * H[i][i];
    if constexpr (DIM == 2) {
      out[2] = mu * (H[0][1] + H[1][0]);
    } else {
      out[3] = mu * (H[0][1] + H[1][0]);  // xy
      out[4] = mu * (H[1][2] + H[2][1]);  // yz
      out[5] = mu * (H[2][0] + H[0][2]);  // zx
    }
#pragma unroll
    for (int c = 0; c < ET<T>::CP; ++c)
#pragma unroll
      for (int q = 0; q < NIP; ++q)
        if (q == ip) sig[q][c] = out[c];
  }
}

// ---- Xu-Needleman (src/m_seplaw.F90:15-53, :57-94); props = smax, dn, dt, q, r, zeta ----
__device__ __forceinline__ void seplaw_1(const double *p, double gn, double gt, double vgn, double dtime,
                                         double &Tn, double &Tt, double (&S)[2][2]) {
  const double E1 = 2.718281828459045;  // dexp(1.d0)
  const double sw = E1 * p[0] * p[1], dn = p[1], dt = p[2], q = p[3], r = p[4], zeta = p[5];
  const double en = exp(-gn / dn), et = exp(-gt * gt / (dt * dt));
  const double x = gn / dn;
  Tn = (sw / dn) * en * (x * et + (1.0 - et) * (1.0 - q) / (r - 1.0) * (r - x)) + zeta * p[0] * vgn / dn;
  Tt = (q + (r - q) / (r - 1.0) * x) * en * et * 2.0 * (dn / dt) * (sw / dn) * gt / dt;
  double c1 = (1.0 - q) / (r - 1.0) * (1.0 - et) * (r + 1.0 - x);
  c1 = (1.0 - x) * et - c1;
  S[0][0] = (sw / (dn * dn)) * en * c1 + zeta * p[0] / dn / dtime;
  c1 = 2.0 * (sw / (dt * dt)) * ((q + x * (r - q) / (r - 1.0)) * en * et);
  S[1][1] = c1 * (1.0 - 2.0 * (gt * gt / (dt * dt)));
  c1 = 2.0 * (sw / (dt * dn)) * ((-x + (1.0 - q) / (r - 1.0) * (r - x)) * en * et);
  S[0][1] = (gt / dt) * c1;
  S[1][0] = S[0][1];
}

// Cohesive element: nodal force (applyTract_1 src/m_global.F90:622-682) and 8x8 tangent
// (applyStiff_1 :762-845) with kinematics of src/m_elems.F90:658-720.  X, U are the four
// nodes' coordinates and (quirk-selected) displacements.
__device__ __forceinline__ void cohesive_element(const double (&X)[4][2], const double (&U)[4][2],
                                                 const double *props, double dt, int quirks, double *force,
                                                 double *stiff) {
  double t[2] = {0.5 * (X[1][0] - X[0][0] + X[2][0] - X[3][0]), 0.5 * (X[1][1] - X[0][1] + X[2][1] - X[3][1])};
  const double mag = sqrt(t[0] * t[0] + t[1] * t[1]);
  t[0] /= mag; t[1] /= mag;
  const double det = mag * 0.5;
  const double nrm[2] = {-t[1], t[0]};
  const double CN = 0.5773502691896260;  // SamPtsCoh src/m_elems.F90:608
  if (force)
    for (int i = 0; i < 8; ++i) force[i] = 0.0;
  if (stiff)
    for (int i = 0; i < 64; ++i) stiff[i] = 0.0;
  for (int ip = 0; ip < 2; ++ip) {
    const double xi = ip ? CN : -CN;
    const double N[4] = {0.5 * (1.0 - xi), 0.5 * (1.0 + xi), 0.5 * (1.0 + xi), 0.5 * (1.0 - xi)};
    double ur[2] = {0.0, 0.0};
    for (int a = 0; a < 2; ++a)
      for (int d = 0; d < 2; ++d) ur[d] = ur[d] + N[a + 2] * U[a + 2][d] - N[a] * U[a][d];
    const double gn = nrm[0] * ur[0] + nrm[1] * ur[1], gt = t[0] * ur[0] + t[1] * ur[1];
    const double vgn = nrm[0] * (ur[0] / dt) + nrm[1] * (ur[1] / dt);
    double Tn, Tt, S[2][2];
    seplaw_1(props, gn, gt, vgn, dt, Tn, Tt, S);
    for (int a = 0; a < 4; ++a) {
      const double sa = a < 2 ? 1.0 : -1.0;
      if (force)
        for (int d = 0; d < 2; ++d) {
          const double v = sa * N[a] * (Tn * nrm[d] + Tt * t[d]) * det;  // w = 1
          if (quirks & FEMX_Q_FORCE_LAST_IP) force[a * 2 + d] = v;        // D8
          else force[a * 2 + d] += v;
        }
      if (stiff)
        for (int b = 0; b < 4; ++b) {
          const double sb = b < 2 ? 1.0 : -1.0;
          for (int d1 = 0; d1 < 2; ++d1)
            for (int d2 = 0; d2 < 2; ++d2) {
              const double t1 = (S[0][0] * nrm[d1] + S[1][0] * t[d1]) * nrm[d2];
              const double t2 = (S[0][1] * nrm[d1] + S[1][1] * t[d1]) * t[d2];
              const double f = sa * sb * N[a] * N[b] * det;
              stiff[(a * 2 + d1) * 8 + b * 2 + d2] +=
                  (quirks & FEMX_Q_TANGENT_PRECEDENCE) ? (t1 + t2 * f) : ((t1 + t2) * f);  // D4
            }
        }
    }
  }
}

}  // namespace femx
