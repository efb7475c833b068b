// oracle.cpp -- CPU restatement of the defmod hot path.  TEST INFRASTRUCTURE ONLY
// (see oracle.h for who may use it).  Every function cites the reference lines it
// follows; loop nests and operation order are the reference's (dense matmul with
// structural zeros, LU inverse per integration point, per-entry sorted insertion).
// Build: g++ -O2 -ffp-contract=off (no FMA contraction, like a plain Fortran build).
//
// Parity: pinned against examples/tmp.txt, the elastic golden VTKs and the UEL
// traction printout (tests/test_oracle_golden.py).  hex8 follows the evident
// intent (8 nodes), since the reference's node-count table says 6 (D1,
// src/m_elems.F90:27) and the element cannot run there.
#include "oracle.h"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

namespace {

// src/m_elems.F90:25-28, :93-105 (nip), :48-61 (dim).  hex nodecount 8 = intent (D1).
const int kNen[6] = {0, 3, 4, 4, 8, 4};
const int kNip[6] = {0, 1, 4, 1, 8, 2};
const int kDim[6] = {0, 2, 2, 3, 3, 2};

// PENALTY_PARAM=1.0e30 is a single-precision literal (src/m_global.F90:15)
const double kPenalty = (double)1.0e30f;

// ---- src/m_utils.F90:45-57 MatInv = dgetrf + dgetri (partial pivoting), n <= 8 ----
void mat_inv(int n, const double *Ain, double *A /*row-major n*n*/) {
  int ipiv[8];
  double work[8];
  for (int i = 0; i < n * n; ++i) A[i] = Ain[i];
  // dgetf2
  for (int j = 0; j < n; ++j) {
    int p = j;
    double amax = std::fabs(A[j * n + j]);
    for (int i = j + 1; i < n; ++i)
      if (std::fabs(A[i * n + j]) > amax) { amax = std::fabs(A[i * n + j]); p = i; }
    ipiv[j] = p;
    if (p != j)
      for (int k = 0; k < n; ++k) std::swap(A[j * n + k], A[p * n + k]);
    double r = 1.0 / A[j * n + j];
    for (int i = j + 1; i < n; ++i) A[i * n + j] *= r;
    for (int k = j + 1; k < n; ++k) {           // rank-1 update, column by column (dger)
      double t = A[j * n + k];
      for (int i = j + 1; i < n; ++i) A[i * n + k] -= A[i * n + j] * t;
    }
  }
  // dtrti2 (upper, non-unit)
  for (int j = 0; j < n; ++j) {
    A[j * n + j] = 1.0 / A[j * n + j];
    double ajj = -A[j * n + j];
    // x = U(0:j,0:j) * A(0:j, j)   (dtrmv upper, no-trans, non-unit)
    for (int jj = 0; jj < j; ++jj) {
      double temp = A[jj * n + j];
      if (temp != 0.0) {
        for (int i = 0; i < jj; ++i) A[i * n + j] += temp * A[i * n + jj];
        A[jj * n + j] *= A[jj * n + jj];
      }
    }
    for (int i = 0; i < j; ++i) A[i * n + j] *= ajj;
  }
  // solve inv(A)*L = inv(U)
  for (int j = n - 1; j >= 0; --j) {
    for (int i = j + 1; i < n; ++i) { work[i] = A[i * n + j]; A[i * n + j] = 0.0; }
    for (int jj = j + 1; jj < n; ++jj) {        // dgemv: A(:,j) -= A(:,jj)*work(jj)
      double t = -work[jj];
      for (int i = 0; i < n; ++i) A[i * n + j] += t * A[i * n + jj];
    }
  }
  for (int j = n - 2; j >= 0; --j) {
    int jp = ipiv[j];
    if (jp != j)
      for (int i = 0; i < n; ++i) std::swap(A[i * n + j], A[i * n + jp]);
  }
}

// src/m_utils.F90:60-66
double mat_det(int n, const double *A) {
  if (n == 2) return A[0] * A[3] - A[1] * A[2];
  return A[0] * (A[4] * A[8] - A[5] * A[7]) - A[1] * (A[3] * A[8] - A[5] * A[6]) +
         A[2] * (A[3] * A[7] - A[4] * A[6]);
}

// src/m_local.F90:175-189 FormdNdetJ.  dN is (dim,nen) row-major.
void form_dn_detj(int type, const double *pt, const double *ecoords, double *dN, double *detj) {
  const int dim = kDim[type], nen = kNen[type];
  double jac[9], inv[9], tmp[24];
  orc_shape_d(type, pt, dN);
  for (int i = 0; i < dim; ++i)
    for (int j = 0; j < dim; ++j) {
      double s = 0.0;
      for (int a = 0; a < nen; ++a) s += dN[i * nen + a] * ecoords[a * dim + j];
      jac[i * dim + j] = s;
    }
  *detj = mat_det(dim, jac);
  mat_inv(dim, jac, inv);
  for (int i = 0; i < dim; ++i)
    for (int a = 0; a < nen; ++a) {
      double s = 0.0;
      for (int j = 0; j < dim; ++j) s += inv[i * dim + j] * dN[j * nen + a];
      tmp[i * nen + a] = s;
    }
  for (int i = 0; i < dim * nen; ++i) dN[i] = tmp[i];
}

// src/m_local.F90:148-171 BMat.  B is (cp, nen*dim) row-major.
void bmat(int dim, int nen, const double *dN, double *B) {
  const int cp = dim == 2 ? 3 : 6, nd = nen * dim;
  for (int i = 0; i < cp * nd; ++i) B[i] = 0.0;
  if (dim == 2) {
    for (int j = 0; j < nen; ++j) {
      B[0 * nd + 2 * j] = dN[0 * nen + j];
      B[1 * nd + 2 * j + 1] = dN[1 * nen + j];
      B[2 * nd + 2 * j] = dN[1 * nen + j];
      B[2 * nd + 2 * j + 1] = dN[0 * nen + j];
    }
  } else {
    for (int j = 0; j < nen; ++j) {
      const double dx = dN[0 * nen + j], dy = dN[1 * nen + j], dz = dN[2 * nen + j];
      B[0 * nd + 3 * j] = dx;
      B[1 * nd + 3 * j + 1] = dy;
      B[2 * nd + 3 * j + 2] = dz;
      B[3 * nd + 3 * j] = dy; B[3 * nd + 3 * j + 1] = dx;
      B[4 * nd + 3 * j + 1] = dz; B[4 * nd + 3 * j + 2] = dy;
      B[5 * nd + 3 * j] = dz; B[5 * nd + 3 * j + 2] = dx;
    }
  }
}

}  // namespace

extern "C" {

int orc_nen(int type) { return kNen[type]; }
int orc_nip(int type) { return kNip[type]; }
int orc_dim(int type) { return kDim[type]; }

// src/m_elems.F90:212-235 (tri), :298-316 (qua), :383-409 (tet), :487-509 (hex), :602-624 (coh)
void orc_sampts(int type, double *pts, double *w) {
  const double g = std::sqrt(1.0 / 3.0);
  switch (type) {
    case ORC_TRI: pts[0] = 1.0 / 3.0; pts[1] = 1.0 / 3.0; w[0] = 0.5; break;
    case ORC_QUA: {
      const double p[8] = {-g, -g, -g, g, g, -g, g, g};
      for (int i = 0; i < 8; ++i) pts[i] = p[i];
      for (int i = 0; i < 4; ++i) w[i] = 1.0;
    } break;
    case ORC_TET:
      pts[0] = pts[1] = pts[2] = 1.0 / 4.0; w[0] = 1.0 / 6.0; break;
    case ORC_HEX: {
      const double p[24] = {-g, -g, -g, g, -g, -g, g, g, -g, -g, g, -g,
                            -g, -g, g,  g, -g, g,  g, g, g,  -g, g, g};
      for (int i = 0; i < 24; ++i) pts[i] = p[i];
      for (int i = 0; i < 8; ++i) w[i] = 1.0;
    } break;
    case ORC_COH: {
      const double CN = 0.5773502691896260;
      pts[0] = -CN; pts[1] = 0.0; pts[2] = CN; pts[3] = 0.0; w[0] = w[1] = 1.0;
    } break;
  }
}

// ShapeFuncPrecomp* src/m_elems.F90:249-270, 332-354, 433-456, 538-566, 626-647
void orc_shape(int type, int ip, double *N) {
  double pts[24], w[8];
  orc_sampts(type, pts, w);
  const int dim = kDim[type];
  const double eta = pts[ip * dim], nu = pts[ip * dim + 1], psi = dim == 3 ? pts[ip * dim + 2] : 0;
  switch (type) {
    case ORC_TRI: N[0] = 1.0 - eta - nu; N[1] = eta; N[2] = nu; break;
    case ORC_QUA:
      N[0] = 0.25 * (1.0 - eta) * (1.0 - nu); N[1] = 0.25 * (1.0 + eta) * (1.0 - nu);
      N[2] = 0.25 * (1.0 + eta) * (1.0 + nu); N[3] = 0.25 * (1.0 - eta) * (1.0 + nu);
      break;
    case ORC_TET: N[0] = 1.0 - eta - nu - psi; N[1] = eta; N[2] = nu; N[3] = psi; break;
    case ORC_HEX: {
      const double c = 0.125;
      N[0] = c * (1 - eta) * (1 - nu) * (1 - psi); N[1] = c * (1 + eta) * (1 - nu) * (1 - psi);
      N[2] = c * (1 + eta) * (1 + nu) * (1 - psi); N[3] = c * (1 - eta) * (1 + nu) * (1 - psi);
      N[4] = c * (1 - eta) * (1 - nu) * (1 + psi); N[5] = c * (1 + eta) * (1 - nu) * (1 + psi);
      N[6] = c * (1 + eta) * (1 + nu) * (1 + psi); N[7] = c * (1 - eta) * (1 + nu) * (1 + psi);
    } break;
    case ORC_COH:
      // 0.5 is a default-real literal promoted to double: exact
      N[0] = 0.5 * (1 - eta); N[1] = 0.5 * (1 + eta); N[2] = N[1]; N[3] = N[0]; break;
  }
}

// ShapeFuncd* src/m_elems.F90:273-279, 357-363, 459-466, 569-580.  No "coh" branch (D5).
void orc_shape_d(int type, const double *pt, double *dN) {
  switch (type) {
    case ORC_TRI: {
      const double d[6] = {-1, 1, 0, -1, 0, 1};
      for (int i = 0; i < 6; ++i) dN[i] = d[i];
    } break;
    case ORC_QUA: {
      const double e = pt[0], n = pt[1];
      dN[0] = 0.25 * -(1 - n); dN[1] = 0.25 * (1 - n); dN[2] = 0.25 * (1 + n); dN[3] = 0.25 * -(1 + n);
      dN[4] = 0.25 * -(1 - e); dN[5] = 0.25 * -(1 + e); dN[6] = 0.25 * (1 + e); dN[7] = 0.25 * (1 - e);
    } break;
    case ORC_TET: {
      const double d[12] = {-1, 1, 0, 0, -1, 0, 1, 0, -1, 0, 0, 1};
      for (int i = 0; i < 12; ++i) dN[i] = d[i];
    } break;
    case ORC_HEX: {
      const double e = pt[0], n = pt[1], s = pt[2], c = 0.125;
      const double a[8] = {-(1 - n) * (1 - s), (1 - n) * (1 - s), (1 + n) * (1 - s), -(1 + n) * (1 - s),
                           -(1 - n) * (1 + s), (1 - n) * (1 + s), (1 + n) * (1 + s), -(1 + n) * (1 + s)};
      const double b[8] = {-(1 - e) * (1 - s), -(1 + e) * (1 - s), (1 + e) * (1 - s), (1 - e) * (1 - s),
                           -(1 - e) * (1 + s), -(1 + e) * (1 + s), (1 + e) * (1 + s), (1 - e) * (1 + s)};
      const double d[8] = {-(1 - e) * (1 - n), -(1 + e) * (1 - n), -(1 + e) * (1 + n), -(1 - e) * (1 + n),
                           (1 - e) * (1 - n),  (1 + e) * (1 - n),  (1 + e) * (1 + n),  (1 - e) * (1 + n)};
      for (int i = 0; i < 8; ++i) { dN[i] = c * a[i]; dN[8 + i] = c * b[i]; dN[16 + i] = c * d[i]; }
    } break;
    default:
      for (int i = 0; i < 8; ++i) dN[i] = 0.0;  // coh: undefined in the reference (D5)
  }
}

// src/m_elems.F90:725-745 NodalStressInv: inverse of N2(j,:) = N(xi_j) when nip == nen.
int orc_nodal_stress_inv(int type, double *N2inv) {
  const int nen = kNen[type], nip = kNip[type];
  if (type == ORC_COH || nip != nen) return 0;
  double N2[64];
  for (int j = 0; j < nip; ++j) orc_shape(type, j, &N2[j * nen]);
  mat_inv(nen, N2, N2inv);
  return 1;
}

// src/m_local.F90:204-228 DMat2d / DMat3d (plane strain / 3-D isotropic)
void orc_dmat(int pdim, double E, double nu, double *D) {
  const int cp = pdim == 2 ? 3 : 6;
  const double c = E / ((1.0 + nu) * (1.0 - 2.0 * nu));
  for (int i = 0; i < cp * cp; ++i) D[i] = 0.0;
  const int nn = pdim;  // normal components
  for (int i = 0; i < nn; ++i)
    for (int j = 0; j < nn; ++j) D[i * cp + j] = c * (i == j ? (1.0 - nu) : nu);
  for (int i = nn; i < cp; ++i) D[i * cp + i] = c * ((1.0 - 2.0 * nu) / 2.0);
}

// src/m_local.F90:21-67 FormElKE: k = sum_ip B^T (D B) * w * detj, dense products.
void orc_form_el_ke(int type, const double *ecoords, double E, double nu, double *k) {
  const int dim = kDim[type], nen = kNen[type], nip = kNip[type];
  const int cp = dim == 2 ? 3 : 6, nd = nen * dim;
  double pts[24], w[8], D[36], dN[24], B[6 * 24], DB[6 * 24], detj;
  for (int i = 0; i < nd * nd; ++i) k[i] = 0.0;
  if (type == ORC_COH) return;  // E = nu = 0 -> D = 0 (src/m_global.F90:250-252); dN undefined (D5)
  orc_sampts(type, pts, w);
  orc_dmat(dim, E, nu, D);
  for (int ip = 0; ip < nip; ++ip) {
    form_dn_detj(type, &pts[ip * dim], ecoords, dN, &detj);
    bmat(dim, nen, dN, B);
    for (int i = 0; i < cp; ++i)
      for (int c = 0; c < nd; ++c) {
        double s = 0.0;
        for (int j = 0; j < cp; ++j) s += D[i * cp + j] * B[j * nd + c];
        DB[i * nd + c] = s;
      }
    for (int c = 0; c < nd; ++c)
      for (int r = 0; r < nd; ++r) {
        double s = 0.0;
        for (int i = 0; i < cp; ++i) s += B[i * nd + r] * DB[i * nd + c];
        k[c * nd + r] += s * w[ip] * detj;
      }
  }
}

// src/m_local.F90:93-123 CalcElStress: strain = B u_e, stress = D strain, per ip.
void orc_calc_el_stress(int type, const double *ecoords, const double *edisp, double E, double nu,
                        double *estress) {
  const int dim = kDim[type], nen = kNen[type], nip = kNip[type];
  const int cp = dim == 2 ? 3 : 6, nd = nen * dim;
  double pts[24], w[8], D[36], dN[24], B[6 * 24], detj, strain[6];
  orc_sampts(type, pts, w);
  orc_dmat(dim, E, nu, D);
  for (int ip = 0; ip < nip; ++ip) {
    form_dn_detj(type, &pts[ip * dim], ecoords, dN, &detj);
    bmat(dim, nen, dN, B);
    for (int i = 0; i < cp; ++i) {
      double s = 0.0;
      for (int c = 0; c < nd; ++c) s += B[i * nd + c] * edisp[c];
      strain[i] = s;
    }
    for (int i = 0; i < cp; ++i) {
      double s = 0.0;
      for (int j = 0; j < cp; ++j) s += D[i * cp + j] * strain[j];
      estress[ip * cp + i] = s;
    }
  }
}

// src/m_global.F90:488-515 RecoverNodalStress
void orc_recover_nodal_stress(int type, const double *estress, double *out) {
  const int dim = kDim[type], nen = kNen[type], nip = kNip[type];
  const int cp = dim == 2 ? 3 : 6;
  if (nip == nen && type != ORC_COH) {
    double inv[64];
    orc_nodal_stress_inv(type, inv);
    for (int c = 0; c < cp; ++c)
      for (int a = 0; a < nen; ++a) {
        double s = 0.0;
        for (int j = 0; j < nip; ++j) s += inv[a * nip + j] * estress[j * cp + c];
        out[a * cp + c] = s;
      }
  } else if (nip == 1) {
    for (int a = 0; a < nen; ++a)
      for (int c = 0; c < cp; ++c) out[a * cp + c] = estress[c];
  } else {
    for (int i = 0; i < nen * cp; ++i) out[i] = 0.0;  // coh: nothing assigned (D5) -> defined as 0
  }
}

// src/m_seplaw.F90:15-53
void orc_seplaw_1_tract(const double *props, const double *gap, const double *vgap, double dtime,
                        double *tract) {
  (void)dtime;
  const double sepwrk = std::exp(1.0) * props[0] * props[1];
  const double dn = props[1], dt = props[2], q = props[3], r = props[4], zeta = props[5];
  double c1, c2;
  c1 = (1.0 - std::exp(-gap[1] * gap[1] / (dt * dt)));
  c1 = c1 * (1.0 - q) / (r - 1.0) * (r - gap[0] / dn);
  c2 = (gap[0] / dn) * std::exp(-gap[1] * gap[1] / (dt * dt));
  tract[0] = (sepwrk / dn) * std::exp(-gap[0] / dn) * (c2 + c1);
  c1 = q + (r - q) / (r - 1.0) * (gap[0] / dn);
  c1 = c1 * std::exp(-gap[0] / dn) * std::exp(-gap[1] * gap[1] / (dt * dt));
  c1 = c1 * 2.0 * (dn / dt) * (sepwrk / dn);
  tract[1] = c1 * gap[1] / dt;
  tract[0] = tract[0] + zeta * props[0] * vgap[0] / dn;
}

// src/m_seplaw.F90:57-94
void orc_seplaw_1_stiff(const double *props, const double *gap, const double *vgap, double dtime,
                        double *stiff) {
  (void)vgap;
  const double sw = std::exp(1.0) * props[0] * props[1];
  const double dn = props[1], dt = props[2], q = props[3], r = props[4], zeta = props[5];
  double c1;
  c1 = (1.0 - q) / (r - 1) * (1.0 - std::exp(-gap[1] * gap[1] / (dt * dt)));
  c1 = c1 * (r + 1.0 - gap[0] / dn);
  c1 = (1.0 - gap[0] / dn) * std::exp(-gap[1] * gap[1] / (dt * dt)) - c1;
  stiff[0] = (sw / (dn * dn)) * std::exp(-gap[0] / dn) * c1;
  c1 = q + (gap[0] / dn) * (r - q) / (r - 1.0);
  c1 = c1 * std::exp(-gap[0] / dn) * std::exp(-gap[1] * gap[1] / (dt * dt));
  c1 = 2.0 * (sw / (dt * dt)) * c1;
  stiff[3] = c1 * (1.0 - 2.0 * (gap[1] * gap[1] / (dt * dt)));
  c1 = -gap[0] / dn + (1.0 - q) / (r - 1.0) * (r - gap[0] / dn);
  c1 = c1 * std::exp(-gap[0] / dn) * std::exp(-gap[1] * gap[1] / (dt * dt));
  c1 = 2.0 * (sw / (dt * dn)) * c1;
  stiff[1] = (gap[1] / dt) * c1;
  stiff[2] = stiff[1];
  stiff[0] = stiff[0] + zeta * props[0] / dn / dtime;
}

// src/m_elems.F90:658-673 getCohValues (+ magnitude src/m_utils.F90:69-81)
void orc_coh_values(const double *x, double *tangent, double *normal, double *det) {
  tangent[0] = 0.5 * (x[2] - x[0] + x[4] - x[6]);
  tangent[1] = 0.5 * (x[3] - x[1] + x[5] - x[7]);
  double m = 0.0;
  for (int i = 0; i < 2; ++i) m = m + tangent[i] * tangent[i];
  m = std::sqrt(m);
  tangent[0] = tangent[0] / m;
  tangent[1] = tangent[1] / m;
  *det = m * 0.5;
  normal[0] = -tangent[1];
  normal[1] = tangent[0];
}

static void coh_gaps(const int *nodes, const double *u_local, int ip, double dt, int quirks,
                     const double *tangent, const double *normal, double *gap, double *vgap,
                     double *N) {
  // getCohRels src/m_elems.F90:676-706.  D3: the reference indexes local_u with the
  // element-local node number 1..4, i.e. the first four nodes of the rank.
  double urel[2] = {0, 0}, vrel[2];
  orc_shape(ORC_COH, ip, N);
  for (int n1 = 0; n1 < 2; ++n1) {
    const int n2 = n1 + 2;
    const int g1 = (quirks & ORC_Q_GATHER_LOCAL14) ? n1 : nodes[n1] - 1;
    const int g2 = (quirks & ORC_Q_GATHER_LOCAL14) ? n2 : nodes[n2] - 1;
    for (int i = 0; i < 2; ++i)
      urel[i] = urel[i] + N[n2] * u_local[2 * g2 + i] - N[n1] * u_local[2 * g1 + i];
  }
  vrel[0] = urel[0] / dt;
  vrel[1] = urel[1] / dt;
  // getCohGaps src/m_elems.F90:709-720
  gap[0] = normal[0] * urel[0] + normal[1] * urel[1];
  gap[1] = tangent[0] * urel[0] + tangent[1] * urel[1];
  vgap[0] = normal[0] * vrel[0] + normal[1] * vrel[1];
  vgap[1] = tangent[0] * vrel[0] + tangent[1] * vrel[1];
}

// src/m_global.F90:622-682 applyTract_1.  D8: `result(...) = ...` inside the ip loop.
void orc_apply_tract_1(const double *ecoords, const int *nodes, const double *u_local, double dt,
                       const double *props, int quirks, double *result, double *per_ip) {
  double tangent[2], normal[2], det, gap[2], vgap[2], f[2], N[4];
  const double wts[2] = {1.0, 1.0};
  orc_coh_values(ecoords, tangent, normal, &det);
  for (int i = 0; i < 8; ++i) result[i] = 0.0;
  for (int ip = 0; ip < 2; ++ip) {
    coh_gaps(nodes, u_local, ip, dt, quirks, tangent, normal, gap, vgap, N);
    orc_seplaw_1_tract(props, gap, vgap, dt, f);
    double sig = 1.0;
    for (int node = 0; node < 4; ++node) {
      if (node + 1 > 2) sig = -1.0;
      for (int d = 0; d < 2; ++d) {
        const double v = sig * N[node] * (f[0] * normal[d] + f[1] * tangent[d]) * wts[ip] * det;
        if (per_ip) per_ip[ip * 8 + node * 2 + d] = v;
        if (quirks & ORC_Q_FORCE_LAST_IP) result[node * 2 + d] = v;
        else result[node * 2 + d] = result[node * 2 + d] + v;
      }
    }
  }
}

// src/m_global.F90:762-845 applyStiff_1.  D4: at :829-831 only the SECOND bracket is
// multiplied by sig1*sig2*N*N*w*det; the UEL multiplies the sum (Cohesive_element.f:169-172).
void orc_apply_stiff_1(const double *ecoords, const int *nodes, const double *u_local, double dt,
                       const double *props, int quirks, double *result) {
  double tangent[2], normal[2], det, gap[2], vgap[2], S[4], N[4];
  const double wts[2] = {1.0, 1.0};
  orc_coh_values(ecoords, tangent, normal, &det);
  for (int i = 0; i < 64; ++i) result[i] = 0.0;
  for (int ip = 0; ip < 2; ++ip) {
    coh_gaps(nodes, u_local, ip, dt, quirks, tangent, normal, gap, vgap, N);
    orc_seplaw_1_stiff(props, gap, vgap, dt, S);  // S[i*2+j] = stiff(i+1,j+1)
    double sig1 = 1.0;
    for (int n1 = 0; n1 < 4; ++n1) {
      if (n1 + 1 > 2) sig1 = -1.0;
      double sig2 = 1.0;
      for (int n2 = 0; n2 < 4; ++n2) {
        if (n2 + 1 > 2) sig2 = -1.0;
        for (int d1 = 0; d1 < 2; ++d1)
          for (int d2 = 0; d2 < 2; ++d2) {
            const double t1 = (S[0] * normal[d1] + S[2] * tangent[d1]) * normal[d2];
            const double t2 = (S[1] * normal[d1] + S[3] * tangent[d1]) * tangent[d2];
            double *r = &result[(n1 * 2 + d1) * 8 + (n2 * 2 + d2)];
            if (quirks & ORC_Q_TANGENT_PRECEDENCE)
              *r = *r + t1 + t2 * sig1 * sig2 * N[n1] * N[n2] * wts[ip] * det;
            else
              *r = *r + (t1 + t2) * sig1 * sig2 * N[n1] * N[n2] * wts[ip] * det;
          }
      }
    }
  }
}

// ---------------------------------------------------------------------------------
// MatSetValues semantics (PETSc SeqAIJ): each row keeps its columns sorted; a new
// (row, col) is inserted in place (explicit zeros are kept); ADD on an existing
// entry accumulates in call order.  Rows start with `prealloc` slots like
// MatCreateAIJ(d_nz) at src/main.F90:153-156; assembly squeezes the unused ones.
// ---------------------------------------------------------------------------------
struct orc_csr {
  int64_t nrows;
  std::vector<std::vector<int32_t>> cols;
  std::vector<std::vector<double>> vals;
};

orc_csr *orc_csr_create(int64_t nrows, int prealloc) {
  orc_csr *m = new orc_csr;
  m->nrows = nrows;
  m->cols.resize(nrows);
  m->vals.resize(nrows);
  if (prealloc > 0)
    for (int64_t i = 0; i < nrows; ++i) { m->cols[i].reserve(prealloc); m->vals[i].reserve(prealloc); }
  return m;
}
void orc_csr_destroy(orc_csr *m) { delete m; }

static inline void csr_set(orc_csr *m, int64_t row, int64_t col, double v, bool add) {
  std::vector<int32_t> &c = m->cols[row];
  std::vector<double> &a = m->vals[row];
  // bisection then short linear scan, as MatSetValues_SeqAIJ does
  int lo = 0, hi = (int)c.size();
  while (hi - lo > 5) {
    int t = (lo + hi) / 2;
    if (c[t] > col) hi = t; else lo = t;
  }
  int i = lo;
  for (; i < hi; ++i) {
    if (c[i] > col) break;
    if (c[i] == col) { if (add) a[i] += v; else a[i] = v; return; }
  }
  c.insert(c.begin() + i, (int32_t)col);
  a.insert(a.begin() + i, v);
}
void orc_csr_add(orc_csr *m, int64_t r, int64_t c, double v) { csr_set(m, r, c, v, true); }
void orc_csr_insert(orc_csr *m, int64_t r, int64_t c, double v) { csr_set(m, r, c, v, false); }
int64_t orc_csr_nnz(const orc_csr *m) {
  int64_t n = 0;
  for (auto &c : m->cols) n += (int64_t)c.size();
  return n;
}
void orc_csr_export(const orc_csr *m, int64_t *rowptr, int32_t *colidx, double *vals) {
  int64_t p = 0;
  for (int64_t i = 0; i < m->nrows; ++i) {
    rowptr[i] = p;
    for (size_t k = 0; k < m->cols[i].size(); ++k) {
      if (colidx) colidx[p] = m->cols[i][k];
      if (vals) vals[p] = m->vals[i][k];
      ++p;
    }
  }
  rowptr[m->nrows] = p;
}
orc_csr *orc_csr_copy(const orc_csr *m) { return new orc_csr(*m); }
void orc_csr_mult(const orc_csr *m, const double *x, double *y) {
  for (int64_t i = 0; i < m->nrows; ++i) {
    double s = 0.0;
    for (size_t k = 0; k < m->cols[i].size(); ++k) s += m->vals[i][k] * x[m->cols[i][k]];
    y[i] = s;
  }
}

static void element_ke_global(const orc_mesh *M, int e, double *k, int *gidx) {
  // FormLocalK src/m_global.F90:239-261 + FormElIndx src/m_local.F90:70-78 + indxmap(:,2)
  const int type = M->etype[e], nen = kNen[type], dim = M->pdim;
  double xe[24];
  const int *nd = &M->eind[M->eptr[e]];
  for (int a = 0; a < nen; ++a)
    for (int d = 0; d < dim; ++d) xe[a * dim + d] = M->coords[(size_t)(nd[a] - 1) * dim + d];
  double E = 0.0, nu = 0.0;
  if (M->emat[e] != 0) { E = M->mat[(M->emat[e] - 1) * 2]; nu = M->mat[(M->emat[e] - 1) * 2 + 1]; }
  orc_form_el_ke(type, xe, E, nu, k);
  for (int a = 0; a < nen; ++a)
    for (int d = 0; d < dim; ++d) gidx[a * dim + d] = dim * (M->nl2g[nd[a] - 1] - 1) + d;
}

static void apply_kbc(const orc_mesh *M, orc_csr *A) {
  // ApplyKBC / UpdateKMatBcs src/m_global.F90:264-299: K(j,j) <- 1e30 INSERT
  for (int i = 0; i < M->nbcs; ++i) {
    const int node = M->bcnode[i * (1 + M->pdim)];
    for (int d = 0; d < M->pdim; ++d)
      if (M->bcnode[i * (1 + M->pdim) + 1 + d] == 0) {
        const int64_t j = (int64_t)M->pdim * node - M->pdim + d;
        orc_csr_insert(A, j, j, kPenalty);
      }
  }
}

// src/main.F90:157-171.  D6: the column-major k is handed to a row-oriented MatSetValues,
// so memory[i*n+j] (= k(j,i)) lands at (indx[i], indx[j]).
orc_csr *orc_assemble_K(const orc_mesh *M, int prealloc, int apply_bc) {
  orc_csr *A = orc_csr_create((int64_t)M->pdim * M->nnds, prealloc);
  for (int e = 0; e < M->nels; ++e) {
    const int n = kNen[M->etype[e]] * M->pdim;
    double *k = (double *)std::malloc(sizeof(double) * n * n);  // allocate(k), allocate(indx)
    int *indx = (int *)std::malloc(sizeof(int) * n);
    element_ke_global(M, e, k, indx);
    for (int i = 0; i < n; ++i)
      for (int j = 0; j < n; ++j) orc_csr_add(A, indx[i], indx[j], k[i * n + j]);
    std::free(k);
    std::free(indx);
  }
  if (apply_bc) apply_kbc(M, A);
  return A;
}

// src/m_global.F90:98-158 CalcJacobian
orc_csr *orc_jacobian(const orc_mesh *M, const orc_csr *K, const double *u_total, double dt,
                      int quirks) {
  orc_csr *J = orc_csr_copy(K);  // MatCopy(Mat_K, Jacobian)
  for (int e = 0; e < M->nels; ++e) {
    if (M->enlmat[e] == 0) continue;
    const int *nd = &M->eind[M->eptr[e]];
    double xe[8], S[64];
    for (int a = 0; a < 4; ++a)
      for (int d = 0; d < 2; ++d) xe[a * 2 + d] = M->coords[(size_t)(nd[a] - 1) * 2 + d];
    orc_apply_stiff_1(xe, nd, u_total, dt, &M->cohprops[(M->enlmat[e] - 1) * 6], quirks, S);
    for (int n1 = 0; n1 < 4; ++n1)
      for (int n2 = 0; n2 < 4; ++n2)
        for (int d1 = 0; d1 < 2; ++d1)
          for (int d2 = 0; d2 < 2; ++d2)
            orc_csr_add(J, 2 * (M->nl2g[nd[n1] - 1] - 1) + d1, 2 * (M->nl2g[nd[n2] - 1] - 1) + d2,
                        S[(n1 * 2 + d1) * 8 + n2 * 2 + d2]);
  }
  apply_kbc(M, J);
  return J;
}

// src/m_global.F90:161-235 CalcResidual (+ EnforceBCForce :439-455)
void orc_residual(const orc_mesh *M, const orc_csr *J, const double *u_total, const double *du,
                  const double *f_ext, double dt, double t_total, int quirks, double *r) {
  const int64_t n = (int64_t)M->pdim * M->nnds;
  for (int64_t i = 0; i < n; ++i) r[i] = f_ext[i];  // VecCopy(Vec_F, Residual)
  for (int e = 0; e < M->nels; ++e) {
    if (M->enlmat[e] == 0) continue;
    const int *nd = &M->eind[M->eptr[e]];
    double xe[8], f[8];
    for (int a = 0; a < 4; ++a)
      for (int d = 0; d < 2; ++d) xe[a * 2 + d] = M->coords[(size_t)(nd[a] - 1) * 2 + d];
    orc_apply_tract_1(xe, nd, u_total, dt, &M->cohprops[(M->enlmat[e] - 1) * 6], quirks, f, nullptr);
    for (int a = 0; a < 4; ++a)
      for (int d = 0; d < 2; ++d) r[2 * (M->nl2g[nd[a] - 1] - 1) + d] += f[a * 2 + d];
  }
  for (int i = 0; i < M->nbcs; ++i) {  // EnforceBCForce: INSERT 1e30*bcval*dt/t
    const int node = M->bcnode[i * (1 + M->pdim)];
    for (int d = 0; d < M->pdim; ++d)
      if (M->bcnode[i * (1 + M->pdim) + 1 + d] == 0)
        r[(int64_t)M->pdim * node - M->pdim + d] = kPenalty * M->bcval[i * M->pdim + d] * dt / t_total;
  }
  std::vector<double> y(n);
  orc_csr_mult(J, du, y.data());
  for (int64_t i = 0; i < n; ++i) r[i] = y[i] + (-1.0 * r[i]);  // VecScale(-1); MatMultAdd
}

// matrix-free equal of the elastic part of MatMultAdd(Jacobian, du, ...): f = sum_e Ke u_e
void orc_fint(const orc_mesh *M, const double *u_local, double *f) {
  const int64_t n = (int64_t)M->pdim * M->nnds;
  for (int64_t i = 0; i < n; ++i) f[i] = 0.0;
  double k[576];
  int gidx[24];
  for (int e = 0; e < M->nels; ++e) {
    const int type = M->etype[e], nd = kNen[type] * M->pdim;
    element_ke_global(M, e, k, gidx);
    const int *en = &M->eind[M->eptr[e]];
    for (int i = 0; i < nd; ++i) {
      double s = 0.0;
      for (int j = 0; j < nd; ++j)
        s += k[j * nd + i] * u_local[(size_t)(en[j / M->pdim] - 1) * M->pdim + j % M->pdim];
      f[gidx[i]] += s;
    }
  }
}

// src/main.F90:253-279 (RecoverStress src/m_global.F90:466-486, RecoverNodalStress :488-515)
void orc_stress(const orc_mesh *M, const double *uu, int quirks, double *stress_node, int *count) {
  const int dim = M->pdim, cp = dim == 2 ? 3 : 6;
  for (size_t i = 0; i < (size_t)M->nlnds * cp; ++i) stress_node[i] = 0.0;
  for (int i = 0; i < M->nlnds; ++i) count[i] = 0;
  for (int e = 0; e < M->nels; ++e) {
    const int type = M->etype[e], nen = kNen[type];
    const int *nd = &M->eind[M->eptr[e]];
    if (type == ORC_COH) {  // D5: contribution defined as 0; the count is still bumped
      if (quirks & ORC_Q_COH_COUNTS_NODE)
        for (int a = 0; a < nen; ++a) count[nd[a] - 1] += 1;
      continue;
    }
    double xe[24], ue[24], sip[48], sn[48];
    for (int a = 0; a < nen; ++a)
      for (int d = 0; d < dim; ++d) {
        xe[a * dim + d] = M->coords[(size_t)(nd[a] - 1) * dim + d];
        ue[a * dim + d] = uu[(size_t)(nd[a] - 1) * dim + d];
      }
    const double E = M->mat[(M->emat[e] - 1) * 2], nu = M->mat[(M->emat[e] - 1) * 2 + 1];
    orc_calc_el_stress(type, xe, ue, E, nu, sip);
    orc_recover_nodal_stress(type, sip, sn);
    for (int a = 0; a < nen; ++a) {
      count[nd[a] - 1] += 1;
      for (int c = 0; c < cp; ++c) stress_node[(size_t)(nd[a] - 1) * cp + c] += sn[a * cp + c];
    }
  }
  for (int i = 0; i < M->nlnds; ++i)
    for (int c = 0; c < cp; ++c) stress_node[(size_t)i * cp + c] /= count[i];
}

// CPU baseline: the assembly loop of src/main.F90:157-166 plus f = sum Ke u_e, per thread on
// a contiguous element range into its own matrix (one "MPI rank" per thread; the interface
// merge is not timed, which favours the CPU).  Every thread's matrix gets the reference's row
// preallocation (src/main.F90:153) on the rows its elements touch, set up before the clock starts.
double orc_time_assembly(const orc_mesh *M, int prealloc, const double *u_local, int nthreads,
                         int64_t *nnz_out) {
  if (nthreads < 1) nthreads = 1;
  std::vector<orc_csr *> mats(nthreads, nullptr);
  std::vector<std::vector<double>> fs(nthreads);
  const int64_t n = (int64_t)M->pdim * M->nnds;
  for (int t = 0; t < nthreads; ++t) {
    mats[t] = orc_csr_create(n, 0);
    fs[t].assign(n, 0.0);
    if (prealloc > 0) {
      const int e0 = (int)((int64_t)M->nels * t / nthreads), e1 = (int)((int64_t)M->nels * (t + 1) / nthreads);
      for (int e = e0; e < e1; ++e)
        for (int q = M->eptr[e]; q < M->eptr[e + 1]; ++q) {
          const int64_t g = (int64_t)M->pdim * (M->nl2g[M->eind[q] - 1] - 1);
          for (int d = 0; d < M->pdim; ++d)
            if (mats[t]->cols[g + d].capacity() < (size_t)prealloc) { mats[t]->cols[g + d].reserve(prealloc); mats[t]->vals[g + d].reserve(prealloc); }
        }
    }
  }
  auto work = [&](int t) {
    const int e0 = (int)((int64_t)M->nels * t / nthreads), e1 = (int)((int64_t)M->nels * (t + 1) / nthreads);
    orc_csr *A = mats[t];
    double *f = fs[t].data();
    for (int e = e0; e < e1; ++e) {
      const int nd = kNen[M->etype[e]] * M->pdim;
      double *k = (double *)std::malloc(sizeof(double) * nd * nd);
      int *indx = (int *)std::malloc(sizeof(int) * nd);
      element_ke_global(M, e, k, indx);
      for (int i = 0; i < nd; ++i)
        for (int j = 0; j < nd; ++j) orc_csr_add(A, indx[i], indx[j], k[i * nd + j]);
      const int *en = &M->eind[M->eptr[e]];
      for (int i = 0; i < nd; ++i) {
        double s = 0.0;
        for (int j = 0; j < nd; ++j)
          s += k[j * nd + i] * u_local[(size_t)(en[j / M->pdim] - 1) * M->pdim + j % M->pdim];
        f[indx[i]] += s;
      }
      std::free(k);
      std::free(indx);
    }
  };
  auto t0 = std::chrono::steady_clock::now();
  if (nthreads == 1) work(0);
  else {
    std::vector<std::thread> th;
    for (int t = 0; t < nthreads; ++t) th.emplace_back(work, t);
    for (auto &x : th) x.join();
  }
  auto t1 = std::chrono::steady_clock::now();
  int64_t nnz = 0;
  for (int t = 0; t < nthreads; ++t) { nnz += orc_csr_nnz(mats[t]); orc_csr_destroy(mats[t]); }
  if (nnz_out) *nnz_out = nnz;
  return std::chrono::duration<double>(t1 - t0).count();
}


// ---- loads and the right-hand side ----
// MatDet of a 3x3 (src/m_utils.F90:60-66) on the columns TriArea builds (src/m_utils.F90:25-33)
static double det3_ones(double a1, double a2, double a3, double b1, double b2, double b3) {
  // A = [a1 b1 1; a2 b2 1; a3 b3 1]
  return a1 * (b2 * 1.0 - 1.0 * b3) - b1 * (a2 * 1.0 - 1.0 * a3) + 1.0 * (a2 * b3 - b2 * a3);
}
static double tri_area(const double *p1, const double *p2, const double *p3) {
  const double d1 = det3_ones(p1[0], p2[0], p3[0], p1[1], p2[1], p3[1]);
  const double d2 = det3_ones(p1[1], p2[1], p3[1], p1[2], p2[2], p3[2]);
  const double d3 = det3_ones(p1[2], p2[2], p3[2], p1[0], p2[0], p3[0]);
  return 0.5 * std::sqrt(std::fabs(d1) * std::fabs(d1) + std::fabs(d2) * std::fabs(d2) + std::fabs(d3) * std::fabs(d3));
}

// EdgeAreaNodes{Tri,Qua,Tet,Hex} (src/m_elems.F90:282-293, 366-378, 469-482, 583-599): local node slots (0-based)
// of side `side` (1-based) and its length / area; returns nodes per side (elNodePerSide, src/m_elems.F90:28)
int orc_edge_area_nodes(int type, const double *ecoords, int side, double *area, int *slots) {
  static const int tri[3][2] = {{0, 1}, {1, 2}, {2, 0}};
  static const int qua[4][2] = {{0, 1}, {1, 2}, {2, 3}, {3, 0}};
  static const int tet[4][3] = {{0, 1, 3}, {1, 2, 3}, {0, 2, 3}, {0, 1, 2}};
  static const int hex[6][4] = {{0, 1, 5, 4}, {1, 2, 6, 5}, {2, 3, 7, 6}, {3, 0, 4, 7}, {0, 1, 2, 3}, {4, 5, 6, 7}};
  if (type == ORC_TRI || type == ORC_QUA) {
    const int *j = type == ORC_TRI ? tri[side - 1] : qua[side - 1];
    slots[0] = j[0]; slots[1] = j[1];
    const double dx = ecoords[j[0] * 2] - ecoords[j[1] * 2], dy = ecoords[j[0] * 2 + 1] - ecoords[j[1] * 2 + 1];
    *area = std::sqrt(dx * dx + dy * dy);
    return 2;
  }
  if (type == ORC_TET) {
    const int *j = tet[side - 1];
    for (int q = 0; q < 3; ++q) slots[q] = j[q];
    *area = tri_area(ecoords + 3 * j[0], ecoords + 3 * j[1], ecoords + 3 * j[2]);
    return 3;
  }
  const int *j = hex[side - 1];
  for (int q = 0; q < 4; ++q) slots[q] = j[q];
  const double ta1 = tri_area(ecoords + 3 * j[0], ecoords + 3 * j[1], ecoords + 3 * j[2]);   // QuadArea, src/m_utils.F90:36-42
  const double ta2 = tri_area(ecoords + 3 * j[0], ecoords + 3 * j[2], ecoords + 3 * j[3]);
  *area = std::fabs(ta1) + std::fabs(ta2);
  return 4;
}

// FormRHS (src/m_global.F90:373-436): Vec_F = 0; time-ramped nodal forces (fnode global 1-based, fval(nfrcs, pdim+2)
// = components, t1, t2); tractions (telsd = local element 1-based / side, tval(ntrcs, pdim+2); NOTE t1, t2 are divided
// by dt here and only here, :414-415) spread by ApplyTraction (:343-368) as vvec*area/nps on the side's nodes;
// then EnforceBCForce (:439-455) INSERTs 1e30*bcval*dt/t on the constrained dofs.  nceqs is not supported by the
// reference itself (cval is never allocated, src/main.F90:108-109).
void orc_form_rhs(const orc_mesh *M, int nfrcs, const int *fnode, const double *fval, int ntrcs, const int *telsd,
                  const double *tval, double t_init, double dt, double t_total, double *f) {
  const int pd = M->pdim, w = pd + 2;
  const int64_t n = (int64_t)pd * M->nnds;
  for (int64_t i = 0; i < n; ++i) f[i] = 0.0;
  const double t_end = t_init + dt;
  for (int i = 0; i < nfrcs; ++i) {
    const double t1 = fval[i * w + pd], t2 = fval[i * w + pd + 1];
    if (t_end < t1 || t_init > t2) continue;
    const double interval = std::fmin(t2, t_end) - std::fmax(t1, t_init);
    for (int d = 0; d < pd; ++d) f[(int64_t)pd * fnode[i] - pd + d] += fval[i * w + d] * interval / (t2 - t1);
  }
  for (int i = 0; i < ntrcs; ++i) {
    const double t1 = tval[i * w + pd] / dt, t2 = tval[i * w + pd + 1] / dt;
    if (t_end < t1 || t_init > t2) continue;
    const double interval = std::fmin(t2, t_end) - std::fmax(t1, t_init);
    const int el = telsd[i * 2], side = telsd[i * 2 + 1];
    if (el == 0) continue;
    const int e = el - 1, type = M->etype[e], nen = orc_nen(type);
    const int *nd = &M->eind[M->eptr[e]];
    double xe[24];
    for (int a = 0; a < nen; ++a)
      for (int d = 0; d < pd; ++d) xe[a * pd + d] = M->coords[(size_t)(nd[a] - 1) * pd + d];
    double area;
    int slots[4];
    const int nps = orc_edge_area_nodes(type, xe, side, &area, slots);
    double vvec[3];
    for (int d = 0; d < pd; ++d) vvec[d] = tval[i * w + d] * (interval / (t2 - t1));
    for (int d = 0; d < pd; ++d) vvec[d] = vvec[d] * area / (double)nps;
    for (int q = 0; q < nps; ++q) {
      const int g = M->nl2g[nd[slots[q]] - 1];
      for (int d = 0; d < pd; ++d) f[(int64_t)pd * g - pd + d] += vvec[d];
    }
  }
  for (int i = 0; i < M->nbcs; ++i) {
    const int node = M->bcnode[i * (1 + pd)];
    for (int d = 0; d < pd; ++d)
      if (M->bcnode[i * (1 + pd) + 1 + d] == 0) f[(int64_t)pd * node - pd + d] = kPenalty * M->bcval[i * pd + d] * dt / t_total;
  }
}

}  // extern "C"
