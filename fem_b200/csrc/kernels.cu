// kernels.cu -- two-pass assembly (variant 1), stress, cohesive overlay, SpMV, pattern export.
//
// Pass 1: one thread per (element, local node a) computes row-block a of Ke in registers and
//         stores it block-major [a][b][i][j] into the packed scratch, plus f_e = Ke u_e.
// Pass 2: one warp per node row sums, for every node block of that row, the incident
//         elements' blocks in ascending element order (deterministic, atomic-free; each CSR
//         value is written exactly once) -- the gather-ordered equivalent of
//         MatSetValues(ADD_VALUES) + MatAssembly (src/main.F90:157-171).
#include "elem_math.cuh"
#include "launch.h"

namespace femx {

__constant__ double c_n2inv_qua[16];
__constant__ double c_n2inv_hex[64];

template <int T>
__device__ __forceinline__ void load_elem(const DevMesh &dm, int e, double (&X)[ET<T>::NEN][ET<T>::DIM],
                                          int (&nd)[ET<T>::NEN]) {
  const int32_t *c = dm.conn0 + dm.eptr[e];
#pragma unroll
  for (int a = 0; a < ET<T>::NEN; ++a) {
    nd[a] = c[a];
#pragma unroll
    for (int d = 0; d < ET<T>::DIM; ++d) X[a][d] = dm.coords[(int64_t)nd[a] * ET<T>::DIM + d];
  }
}

// ---------------------------------------------------------------- pass 1
template <int T>
__global__ void __launch_bounds__(128) elem_rows_kernel(DevMesh dm, const int32_t *__restrict__ elems, int n,
                                                        const double *__restrict__ u, double *__restrict__ ke,
                                                        double *__restrict__ fe) {
  constexpr int NEN = ET<T>::NEN, DIM = ET<T>::DIM, D2 = DIM * DIM;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= (int64_t)n * NEN) return;
  const int e = elems[tid / NEN], a = (int)(tid % NEN);
  double *kout = ke + dm.ke_off[e] + (int64_t)a * NEN * D2;
  if constexpr (T == FEMX_COH) {
    // E = nu = 0 -> D = 0 (src/m_global.F90:250-252): an explicit zero block that still
    // enters the sparsity pattern (src/main.F90:162)
    for (int q = 0; q < NEN * D2; ++q) kout[q] = 0.0;
    if (fe)
      for (int i = 0; i < DIM; ++i) fe[dm.fe_off[e] + a * DIM + i] = 0.0;
  } else {
    double X[NEN][DIM];
    int nd[NEN];
    load_elem<T>(dm, e, X, nd);
    const int m = dm.emat[e] - 1;
    const double lam = dm.lam_mu[2 * m], mu = dm.lam_mu[2 * m + 1];
    double acc[NEN][D2];
    ke_row_blocks<T>(X, a, lam, mu, acc);
#pragma unroll
    for (int b = 0; b < NEN; ++b)
#pragma unroll
      for (int q = 0; q < D2; ++q) kout[b * D2 + q] = acc[b][q];
    if (fe) {
      double f[DIM];
#pragma unroll
      for (int i = 0; i < DIM; ++i) f[i] = 0.0;
#pragma unroll
      for (int b = 0; b < NEN; ++b)
#pragma unroll
        for (int j = 0; j < DIM; ++j) {
          const double ub = u[(int64_t)nd[b] * DIM + j];
#pragma unroll
          for (int i = 0; i < DIM; ++i) f[i] += acc[b][i * DIM + j] * ub;
        }
#pragma unroll
      for (int i = 0; i < DIM; ++i) fe[dm.fe_off[e] + a * DIM + i] = f[i];
    }
  }
}

// ---------------------------------------------------------------- pass 2
template <int PD>
__global__ void __launch_bounds__(256) gather_rows_kernel(DevMesh dm, const double *__restrict__ ke,
                                                          const double *__restrict__ fe, double *__restrict__ vals,
                                                          double *__restrict__ fint) {
  constexpr int D2 = PD * PD;
  const int lane = threadIdx.x & 31;
  const int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (r >= dm.nlnds) return;
  const int64_t b0 = dm.bptr[r];
  const int nb = (int)(dm.bptr[r + 1] - b0);
  const int64_t p0 = dm.n2e_ptr[r], p1 = dm.n2e_ptr[r + 1];
  double *vrow = vals + D2 * b0;
  for (int chunk = 0; chunk < nb; chunk += 32) {
    const int k = chunk + lane;
    double acc[D2];
#pragma unroll
    for (int q = 0; q < D2; ++q) acc[q] = 0.0;
    for (int64_t p = p0; p < p1; ++p) {
      const int e = dm.n2e_elem[p], a = dm.n2e_a[p];
      const int nen = type_nen(dm.etype[e]);
      const int my = lane < nen ? (int)dm.bpos[dm.eb_ptr[e] + a * nen + lane] : -1;
      const double *kb = ke + dm.ke_off[e] + (int64_t)a * nen * D2;
      for (int b = 0; b < nen; ++b) {
        const int kk = __shfl_sync(0xffffffffu, my, b);
        if (kk == k) {
#pragma unroll
          for (int q = 0; q < D2; ++q) acc[q] += kb[b * D2 + q];
        }
      }
    }
    if (k < nb) {
#pragma unroll
      for (int i = 0; i < PD; ++i)
#pragma unroll
        for (int j = 0; j < PD; ++j) vrow[(int64_t)i * PD * nb + PD * k + j] = acc[i * PD + j];
    }
  }
  if (fint && lane < PD) {
    double s = 0.0;
    for (int64_t p = p0; p < p1; ++p) s += fe[dm.fe_off[dm.n2e_elem[p]] + dm.n2e_a[p] * PD + lane];
    fint[(int64_t)PD * (dm.nl2g[r] - 1) + lane] = s;
  }
}

// ---------------------------------------------------------------- unit-level element API
template <int T>
__global__ void element_ke_kernel(int n, const double *__restrict__ ecoords, const double *__restrict__ E_nu,
                                  double *__restrict__ out) {
  constexpr int NEN = ET<T>::NEN, DIM = ET<T>::DIM, ND = NEN * DIM;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= (int64_t)n * NEN) return;
  const int64_t e = tid / NEN;
  const int a = (int)(tid % NEN);
  double X[NEN][DIM];
  for (int b = 0; b < NEN; ++b)
    for (int d = 0; d < DIM; ++d) X[b][d] = ecoords[(e * NEN + b) * DIM + d];
  double lam, mu;
  lame(E_nu[2 * e], E_nu[2 * e + 1], &lam, &mu);
  double acc[NEN][DIM * DIM];
  ke_row_blocks<T>(X, a, lam, mu, acc);
  double *k = out + e * ND * ND;  // column-major k(row, col), src/m_local.F90:43
#pragma unroll
  for (int b = 0; b < NEN; ++b)
#pragma unroll
    for (int i = 0; i < DIM; ++i)
#pragma unroll
      for (int j = 0; j < DIM; ++j) k[(b * DIM + j) * ND + a * DIM + i] = acc[b][i * DIM + j];
}

// nodal extrapolation: RecoverNodalStress src/m_global.F90:488-515 with the memoised inverse of
// N2(j,:) = N(xi_j) (NodalStressInv src/m_elems.F90:725-745) held in __constant__
template <int T>
__device__ __forceinline__ void nodal_from_ip(const double (&sig)[ET<T>::NIP][ET<T>::CP],
                                              double (&sn)[ET<T>::NEN][ET<T>::CP]) {
  constexpr int NEN = ET<T>::NEN, NIP = ET<T>::NIP, CP = ET<T>::CP;
  if constexpr (NIP == 1) {
#pragma unroll
    for (int a = 0; a < NEN; ++a)
#pragma unroll
      for (int c = 0; c < CP; ++c) sn[a][c] = sig[0][c];
  } else {
    const double *inv = T == FEMX_QUA ? c_n2inv_qua : c_n2inv_hex;
#pragma unroll
    for (int a = 0; a < NEN; ++a)
#pragma unroll
      for (int c = 0; c < CP; ++c) {
        double s = 0.0;
#pragma unroll
        for (int j = 0; j < NIP; ++j) s += inv[a * NIP + j] * sig[j][c];
        sn[a][c] = s;
      }
  }
}

template <int T>
__global__ void element_stress_kernel(int n, const double *__restrict__ ecoords, const double *__restrict__ edisp,
                                      const double *__restrict__ E_nu, double *__restrict__ estress,
                                      double *__restrict__ enodal) {
  constexpr int NEN = ET<T>::NEN, DIM = ET<T>::DIM, NIP = ET<T>::NIP, CP = ET<T>::CP;
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n) return;
  double X[NEN][DIM], U[NEN][DIM];
  for (int b = 0; b < NEN; ++b)
    for (int d = 0; d < DIM; ++d) {
      X[b][d] = ecoords[(e * NEN + b) * DIM + d];
      U[b][d] = edisp[(e * NEN + b) * DIM + d];
    }
  double lam, mu;
  lame(E_nu[2 * e], E_nu[2 * e + 1], &lam, &mu);
  double sig[NIP][CP], sn[NEN][CP];
  ip_stress<T>(X, U, lam, mu, sig);
  nodal_from_ip<T>(sig, sn);
  for (int q = 0; q < NIP; ++q)
    for (int c = 0; c < CP; ++c) estress[(e * NIP + q) * CP + c] = sig[q][c];
  if (enodal)
    for (int a = 0; a < NEN; ++a)
      for (int c = 0; c < CP; ++c) enodal[(e * NEN + a) * CP + c] = sn[a][c];
}

// ---------------------------------------------------------------- stress, mesh level
// ONEIP (pure tri3 / tet4 mesh): the nodal value of a one-point element is its single stress, stored once per element
template <int T, bool ONEIP>
__global__ void __launch_bounds__(128) mesh_stress_kernel(DevMesh dm, const int32_t *__restrict__ elems, int n,
                                                          const double *__restrict__ uu, double *__restrict__ sn_out) {
  constexpr int NEN = ET<T>::NEN, DIM = ET<T>::DIM, NIP = ET<T>::NIP, CP = ET<T>::CP;
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n) return;
  const int e = elems[idx];
  double *out = sn_out + (ONEIP ? (int64_t)e : (int64_t)dm.eptr[e]) * CP;   // ONEIP: dense [nels][cp]
  if constexpr (T == FEMX_COH) {
    for (int q = 0; q < NEN * CP; ++q) out[q] = 0.0;  // D5: contribution defined as zero
  } else {
    double X[NEN][DIM], U[NEN][DIM];
    int nd[NEN];
    load_elem<T>(dm, e, X, nd);
#pragma unroll
    for (int a = 0; a < NEN; ++a)
#pragma unroll
      for (int d = 0; d < DIM; ++d) U[a][d] = uu[(int64_t)nd[a] * DIM + d];
    const int m = dm.emat[e] - 1;
    double sig[NIP][CP], sn[NEN][CP];
    ip_stress<T>(X, U, dm.lam_mu[2 * m], dm.lam_mu[2 * m + 1], sig);
    nodal_from_ip<T>(sig, sn);
#pragma unroll
    for (int a = 0; a < (ONEIP ? 1 : NEN); ++a)
#pragma unroll
      for (int c = 0; c < CP; ++c) out[a * CP + c] = sn[a][c];
  }
}

// hex8 form of mesh_stress_kernel: eight lanes per element.  Lane k first acts as integration point k (CalcElStress,
// src/m_local.F90:93-123: gradients at the point, strain from the element's displacements, sigma = D eps), then as local
// node k of the extrapolation (RecoverNodalStress, src/m_global.F90:488-515: sn[k] = sum_j N2inv[k][j] sigma[j], j ascending).
// The element's coordinates and displacements are gathered once (lane k fetches node k) and shared through shared memory,
// which then carries the eight stresses; the per-(element, node) results leave as one contiguous 384-byte run per element.
__device__ __forceinline__ void cp_async8(void *smem, const void *gmem) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((unsigned)__cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}
__global__ void __launch_bounds__(256, 2) hex_stress_ip_kernel(DevMesh dm, const int32_t *__restrict__ elems, int n, const double *__restrict__ uu,
                                                               double *__restrict__ sn_out) {
  constexpr int NEN = 8, CP = 6, IP = 9;             // IP: row pitch of the inverse table (8 lanes read 8 rows: distinct banks)
  constexpr int EP = NEN * 6 + 2;                    // + 2: the four elements of a warp start 4 banks apart (128-bit loads without conflicts)
  __shared__ __align__(16) double sh[8][2][4][EP];   // [warp][stage][element of the warp][node][x0 x1 x2 u0 u1 u2]
  __shared__ double inv_s[NEN * IP];
  if (threadIdx.x < 64) inv_s[(threadIdx.x >> 3) * IP + (threadIdx.x & 7)] = c_n2inv_hex[threadIdx.x];
  __syncthreads();
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, sub = lane >> 3, k = lane & 7;
  // a block walks a contiguous chunk of the element list, 32 elements per iteration: consecutive iterations touch
  // neighbouring elements, whose shared nodes are still in L1
  const int64_t chunk = ((((int64_t)n + gridDim.x - 1) / gridDim.x) + 31) / 32 * 32, stride = 32;
  const int64_t first = (int64_t)blockIdx.x * chunk + w * 4 + sub;
  n = (int)min((int64_t)n, ((int64_t)blockIdx.x + 1) * chunk);
  // the dependent chain element id -> connectivity offset -> node id -> coordinates / displacements is software-pipelined:
  // every level is fetched one iteration before the next level needs it (the last one by cp.async straight into the other
  // shared-memory stage), so a warp never waits for a load it just issued
  auto elem_at = [&](int64_t i) { return i < n ? elems[i] : -1; };
  auto fetch = [&](int stage, int nd) {   // node nd's coordinates and displacements -> this lane's slot of the stage
    double *dst = &sh[w][stage][sub][k * 6];
    if (nd >= 0) {
#pragma unroll
      for (int d = 0; d < 3; ++d) { cp_async8(dst + d, dm.coords + (int64_t)nd * 3 + d); cp_async8(dst + 3 + d, uu + (int64_t)nd * 3 + d); }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  int e3 = elem_at(first + 3 * stride), e2 = elem_at(first + 2 * stride), e1 = elem_at(first + stride), e0 = elem_at(first);
  int o2 = e2 >= 0 ? dm.eptr[e2] : 0, o1 = e1 >= 0 ? dm.eptr[e1] : 0, o0 = e0 >= 0 ? dm.eptr[e0] : 0;   // connectivity offsets
  int nd1 = e1 >= 0 ? dm.conn0[o1 + k] : -1;
  fetch(0, e0 >= 0 ? dm.conn0[o0 + k] : -1);
  int it = 0;
  for (int64_t idx = first; idx - sub < n; idx += stride, ++it) {   // warp-uniform trip count (idx - sub = the warp's first element)
    const bool on = idx < n;
    const double *S = &sh[w][it & 1][sub][0];
    double *Sw = &sh[w][it & 1][sub][0];
    const int64_t ecur = o0;
    const int m = on ? dm.emat[e0] - 1 : 0;
    // next levels (each consumed one iteration from now)
    fetch((it + 1) & 1, nd1);
    const int e4 = elem_at(idx + 4 * stride);
    const int o3 = e3 >= 0 ? dm.eptr[e3] : 0;
    nd1 = e2 >= 0 ? dm.conn0[o2 + k] : -1;
    e0 = e1; e1 = e2; e2 = e3; e3 = e4;
    o0 = o1; o1 = o2; o2 = o3;
    asm volatile("cp.async.wait_group 1;" ::: "memory");   // this iteration's stage has landed (the next one may be in flight)
    __syncwarp();
    const double lam = dm.lam_mu[2 * m], mu = dm.lam_mu[2 * m + 1];
    double out[CP];
    {
      double X[NEN][3];
#pragma unroll
      for (int a = 0; a < NEN; ++a)
#pragma unroll
        for (int d = 0; d < 3; ++d) X[a][d] = S[a * 6 + d];
      double g[NEN][3];
      phys_grad<FEMX_HEX>(X, k, g);
      double U[NEN][3];
#pragma unroll
      for (int a = 0; a < NEN; ++a)
#pragma unroll
        for (int d = 0; d < 3; ++d) U[a][d] = S[a * 6 + 3 + d];
      double H[3][3];  // H[i][j] = du_i/dx_j
#pragma unroll
      for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) {
          double t = 0.0;
#pragma unroll
          for (int a = 0; a < NEN; ++a) t += U[a][i] * g[a][j];
          H[i][j] = t;
        }
      const double tr = H[0][0] + H[1][1] + H[2][2];
#pragma unroll
      for (int i = 0; i < 3; ++i) out[i] = lam * tr + 2.0 * mu * H[i][i];
      out[3] = mu * (H[0][1] + H[1][0]);  // xy
      out[4] = mu * (H[1][2] + H[2][1]);  // yz
      out[5] = mu * (H[2][0] + H[0][2]);  // zx
    }
    __syncwarp();
#pragma unroll
    for (int c = 0; c < CP; ++c) Sw[k * 6 + c] = out[c];
    __syncwarp();
    double sn[CP];
#pragma unroll
    for (int c = 0; c < CP; ++c) sn[c] = 0.0;
#pragma unroll
    for (int j = 0; j < NEN; ++j) {
      const double f = inv_s[k * IP + j];
#pragma unroll
      for (int c = 0; c < CP; ++c) sn[c] += f * S[j * 6 + c];
    }
    if (on) {
      double *o = sn_out + (ecur + k) * CP;
#pragma unroll
      for (int c = 0; c < CP; ++c) o[c] = sn[c];
    }
    __syncwarp();   // the stage is rewritten by the fetch of the iteration after next
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
}

// nodal sum in ascending element order, then the average (src/main.F90:258-279).  One thread per (node, component);
// the incidence loop runs in batches of four so that the dependent loads (element id -> value) of a batch are in flight
// together while the adds keep the reference's order.  ONEIP: one stored stress per element, nen nodes per element.
template <bool ONEIP>
__global__ void __launch_bounds__(256) node_stress_gather_kernel(DevMesh dm, const double *__restrict__ sn, double *__restrict__ out,
                                                                 int col_major, int nen, int average) {
  const int cp = dm.cp;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= (int64_t)dm.nlnds * cp) return;
  const int64_t r = tid / cp;
  const int c = (int)(tid % cp);
  const int64_t p1 = dm.n2e_ptr[r + 1];
  int64_t p = dm.n2e_ptr[r];
  auto at = [&](int64_t q) -> const double * {
    if constexpr (ONEIP) return sn + (int64_t)dm.n2e_elem[q] * cp + c;
    else return sn + ((int64_t)dm.eptr[dm.n2e_elem[q]] + dm.n2e_a[q]) * cp + c;
  };
  double s = 0.0;
  for (; p + 4 <= p1; p += 4) {
    const double *a0 = at(p), *a1 = at(p + 1), *a2 = at(p + 2), *a3 = at(p + 3);
    const double v0 = *a0, v1 = *a1, v2 = *a2, v3 = *a3;
    s += v0; s += v1; s += v2; s += v3;
  }
  for (; p < p1; ++p) s += *at(p);
  if (average) s = s / (double)dm.node_count[r];   // multi-rank hosts divide by the global count themselves
  out[col_major ? (int64_t)c * dm.nlnds + r : tid] = s;
}

// ---------------------------------------------------------------- pattern / BC / SpMV
template <int PD>
__global__ void colidx_kernel(DevMesh dm, int32_t *__restrict__ colidx) {
  const int64_t blk = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t nblk = dm.bptr[dm.nlnds];
  if (blk >= nblk) return;
  // find the row by bisection over bptr
  int lo = 0, hi = dm.nlnds;
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (dm.bptr[mid] <= blk) lo = mid; else hi = mid;
  }
  const int64_t b0 = dm.bptr[lo];
  const int nb = (int)(dm.bptr[lo + 1] - b0), k = (int)(blk - b0);
  const int gcol = PD * (dm.nl2g[dm.bcol[blk]] - 1);
  for (int i = 0; i < PD; ++i)
    for (int j = 0; j < PD; ++j) colidx[PD * PD * b0 + (int64_t)i * PD * nb + PD * k + j] = gcol + j;
}

// plan-time check for the hex8 tile kernels: number of (element, integration point) pairs with detJ <= 0
__global__ void hex_detj_kernel(DevMesh dm, int *__restrict__ bad) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= dm.nels * 8) return;
  const int e = t >> 3, ip = t & 7;
  double X[8][3];
  const int32_t *nd = dm.conn0 + dm.eptr[e];
#pragma unroll
  for (int b = 0; b < 8; ++b)
#pragma unroll
    for (int d = 0; d < 3; ++d) X[b][d] = dm.coords[(size_t)nd[b] * 3 + d];
  double g[8][3];
  const double wdet = phys_grad<FEMX_HEX>(X, ip, g);
  if (!(wdet > 0.0)) atomicAdd(bad, 1);
}

// COO (i, j) of every value in femx_get_pattern order: the arrays MatSetPreallocationCOO takes
template <int PD>
__global__ void coo_kernel(DevMesh dm, int32_t *__restrict__ coo_i, int32_t *__restrict__ coo_j) {
  const int64_t blk = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t nblk = dm.bptr[dm.nlnds];
  if (blk >= nblk) return;
  int lo = 0, hi = dm.nlnds;
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (dm.bptr[mid] <= blk) lo = mid; else hi = mid;
  }
  const int64_t b0 = dm.bptr[lo];
  const int nb = (int)(dm.bptr[lo + 1] - b0), k = (int)(blk - b0);
  const int grow = PD * (dm.nl2g[lo] - 1), gcol = PD * (dm.nl2g[dm.bcol[blk]] - 1);
  for (int i = 0; i < PD; ++i)
    for (int j = 0; j < PD; ++j) {
      const int64_t s = PD * PD * b0 + (int64_t)i * PD * nb + PD * k + j;
      coo_i[s] = grow + i;
      coo_j[s] = gcol + j;
    }
}

__global__ void apply_kbc_kernel(int n, const int64_t *__restrict__ diag, double *__restrict__ vals) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) vals[diag[i]] = kPenalty;  // MatSetValue(j, j, 1e30, INSERT), src/m_global.F90:296
}

__global__ void bc_force_kernel(int n, const int64_t *__restrict__ dof, const double *__restrict__ val,
                                double dt, double t_total, double *__restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  // vvec = PENALTY_PARAM * bcval * dt / t, INSERT (src/m_global.F90:451), evaluated left to right
  if (i < n) out[dof[i]] = __ddiv_rn(__dmul_rn(__dmul_rn(kPenalty, val[i]), dt), t_total);
}

// ---- FormRHS (src/m_global.F90:373-436).  Explicit _rn intrinsics: no FMA contraction, the sums round like the
// reference's scalar code.
// MatDet (src/m_utils.F90:60-66) of the matrices TriArea builds (:25-33): columns a, b, (1, 1, 1)
__device__ __forceinline__ double det3_ones(double a1, double a2, double a3, double b1, double b2, double b3) {
  const double t1 = __dmul_rn(a1, __dsub_rn(b2, b3)), t2 = __dmul_rn(b1, __dsub_rn(a2, a3));
  const double t3 = __dsub_rn(__dmul_rn(a2, b3), __dmul_rn(b2, a3));
  return __dadd_rn(__dsub_rn(t1, t2), t3);
}
__device__ __forceinline__ double tri_area(const double *p1, const double *p2, const double *p3) {
  const double d1 = fabs(det3_ones(p1[0], p2[0], p3[0], p1[1], p2[1], p3[1]));
  const double d2 = fabs(det3_ones(p1[1], p2[1], p3[1], p1[2], p2[2], p3[2]));
  const double d3 = fabs(det3_ones(p1[2], p2[2], p3[2], p1[0], p2[0], p3[0]));
  return __dmul_rn(0.5, __dsqrt_rn(__dadd_rn(__dadd_rn(__dmul_rn(d1, d1), __dmul_rn(d2, d2)), __dmul_rn(d3, d3))));
}
// EdgeAreaNodes{Tri,Qua,Tet,Hex} (src/m_elems.F90:282-293, 366-378, 469-482, 583-599): length / area of a side
__device__ double side_area(const DevMesh &dm, int e, int side, int *nps) {
  const int t = dm.etype[e];
  const int32_t *cn = dm.conn0 + dm.eptr[e];
  if (t == FEMX_TRI || t == FEMX_QUA) {
    const int nen = t == FEMX_TRI ? 3 : 4, a = side - 1, b = side % nen;
    const double *x = dm.coords + 2 * (int64_t)cn[a], *y = dm.coords + 2 * (int64_t)cn[b];
    const double dx = __dsub_rn(x[0], y[0]), dy = __dsub_rn(x[1], y[1]);
    *nps = 2;
    return __dsqrt_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)));
  }
  if (t == FEMX_TET) {
    const int j[4][3] = {{0, 1, 3}, {1, 2, 3}, {0, 2, 3}, {0, 1, 2}};
    *nps = 3;
    return tri_area(dm.coords + 3 * (int64_t)cn[j[side - 1][0]], dm.coords + 3 * (int64_t)cn[j[side - 1][1]],
                    dm.coords + 3 * (int64_t)cn[j[side - 1][2]]);
  }
  const int j[6][4] = {{0, 1, 5, 4}, {1, 2, 6, 5}, {2, 3, 7, 6}, {3, 0, 4, 7}, {0, 1, 2, 3}, {4, 5, 6, 7}};
  const double *p0 = dm.coords + 3 * (int64_t)cn[j[side - 1][0]], *p1 = dm.coords + 3 * (int64_t)cn[j[side - 1][1]];
  const double *p2 = dm.coords + 3 * (int64_t)cn[j[side - 1][2]], *p3 = dm.coords + 3 * (int64_t)cn[j[side - 1][3]];
  *nps = 4;
  return __dadd_rn(fabs(tri_area(p0, p1, p2)), fabs(tri_area(p0, p2, p3)));   // QuadArea, src/m_utils.F90:36-42
}

// one thread per loaded dof; contributions added in the reference's call order (nodal forces, then tractions)
__global__ void form_rhs_kernel(DevMesh dm, DevLoads L, double t_init, double t_end, double dt, double *__restrict__ f) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= L.ntouched) return;
  const int pd = dm.pdim, w = pd + 2;
  double s = 0.0;   // VecZeroEntries
  for (int c = L.ptr[i]; c < L.ptr[i + 1]; ++c) {
    const int code = L.src[c], comp = code & 3, idx = code >> 3;
    if (!(code & 4)) {   // nodal force: fval * applied_interval / (t2 - t1), src/m_global.F90:399-409
      const double *fv = L.fval + (int64_t)idx * w;
      const double t1 = fv[pd], t2 = fv[pd + 1];
      if (t_end < t1 || t_init > t2) continue;
      const double interval = __dsub_rn(fmin(t2, t_end), fmax(t1, t_init));
      s = __dadd_rn(s, __ddiv_rn(__dmul_rn(fv[comp], interval), __dsub_rn(t2, t1)));
    } else {             // traction: tval * (applied_interval / (t2 - t1)), t1 and t2 divided by dt (:414-415), then * area / nps (:360)
      const double *tv = L.tval + (int64_t)idx * w;
      const double t1 = __ddiv_rn(tv[pd], dt), t2 = __ddiv_rn(tv[pd + 1], dt);
      if (t_end < t1 || t_init > t2) continue;
      const double interval = __dsub_rn(fmin(t2, t_end), fmax(t1, t_init));
      int nps;
      const double area = side_area(dm, L.tel[2 * idx], L.tel[2 * idx + 1], &nps);
      const double v = __dmul_rn(tv[comp], __ddiv_rn(interval, __dsub_rn(t2, t1)));
      s = __dadd_rn(s, __ddiv_rn(__dmul_rn(v, area), (double)nps));
    }
  }
  f[L.dof[i]] = s;
}

// y[g] = sum_k J[row, k] * x[col] - sub[g] over the local node rows (MatMultAdd, src/m_global.F90:225-226).
// G lanes own one node: its PD matrix rows are PD consecutive runs of rowlen = PD * nb values, read lane <-> consecutive
// value (coalesced); 32 / G nodes per warp keep several dependent column -> x gathers in flight.  Fixed summation order:
// lane partials over v = gl, gl + G, ..., then a butterfly over the group.
template <int PD, int G>
__global__ void __launch_bounds__(256) spmv_sub_kernel(DevMesh dm, const double *__restrict__ vals,
                                                       const double *__restrict__ x, const double *__restrict__ sub,
                                                       double *__restrict__ y, int ident) {
  constexpr int NPW = 32 / G;
  const int lane = threadIdx.x & 31, gl = lane % G, gi = lane / G;
  const int64_t r = (((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5) * NPW + gi;
  const bool on = r < dm.nlnds;
  double acc[PD];
#pragma unroll
  for (int i = 0; i < PD; ++i) acc[i] = 0.0;
  if (on) {
    const int64_t b0 = dm.bptr[r];
    const int nb = (int)(dm.bptr[r + 1] - b0), rowlen = PD * nb;
    const double *row = vals + PD * PD * b0;
    for (int rem = gl; rem < rowlen; rem += G) {
      const int k = rem / PD, j = rem - k * PD;
      const int col = dm.bcol[b0 + k];
      const int64_t gc = (int64_t)PD * (ident ? col : dm.nl2g[col] - 1) + j;
      const double xv = x[gc];
#pragma unroll
      for (int i = 0; i < PD; ++i) acc[i] = fma(row[(int64_t)i * rowlen + rem], xv, acc[i]);
    }
  }
#pragma unroll
  for (int o = G / 2; o > 0; o >>= 1)
#pragma unroll
    for (int i = 0; i < PD; ++i) acc[i] += __shfl_xor_sync(0xffffffffu, acc[i], o);
  if (on && gl < PD) {
    const int64_t g = (int64_t)PD * (dm.nl2g[r] - 1) + gl;
    y[g] = (gl == 0 ? acc[0] : (gl == 1 ? acc[1] : acc[PD - 1])) - sub[g];
  }
}

// ---------------------------------------------------------------- cohesive
__global__ void cohesive_mesh_kernel(DevMesh dm, const int32_t *__restrict__ coh_elems, int n,
                                     const double *__restrict__ u, double dt, int quirks,
                                     double *__restrict__ force, double *__restrict__ stiff) {
  const int ci = blockIdx.x * blockDim.x + threadIdx.x;
  if (ci >= n) return;
  const int e = coh_elems[ci];
  const int32_t *c = dm.conn0 + dm.eptr[e];
  double X[4][2], U[4][2];
  for (int a = 0; a < 4; ++a) {
    // D3: the reference reads the rank's local nodes 1..4 (src/m_elems.F90:700-702)
    const int un = (quirks & FEMX_Q_GATHER_LOCAL14) ? a : c[a];
    for (int d = 0; d < 2; ++d) {
      X[a][d] = dm.coords[(int64_t)c[a] * 2 + d];
      U[a][d] = u[(int64_t)un * 2 + d];
    }
  }
  cohesive_element(X, U, dm.cohprops + 6 * (dm.enlmat[e] - 1), dt, quirks, force ? force + 8 * ci : nullptr,
                   stiff ? stiff + 64 * ci : nullptr);
}

__global__ void cohesive_unit_kernel(int n, const double *__restrict__ ecoords, const int32_t *__restrict__ nodes,
                                     const double *__restrict__ u, const double *__restrict__ props, double dt,
                                     int quirks, double *__restrict__ force, double *__restrict__ stiff) {
  const int ci = blockIdx.x * blockDim.x + threadIdx.x;
  if (ci >= n) return;
  double X[4][2], U[4][2];
  for (int a = 0; a < 4; ++a) {
    const int un = (quirks & FEMX_Q_GATHER_LOCAL14) ? a : nodes[4 * ci + a] - 1;
    for (int d = 0; d < 2; ++d) {
      X[a][d] = ecoords[(ci * 4 + a) * 2 + d];
      U[a][d] = u[(int64_t)un * 2 + d];
    }
  }
  cohesive_element(X, U, props + 6 * ci, dt, quirks, force + 8 * ci, stiff + 64 * ci);
}

// J[slot] = K[slot] + sum of cohesive tangents in ascending element order, per touched block
__global__ void jacobian_overlay_kernel(DevMesh dm, int n, const int64_t *__restrict__ blk_id,
                                        const int32_t *__restrict__ blk_row, const int32_t *__restrict__ ptr,
                                        const int32_t *__restrict__ src, const double *__restrict__ coh_ke,
                                        const double *__restrict__ kvals, double *__restrict__ jvals) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  const int r = blk_row[t];
  const int64_t b0 = dm.bptr[r];
  const int nb = (int)(dm.bptr[r + 1] - b0), k = (int)(blk_id[t] - b0);
  for (int i = 0; i < 2; ++i)
    for (int j = 0; j < 2; ++j) {
      const int64_t slot = 4 * b0 + (int64_t)i * 2 * nb + 2 * k + j;
      double s = kvals[slot];
      for (int q = ptr[t]; q < ptr[t + 1]; ++q) {
        const int ci = src[q] >> 8, a = (src[q] >> 4) & 15, b = src[q] & 15;
        s += coh_ke[64 * ci + (a * 2 + i) * 8 + b * 2 + j];
      }
      jvals[slot] = s;
    }
}

// out[g] += sum of cohesive nodal forces in ascending element order (src/m_global.F90:203-205)
__global__ void cohesive_force_gather_kernel(DevMesh dm, int n, const int32_t *__restrict__ node_id,
                                             const int32_t *__restrict__ ptr, const int32_t *__restrict__ src,
                                             const double *__restrict__ coh_fe, double *__restrict__ out) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  const int64_t g = 2 * (int64_t)(dm.nl2g[node_id[t]] - 1);
  for (int d = 0; d < 2; ++d) {
    double s = out[g + d];
    for (int q = ptr[t]; q < ptr[t + 1]; ++q) s += coh_fe[8 * (src[q] >> 4) + (src[q] & 15) * 2 + d];
    out[g + d] = s;
  }
}

// ordered merge of own + received contributions per owned nonzero (receiver adds in
// (source rank, position) order)
__global__ void gather_sum_kernel(const double *__restrict__ a, int64_t na, const double *__restrict__ b,
                                  const int64_t *__restrict__ ptr, const int64_t *__restrict__ idx, int64_t nout,
                                  double *__restrict__ out) {
  const int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= nout) return;
  double v = 0.0;
  for (int64_t q = ptr[s]; q < ptr[s + 1]; ++q) {
    const int64_t i = idx[q];
    v += i < na ? a[i] : b[i - na];
  }
  out[s] = v;
}

// ================================================================= launchers
static inline int nblocks(int64_t n, int bs) { return n > 0 ? (int)((n + bs - 1) / bs) : 1; }   // an empty mesh launches one block that finds nothing to do

void set_n2inv(const double *qua16, const double *hex64) {
  cudaMemcpyToSymbol(c_n2inv_qua, qua16, sizeof(double) * 16);
  cudaMemcpyToSymbol(c_n2inv_hex, hex64, sizeof(double) * 64);
}

int launch_elem_rows(femx_ctx *c, const double *d_u, double *d_ke, double *d_fe) {
  int launches = 0;
  for (auto &b : c->buckets) {
    if (!b.n) continue;
    const int64_t nt = (int64_t)b.n * type_nen(b.type);
    const int bs = 128, g = nblocks(nt, bs);
    switch (b.type) {
      case FEMX_TRI: elem_rows_kernel<FEMX_TRI><<<g, bs, 0, c->stream>>>(c->dm, b.d_elem, b.n, d_u, d_ke, d_fe); break;
      case FEMX_QUA: elem_rows_kernel<FEMX_QUA><<<g, bs, 0, c->stream>>>(c->dm, b.d_elem, b.n, d_u, d_ke, d_fe); break;
      case FEMX_TET: elem_rows_kernel<FEMX_TET><<<g, bs, 0, c->stream>>>(c->dm, b.d_elem, b.n, d_u, d_ke, d_fe); break;
      case FEMX_HEX: elem_rows_kernel<FEMX_HEX><<<g, bs, 0, c->stream>>>(c->dm, b.d_elem, b.n, d_u, d_ke, d_fe); break;
      case FEMX_COH: elem_rows_kernel<FEMX_COH><<<g, bs, 0, c->stream>>>(c->dm, b.d_elem, b.n, d_u, d_ke, d_fe); break;
    }
    ++launches;
  }
  return launches;
}

int launch_gather_rows(femx_ctx *c, const double *d_ke, const double *d_fe, double *d_vals, double *d_fint) {
  const int bs = 256, g = nblocks((int64_t)c->nlnds * 32, bs);
  if (c->pdim == 2) gather_rows_kernel<2><<<g, bs, 0, c->stream>>>(c->dm, d_ke, d_fe, d_vals, d_fint);
  else gather_rows_kernel<3><<<g, bs, 0, c->stream>>>(c->dm, d_ke, d_fe, d_vals, d_fint);
  return 1;
}

int launch_element_ke(femx_ctx *c, int type, int n, const double *d_x, const double *d_enu, double *d_out) {
  const int bs = 128, g = nblocks((int64_t)n * type_nen(type), bs);
  switch (type) {
    case FEMX_TRI: element_ke_kernel<FEMX_TRI><<<g, bs, 0, c->stream>>>(n, d_x, d_enu, d_out); break;
    case FEMX_QUA: element_ke_kernel<FEMX_QUA><<<g, bs, 0, c->stream>>>(n, d_x, d_enu, d_out); break;
    case FEMX_TET: element_ke_kernel<FEMX_TET><<<g, bs, 0, c->stream>>>(n, d_x, d_enu, d_out); break;
    case FEMX_HEX: element_ke_kernel<FEMX_HEX><<<g, bs, 0, c->stream>>>(n, d_x, d_enu, d_out); break;
    default: return 0;
  }
  return 1;
}

int launch_element_stress(femx_ctx *c, int type, int n, const double *d_x, const double *d_u, const double *d_enu,
                          double *d_sig, double *d_sn) {
  const int bs = 128, g = nblocks(n, bs);
  switch (type) {
    case FEMX_TRI: element_stress_kernel<FEMX_TRI><<<g, bs, 0, c->stream>>>(n, d_x, d_u, d_enu, d_sig, d_sn); break;
    case FEMX_QUA: element_stress_kernel<FEMX_QUA><<<g, bs, 0, c->stream>>>(n, d_x, d_u, d_enu, d_sig, d_sn); break;
    case FEMX_TET: element_stress_kernel<FEMX_TET><<<g, bs, 0, c->stream>>>(n, d_x, d_u, d_enu, d_sig, d_sn); break;
    case FEMX_HEX: element_stress_kernel<FEMX_HEX><<<g, bs, 0, c->stream>>>(n, d_x, d_u, d_enu, d_sig, d_sn); break;
    default: return 0;
  }
  return 1;
}

int launch_mesh_stress(femx_ctx *c, const double *d_uu, double *d_sn, double *d_out, int col_major, int average) {
  int launches = 0;
  bool oneip = !c->buckets.empty();   // pure tri3 or pure tet4 mesh: one stress per element, uniform element size
  for (auto &b : c->buckets) oneip = oneip && (b.n == 0 || b.type == FEMX_TRI || b.type == FEMX_TET);
  for (auto &b : c->buckets) {
    if (!b.n) continue;
    const int bs = 128, g = nblocks(b.n, bs);
    switch (b.type) {
      case FEMX_TRI:
        if (oneip) mesh_stress_kernel<FEMX_TRI, true><<<g, bs, 0, c->stream>>>(c->dm, b.d_elem, b.n, d_uu, d_sn);
        else mesh_stress_kernel<FEMX_TRI, false><<<g, bs, 0, c->stream>>>(c->dm, b.d_elem, b.n, d_uu, d_sn);
        break;
      case FEMX_QUA: mesh_stress_kernel<FEMX_QUA, false><<<g, bs, 0, c->stream>>>(c->dm, b.d_elem, b.n, d_uu, d_sn); break;
      case FEMX_TET:
        if (oneip) mesh_stress_kernel<FEMX_TET, true><<<g, bs, 0, c->stream>>>(c->dm, b.d_elem, b.n, d_uu, d_sn);
        else mesh_stress_kernel<FEMX_TET, false><<<g, bs, 0, c->stream>>>(c->dm, b.d_elem, b.n, d_uu, d_sn);
        break;
      case FEMX_HEX: hex_stress_ip_kernel<<<std::min(nblocks(b.n, 32), 148 * 2 * 8), 256, 0, c->stream>>>(c->dm, b.d_elem, b.n, d_uu, d_sn); break;
      case FEMX_COH: mesh_stress_kernel<FEMX_COH, false><<<g, bs, 0, c->stream>>>(c->dm, b.d_elem, b.n, d_uu, d_sn); break;
    }
    ++launches;
  }
  const int bs = 256, g = nblocks((int64_t)c->nlnds * c->dm.cp, bs);
  if (oneip) node_stress_gather_kernel<true><<<g, bs, 0, c->stream>>>(c->dm, d_sn, d_out, col_major, c->pdim + 1, average);
  else node_stress_gather_kernel<false><<<g, bs, 0, c->stream>>>(c->dm, d_sn, d_out, col_major, 0, average);
  return launches + 1;
}

int launch_colidx(femx_ctx *c, int32_t *d_colidx) {
  const int bs = 256, g = nblocks(c->plan.nblk, bs);
  if (!g) return 0;
  if (c->pdim == 2) colidx_kernel<2><<<g, bs, 0, c->stream>>>(c->dm, d_colidx);
  else colidx_kernel<3><<<g, bs, 0, c->stream>>>(c->dm, d_colidx);
  return 1;
}

int launch_coo(femx_ctx *c, int32_t *d_i, int32_t *d_j) {
  const int bs = 256, g = nblocks(c->plan.nblk, bs);
  if (!g) return 0;
  if (c->pdim == 2) coo_kernel<2><<<g, bs, 0, c->stream>>>(c->dm, d_i, d_j);
  else coo_kernel<3><<<g, bs, 0, c->stream>>>(c->dm, d_i, d_j);
  return 1;
}

int launch_hex_detj(femx_ctx *c, int *d_bad) {
  const int n = c->nels * 8;
  if (!n) return 0;
  hex_detj_kernel<<<nblocks(n, 128), 128, 0, c->stream>>>(c->dm, d_bad);
  return 1;
}

int launch_apply_kbc(femx_ctx *c, double *d_vals) {
  if (!c->nbc_dofs) return 0;
  apply_kbc_kernel<<<nblocks(c->nbc_dofs, 128), 128, 0, c->stream>>>(c->nbc_dofs, c->d_bc_diag, d_vals);
  return 1;
}

int launch_bc_force(femx_ctx *c, double dt, double t_total, double *d_out) {
  if (!c->nbc_dofs) return 0;
  bc_force_kernel<<<nblocks(c->nbc_dofs, 128), 128, 0, c->stream>>>(c->nbc_dofs, c->d_bc_dof, c->d_bc_val, dt, t_total, d_out);
  return 1;
}

int launch_form_rhs(femx_ctx *c, double t_init, double dt, double *d_f) {
  if (!c->loads.ntouched) return 0;
  form_rhs_kernel<<<nblocks(c->loads.ntouched, 128), 128, 0, c->stream>>>(c->dm, c->loads, t_init, t_init + dt, dt, d_f);
  return 1;
}

int launch_spmv_sub(femx_ctx *c, const double *d_vals, const double *d_x, const double *d_sub, double *d_y) {
  const int bs = 256;
  const bool ident = c->nl2g_identity;   // single rank: skip the nl2g indirection
  if (c->pdim == 2) {   // quad / tri rows hold ~9 blocks = 18 values per run: 8 lanes per node
    spmv_sub_kernel<2, 8><<<nblocks(((int64_t)c->nlnds + 3) / 4 * 32, bs), bs, 0, c->stream>>>(c->dm, d_vals, d_x, d_sub, d_y, ident);
  } else {              // tet ~15 blocks = 45 values, hex 27 blocks = 81 values per run: 16 lanes per node
    spmv_sub_kernel<3, 16><<<nblocks(((int64_t)c->nlnds + 1) / 2 * 32, bs), bs, 0, c->stream>>>(c->dm, d_vals, d_x, d_sub, d_y, ident);
  }
  return 1;
}

int launch_cohesive_mesh(femx_ctx *c, const double *d_u, double dt, double *d_force, double *d_stiff) {
  const int n = (int)c->plan.coh_elems.size();
  if (!n) return 0;
  cohesive_mesh_kernel<<<nblocks(n, 64), 64, 0, c->stream>>>(c->dm, c->d_coh_elems, n, d_u, dt, c->opts.quirks, d_force, d_stiff);
  return 1;
}

int launch_cohesive_unit(femx_ctx *c, int n, const double *d_x, const int32_t *d_nodes, const double *d_u,
                         const double *d_props, double dt, double *d_force, double *d_stiff) {
  cohesive_unit_kernel<<<nblocks(n, 64), 64, 0, c->stream>>>(n, d_x, d_nodes, d_u, d_props, dt, c->opts.quirks, d_force, d_stiff);
  return 1;
}

int launch_jacobian_overlay(femx_ctx *c, const double *d_kvals, double *d_jvals) {
  const int n = (int)c->plan.cblk_id.size();
  if (!n) return 0;
  jacobian_overlay_kernel<<<nblocks(n, 128), 128, 0, c->stream>>>(c->dm, n, c->d_cblk_id, c->d_cblk_row, c->d_cblk_ptr,
                                                                 c->d_cblk_src, c->d_coh_ke, d_kvals, d_jvals);
  return 1;
}

int launch_gather_sum(femx_ctx *c, const double *a, int64_t na, const double *b, const int64_t *ptr, const int64_t *idx,
                      int64_t nout, double *out) {
  if (nout <= 0) return 0;
  gather_sum_kernel<<<nblocks(nout, 256), 256, 0, c->stream>>>(a, na, b, ptr, idx, nout, out);
  return 1;
}

int launch_cohesive_force_gather(femx_ctx *c, double *d_out) {
  const int n = (int)c->plan.cnode_id.size();
  if (!n) return 0;
  cohesive_force_gather_kernel<<<nblocks(n, 128), 128, 0, c->stream>>>(c->dm, n, c->d_cnode_id, c->d_cnode_ptr,
                                                                      c->d_cnode_src, c->d_coh_fe, d_out);
  return 1;
}

}  // namespace femx
