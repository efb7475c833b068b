// rows_simplex.cu -- assembly variant 11: fused gather-ordered assembly for pure tri3 / tet4 meshes (one integration
// point per element), the lean form of rows_fused.cu.
//
// A simplex element's stiffness is an outer product of its constant shape-function gradients:
//   K_ab(i,j) = w detJ (lambda g_a,i g_b,j + mu g_a,j g_b,i + delta_ij mu g_a . g_b)      (FormElKE, src/m_local.F90:61-65)
// so instead of parking the 4 x 9 values of a row block per (node, incident element) in shared memory (39 doubles) and
// reading every one of them back once per CSR value, the incident lanes park the GRADIENTS (12 doubles + 2
// coefficients) and the CSR blocks are formed where they are summed:
//   phase A  lane i of the node's group: gradients of incident element i -> shared memory slot (15 doubles, odd pitch);
//   phase B  lane k: block k of the node's row = sum over the contributing (element, a, b) of the plan's list, in
//            ascending element order (MatSetValues ADD semantics, src/main.F90:157-171, without atomics); the diagonal
//            block (one contribution per incident element) is cut into chunks taken by otherwise idle lanes and
//            recombined in chunk order; the lane multiplies its block with u of the column node and the group
//            butterfly-sums the internal force.
// Per contribution: 8 shared loads + ~30 FP64 operations for 9 matrix entries (rows_fused.cu: 9 loads + 9 index
// loads + 9 adds), no per-value index arithmetic.
#include "elem_math.cuh"
#include "launch.h"

namespace femx {

namespace {

constexpr int kWarps = 8;
constexpr int kChunks = 4;   // the diagonal block is summed by up to this many lanes

template <int PD, int G, bool PEER>
__global__ void __launch_bounds__(kWarps * 32, 3)
rows_simplex_kernel(DevMesh dm, const double *__restrict__ u, double *__restrict__ vals, double *__restrict__ fint, const PeerTab pt) {
  constexpr int T = PD == 2 ? FEMX_TRI : FEMX_TET;
  constexpr int NEN = PD + 1, D2 = PD * PD, NPW = 32 / G;
  constexpr int SLOT = (NEN * PD + 2) | 1;          // gradients, w detJ lambda, w detJ mu; odd pitch
  constexpr int DPART = kChunks * D2;               // diagonal partials of one group
  __shared__ double sm_slot[kWarps * 32 * SLOT];
  __shared__ double sm_diag[kWarps * NPW * DPART];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int gl = lane % G, gi = lane / G;
  double *mine = sm_slot + ((size_t)w * 32 + lane) * SLOT;
  const double *grp = sm_slot + ((size_t)w * 32 + gi * G) * SLOT;
  double *dpart = sm_diag + ((size_t)w * NPW + gi) * DPART;
  const int64_t ngroups = ((int64_t)dm.nlnds + NPW - 1) / NPW;
  const int64_t stride = (int64_t)gridDim.x * kWarps;
  // row metadata of the NEXT group is fetched while the current one is processed (one level of the dependent-load
  // chain off the critical path); the incidence records hold the element's nodes directly (plan: sx_inc)
  struct Meta { int64_t p0, b0; int ninc, nb, kd; };
  auto fetch = [&](int64_t g) {
    Meta m{0, 0, 0, 0, 0};
    const int64_t r = g * NPW + gi;
    if (g < ngroups && r < dm.nlnds) {
      m.p0 = dm.n2e_ptr[r]; m.ninc = (int)(dm.n2e_ptr[r + 1] - m.p0);
      m.b0 = dm.bptr[r]; m.nb = (int)(dm.bptr[r + 1] - m.b0);
      m.kd = dm.bdiag[r];
    }
    return m;
  };
  int64_t wg = (int64_t)blockIdx.x * kWarps + w;
  Meta nxt = fetch(wg);
  for (; wg < ngroups; wg += stride) {
    const int64_t r = wg * NPW + gi;
    const bool node_on = r < dm.nlnds;
    const Meta cur = nxt;
    const int ninc = cur.ninc;
    // ---- phase A: gradients of one incident element per lane
    int4 nd4 = make_int4(0, 0, 0, 0);
    int mat = 0;
    if (gl < ninc) {
      nd4 = *reinterpret_cast<const int4 *>(dm.sx_inc + 4 * (cur.p0 + gl));
      mat = dm.sx_mat[cur.p0 + gl];
    }
    // first task of the lane in phase B: its list bounds travel together with the incidence record
    const uint16_t *off = dm.rb_off + cur.b0 + r;
    int t_off0 = 0, t_off1 = 0, d_off0 = 0, d_off1 = 0;
    if (node_on) {
      if (gl < cur.nb) { t_off0 = off[gl]; t_off1 = off[gl + 1]; }
      d_off0 = off[cur.kd]; d_off1 = off[cur.kd + 1];
    }
    nxt = fetch(wg + stride);
    if (gl < ninc) {
      const int nds[4] = {nd4.x, nd4.y, nd4.z, nd4.w};
      double X[NEN][PD];
#pragma unroll
      for (int b = 0; b < NEN; ++b)
#pragma unroll
        for (int d = 0; d < PD; ++d) X[b][d] = dm.coords[(int64_t)nds[b] * PD + d];
      const double lam = dm.lam_mu[2 * mat], mu = dm.lam_mu[2 * mat + 1];
      double g[NEN][PD];
      const double wd = phys_grad<T>(X, 0, g);
#pragma unroll
      for (int b = 0; b < NEN; ++b)
#pragma unroll
        for (int d = 0; d < PD; ++d) mine[b * PD + d] = g[b][d];
      mine[NEN * PD] = wd * lam;
      mine[NEN * PD + 1] = wd * mu;
    }
    __syncwarp();
    // ---- phase B: one block (or one chunk of the diagonal block) per lane
    double fpart[PD];
#pragma unroll
    for (int i = 0; i < PD; ++i) fpart[i] = 0.0;
    int nchunk = 1;
    if (node_on) {
      const int64_t b0 = cur.b0;
      const int nb = cur.nb;
      const int rowlen = PD * nb;
      const uint16_t *ent = dm.sx_ent + NEN * cur.p0;   // every element lists NEN contributions: rbe_ptr = NEN * n2e_ptr
      const int kd = cur.kd;
      // diagonal chunks: as many as there are idle lanes in the last pass over the blocks, at most kChunks
      const int spare = (G - nb % G) % G;
      nchunk = min(kChunks, 1 + spare);
      const int dlo = d_off0, dn = d_off1 - dlo, clen = (dn + nchunk - 1) / nchunk;
      double *prow[PD];
#pragma unroll
      for (int i = 0; i < PD; ++i) prow[i] = vals + (D2 * b0 + (int64_t)i * rowlen);
      if constexpr (PEER) {
        if (!peer_local(pt, D2 * b0, D2 * b0 + (int64_t)PD * rowlen)) {
#pragma unroll
          for (int i = 0; i < PD; ++i) prow[i] = peer_row(pt, vals, D2 * b0 + (int64_t)i * rowlen);
        }
      }
      for (int t = gl; t < nb + nchunk - 1; t += G) {
        // task t: block t, or chunk (t - nb + 1) of the diagonal block
        const int k = t < nb ? t : kd;
        int c0, c1, chunk = -1;
        if (t == gl) { c0 = t_off0; c1 = t_off1; } else { c0 = off[k]; c1 = off[k + 1]; }
        if (k == kd) {
          chunk = t < nb ? 0 : t - nb + 1;
          c0 = dlo + chunk * clen;
          c1 = min(dlo + dn, c0 + clen);
        }
        double acc[D2], tr = 0.0;
#pragma unroll
        for (int q = 0; q < D2; ++q) acc[q] = 0.0;
        for (int c = c0; c < c1; ++c) {
          const int en = ent[c];
          const double *sl = grp + (en >> 4) * SLOT;
          const double *ga = sl + ((en >> 2) & 3) * PD, *gb = sl + (en & 3) * PD;
          const double cl = sl[NEN * PD], cm = sl[NEN * PD + 1];
          double la[PD], ma[PD], vb[PD];
#pragma unroll
          for (int i = 0; i < PD; ++i) { const double a = ga[i]; la[i] = cl * a; ma[i] = cm * a; vb[i] = gb[i]; }
#pragma unroll
          for (int i = 0; i < PD; ++i) tr = fma(ma[i], vb[i], tr);
#pragma unroll
          for (int i = 0; i < PD; ++i)
#pragma unroll
            for (int j = 0; j < PD; ++j) acc[i * PD + j] = fma(ma[j], vb[i], fma(la[i], vb[j], acc[i * PD + j]));
        }
#pragma unroll
        for (int i = 0; i < PD; ++i) acc[i * PD + i] += tr;   // delta_ij mu g_a . g_b, summed over the contributions
        if (fint) {   // this block's (or chunk's) share of f_r = sum_k K_rk u_k
          const int64_t col = dm.bcol[b0 + k];
#pragma unroll
          for (int j = 0; j < PD; ++j) {
            const double uj = u[col * PD + j];
#pragma unroll
            for (int i = 0; i < PD; ++i) fpart[i] += acc[i * PD + j] * uj;
          }
        }
        if (chunk < 0 || nchunk == 1) {
#pragma unroll
          for (int i = 0; i < PD; ++i)
#pragma unroll
            for (int j = 0; j < PD; ++j) prow[i][PD * k + j] = acc[i * PD + j];
        } else {
#pragma unroll
          for (int q = 0; q < D2; ++q) dpart[chunk * D2 + q] = acc[q];
        }
      }
    }
    __syncwarp();
    if (node_on && nchunk > 1) {   // diagonal block = its chunks in order, one entry per lane
      const int64_t b0 = cur.b0;
      const int nb = cur.nb, kd = cur.kd;
      for (int q = gl; q < D2; q += G) {
        const int i = q / PD, j = q % PD;
        double s = dpart[q];
        for (int ch = 1; ch < nchunk; ++ch) s += dpart[ch * D2 + q];
        double *row = vals + (D2 * b0 + (int64_t)i * PD * nb);
        if constexpr (PEER) {
          if (!peer_local(pt, D2 * b0, D2 * b0 + (int64_t)D2 * nb)) row = peer_row(pt, vals, D2 * b0 + (int64_t)i * PD * nb);
        }
        row[PD * kd + j] = s;
      }
    }
    if (fint) {   // butterfly over the group (fixed order), lanes 0..PD-1 write
#pragma unroll
      for (int d = G / 2; d >= 1; d >>= 1)
#pragma unroll
        for (int i = 0; i < PD; ++i) fpart[i] += __shfl_xor_sync(0xffffffffu, fpart[i], d);
      if (node_on && gl < PD) {
        const double mine_f = gl == 0 ? fpart[0] : (gl == 1 ? fpart[1] : fpart[PD - 1]);
        double *fp = fint + ((int64_t)PD * (dm.nl2g[r] - 1) + gl);
        if constexpr (PEER) {
          if (!peer_local_f(pt, (int64_t)PD * r + gl)) fp = peer_f(pt, fp, (int64_t)PD * r + gl);
        }
        *fp = mine_f;
      }
    }
    __syncwarp();
  }
}

template <int PD, int G>
int launch_g(femx_ctx *c, const double *d_u, double *d_vals, double *d_fint) {
  const PeerTab pt = c->peer.in_kernel ? peer_tab(c) : PeerTab{};
  int nsm = 148, per_sm = 1;
  cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, c->device);
  auto go = [&](auto kern) {
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kWarps * 32, 0);
    if (per_sm < 1) per_sm = 1;
    const int64_t ngroups = ((int64_t)c->nlnds + 32 / G - 1) / (32 / G);
    int64_t grid = (ngroups + kWarps - 1) / kWarps;
    const int64_t cap = (int64_t)nsm * per_sm * 4;
    if (grid > cap) grid = cap;
    if (grid < 1) grid = 1;
    kern<<<(unsigned)grid, kWarps * 32, 0, c->stream>>>(c->dm, d_u, d_vals, d_fint, pt);
  };
  if (pt.n > 0) go(rows_simplex_kernel<PD, G, true>); else go(rows_simplex_kernel<PD, G, false>);
  return 1;
}

template <int PD>
int launch_pd(femx_ctx *c, const double *d_u, double *d_vals, double *d_fint, int g) {
  switch (g) {
    case 8: return launch_g<PD, 8>(c, d_u, d_vals, d_fint);
    case 16: return launch_g<PD, 16>(c, d_u, d_vals, d_fint);
    default: return launch_g<PD, 32>(c, d_u, d_vals, d_fint);
  }
}

}  // namespace

bool rows_simplex_ok(const femx_ctx *c) { return c->plan.simplex && c->plan.max_inc >= 1 && c->plan.max_inc <= 32; }

int launch_rows_simplex(femx_ctx *c, const double *d_u, double *d_vals, double *d_fint) {
  int g = 8;   // groups of at least 8 lanes: room for the diagonal chunks
  while (g < c->plan.max_inc) g *= 2;
  return c->pdim == 2 ? launch_pd<2>(c, d_u, d_vals, d_fint, g) : launch_pd<3>(c, d_u, d_vals, d_fint, g);
}

}  // namespace femx
