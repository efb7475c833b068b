// rows_simplex.cu -- assembly variant 11: fused gather-ordered assembly for pure tri3 / tet4 meshes (one integration
// point per element), the lean form of rows_fused.cu.
//
// A simplex element's stiffness is an outer product of its constant shape-function gradients:
//   K_ab(i,j) = w detJ (lambda g_a,i g_b,j + mu g_a,j g_b,i + delta_ij mu g_a . g_b)      (FormElKE, src/m_local.F90:61-65)
// so instead of parking the 4 x 9 values of a row block per (node, incident element) in shared memory (39 doubles) and
// reading every one of them back once per CSR value, the incident lanes park the GRADIENTS (12 doubles + 2
// coefficients) and the CSR blocks are formed where they are summed:
//   phase 0  lane k: coordinates of column node k of the row -> shared memory (every node of every incident element is one of
//            the row's column nodes; the plan names them by position, sx_loc), loaded one group ahead;
//   phase A  the lanes of the node's group: gradients of the incident elements, scaled by sqrt(w detJ mu), and
//            rho = lambda / mu -> one shared-memory slot per element (K_ab = rho s_a s_b^T + s_b s_a^T + (s_a . s_b) I;
//            the plan admits a mesh to this kernel only if every element has detJ > 0 and mu > 0);
//   phase B  lane k: block k of the node's row = sum over the contributing (element, a, b) of the plan's list, in
//            ascending element order (MatSetValues ADD semantics, src/main.F90:157-171, without atomics); the diagonal
//            block (one contribution per incident element) is cut into chunks taken by otherwise idle lanes and
//            recombined in chunk order; the lane multiplies its block with u of the column node and the group
//            butterfly-sums the internal force.
// Per contribution: 8 shared loads + ~30 FP64 operations for 9 matrix entries (rows_fused.cu: 9 loads + 9 index
// loads + 9 adds), no per-value index arithmetic.
#include <cstdio>

#include "elem_math.cuh"
#include "launch.h"

namespace femx {

namespace {

#ifndef FEMX_SX_MINB
#define FEMX_SX_MINB 4       // resident CTAs per SM the register allocation aims at
#endif
#ifndef FEMX_SX_BAL
#define FEMX_SX_BAL 1        // balanced phase B (rows of at most G blocks): own block, then a share of the diagonal block
#endif
#ifndef FEMX_SX_SLOT3
#define FEMX_SX_SLOT3 18     // tet4: doubles per incident-element slot (16 used; pitch = odd multiple of 16 bytes)
#endif
constexpr int kChunks = 4;   // unbalanced form: the diagonal block is cut into up to this many tasks
constexpr int kChunkLen = 6; // ... of about this many contributions (the typical length of an off-diagonal list)

template <int PD, int G, int WARPS, bool PEER>
__global__ void __launch_bounds__(WARPS * 32, FEMX_SX_MINB)
rows_simplex_kernel(DevMesh dm, const double *__restrict__ u, double *__restrict__ vals, double *__restrict__ fint, int mslots,
                    int xslots, const PeerTab pt) {
  constexpr int T = PD == 2 ? FEMX_TRI : FEMX_TET;
  constexpr int NEN = PD + 1, D2 = PD * PD, NPW = 32 / G;
  // slot of one incident element: per node a 16-byte aligned scaled gradient (tet: s0 s1 s2 rho, tri: s0 s1 and rho
  // in a quad of its own); pitch = odd multiple of 16 bytes so that the lanes' 128-bit stores spread over the banks
  constexpr int NS = PD == 3 ? 4 : 2;               // doubles per node
  constexpr int COEF = NEN * NS;                    // tri: offset of rho
  constexpr int SLOT = PD == 3 ? FEMX_SX_SLOT3 : 10;
  constexpr int DPART = (FEMX_SX_BAL && G > kChunks ? G : kChunks) * D2;   // diagonal partials of one group
  extern __shared__ __align__(16) double sm_dyn[];
  // G lanes own one node row; a warp works on NPW = 32 / G rows at once (independent dependent-load chains in flight);
  // a group holds mslots = max elements around a node slots, its lanes take the incident elements / the row's tasks
  // round-robin
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int gl = lane % G, gi = lane / G;
  double *grp = sm_dyn + ((size_t)(w * NPW + gi) * mslots) * SLOT;
  double *dpart = sm_dyn + (size_t)WARPS * NPW * mslots * SLOT + ((size_t)w * NPW + gi) * DPART;
  double *xg = sm_dyn + (size_t)WARPS * NPW * (mslots * SLOT + DPART) + ((size_t)w * NPW + gi) * xslots * PD;
  const int64_t ngroups = ((int64_t)dm.nlnds + NPW - 1) / NPW;
  const int64_t stride = (int64_t)gridDim.x * WARPS;
  // Every element around node r consists of column nodes of r's block row, and the plan names them by their position there
  // (sx_loc, one byte per node).  So the group fetches the coordinates of the row's column nodes ONCE (lane k: column k) into
  // shared memory and the incident elements pick their nodes from there, instead of 4 x PD scattered loads per incidence.
  // The dependent chain row -> (column, list bounds, incidence records) -> coordinates is software-pipelined over the
  // groups a warp visits: row metadata two groups ahead, the lane's record one group ahead (after phase A), the column
  // node's coordinates after phase B.
  struct Meta { int64_t p0, b0; int ninc, nb, kd; };
  struct Rec { int col, off0, off1, d0, d1, mat0, mat1; uint32_t loc0, loc1; };
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
  auto load_rec = [&](const Meta &m, int64_t g) {   // first task of the lane in phase B and its incidences of the first two passes
    Rec R{0, 0, 0, 0, 0, 0, 0, 0u, 0u};
    const int64_t r = g * NPW + gi;
    if (g < ngroups && r < dm.nlnds) {
      const uint16_t *off = dm.rb_off + m.b0 + r;
      if (gl < m.nb) { R.off0 = off[gl]; R.off1 = off[gl + 1]; R.col = dm.bcol[m.b0 + gl]; }
      R.d0 = off[m.kd]; R.d1 = off[m.kd + 1];
      if (gl < m.ninc) { R.loc0 = dm.sx_loc[m.p0 + gl]; R.mat0 = dm.sx_mat[m.p0 + gl]; }
      if (gl + G < m.ninc) { R.loc1 = dm.sx_loc[m.p0 + gl + G]; R.mat1 = dm.sx_mat[m.p0 + gl + G]; }
    }
    return R;
  };
  auto load_x = [&](const Meta &m, const Rec &R, double (&x)[PD]) {
#pragma unroll
    for (int d = 0; d < PD; ++d) x[d] = gl < m.nb ? dm.coords[(int64_t)R.col * PD + d] : 0.0;
  };
  int64_t wg = (int64_t)blockIdx.x * WARPS + w;
  Meta cur = fetch(wg);
  Rec R = load_rec(cur, wg);
  double xk[PD];
  load_x(cur, R, xk);
  Meta nxt = fetch(wg + stride);
  for (; wg < ngroups; wg += stride) {
    const int64_t r = wg * NPW + gi;
    const bool node_on = r < dm.nlnds;
    const int ninc = cur.ninc, nb = cur.nb, kd = cur.kd;
    const int64_t b0 = cur.b0;
    const uint16_t *off = dm.rb_off + b0 + r;
    const int t_off0 = R.off0, t_off1 = R.off1, d_off0 = R.d0, d_off1 = R.d1, t_col = R.col;
    const Meta nx2 = fetch(wg + 2 * stride);
    if (gl < nb) {
#pragma unroll
      for (int d = 0; d < PD; ++d) xg[gl * PD + d] = xk[d];
    }
    for (int k = gl + G; k < nb; k += G) {   // rows wider than the group
      const int col = dm.bcol[b0 + k];
#pragma unroll
      for (int d = 0; d < PD; ++d) xg[k * PD + d] = dm.coords[(int64_t)col * PD + d];
    }
    __syncwarp();
    // ---- phase A: gradients of the incident elements, one per lane and pass
    for (int q = gl, pass = 0; q < ninc; q += G, ++pass) {
      const uint32_t loc = pass == 0 ? R.loc0 : (pass == 1 ? R.loc1 : dm.sx_loc[cur.p0 + q]);
      const int mat = pass == 0 ? R.mat0 : (pass == 1 ? R.mat1 : dm.sx_mat[cur.p0 + q]);
      double X[NEN][PD];
#pragma unroll
      for (int b = 0; b < NEN; ++b) {
        const double *xs = xg + ((loc >> (8 * b)) & 255u) * PD;
#pragma unroll
        for (int d = 0; d < PD; ++d) X[b][d] = xs[d];
      }
      const double lam = dm.lam_mu[2 * mat], mu = dm.lam_mu[2 * mat + 1];
      double g[NEN][PD];
      const double wd = phys_grad<T>(X, 0, g);
      double *mine = grp + q * SLOT;
      const double sc = sqrt(wd * mu), rho = lam / mu;
#pragma unroll
      for (int b = 0; b < NEN; ++b) {
        double2 *qd = reinterpret_cast<double2 *>(mine + b * NS);
        qd[0] = make_double2(g[b][0] * sc, g[b][1] * sc);
        if constexpr (PD == 3) qd[1] = make_double2(g[b][2] * sc, rho);
      }
      if constexpr (PD == 2) *reinterpret_cast<double2 *>(mine + COEF) = make_double2(rho, 0.0);
    }
    const Rec Rn = load_rec(nxt, wg + stride);
    __syncwarp();
    // ---- phase B.  One list of the plan = the contributions (element slot, a, b) to one block, ascending elements:
    auto run_list = [&](const uint16_t *ent, int c0, int c1, double (&acc)[D2]) {
      double tr = 0.0;
#pragma unroll
      for (int q = 0; q < D2; ++q) acc[q] = 0.0;
      int en = c0 < c1 ? ent[c0] : 0;
      for (int c = c0; c < c1; ++c) {
        const int en_next = c + 1 < c1 ? ent[c + 1] : 0;   // the next entry travels while this one is used
        const double *sl = grp + (en >> 4) * SLOT;
        const double2 *qa = reinterpret_cast<const double2 *>(sl + ((en >> 2) & 3) * NS);
        const double2 *qb = reinterpret_cast<const double2 *>(sl + (en & 3) * NS);
        double ga[PD], vb[PD], la[PD], rho;
        { const double2 t2 = qa[0], v2 = qb[0]; ga[0] = t2.x; ga[1] = t2.y; vb[0] = v2.x; vb[1] = v2.y; }
        if constexpr (PD == 3) { const double2 t3 = qa[1]; ga[2] = t3.x; rho = t3.y; vb[2] = qb[1].x; }
        else rho = sl[COEF];
#pragma unroll
        for (int i = 0; i < PD; ++i) la[i] = rho * ga[i];
#pragma unroll
        for (int i = 0; i < PD; ++i) tr = fma(ga[i], vb[i], tr);
#pragma unroll
        for (int i = 0; i < PD; ++i)
#pragma unroll
          for (int j = 0; j < PD; ++j) acc[i * PD + j] = fma(ga[j], vb[i], fma(la[i], vb[j], acc[i * PD + j]));
        en = en_next;
      }
#pragma unroll
      for (int i = 0; i < PD; ++i) acc[i * PD + i] += tr;   // delta_ij s_a . s_b, summed over the contributions
    };
    double fpart[PD];
#pragma unroll
    for (int i = 0; i < PD; ++i) fpart[i] = 0.0;
    auto add_f = [&](const double (&acc)[D2], int col) {   // this block's (or chunk's) share of f_r = sum_k K_rk u_k
      double uk[PD];
#pragma unroll
      for (int j = 0; j < PD; ++j) uk[j] = u[(int64_t)col * PD + j];
#pragma unroll
      for (int j = 0; j < PD; ++j)
#pragma unroll
        for (int i = 0; i < PD; ++i) fpart[i] = fma(acc[i * PD + j], uk[j], fpart[i]);
    };
    const int dlo = d_off0, dn = d_off1 - d_off0;
    const int rowlen = PD * nb;
    const uint16_t *ent = dm.sx_ent + NEN * cur.p0;   // every element lists NEN contributions: rbe_ptr = NEN * n2e_ptr
    bool remote = false;
    if constexpr (PEER) remote = node_on && !peer_local(pt, D2 * b0, D2 * b0 + (int64_t)PD * rowlen);
    auto store_block = [&](const double (&acc)[D2], int k) {
#pragma unroll
      for (int i = 0; i < PD; ++i) {
        double *row = vals + (D2 * b0 + (int64_t)i * rowlen);
        if constexpr (PEER) { if (remote) row = peer_row(pt, vals, D2 * b0 + (int64_t)i * rowlen); }
#pragma unroll
        for (int j = 0; j < PD; ++j) row[PD * k + j] = acc[i * PD + j];
      }
    };
    // Balanced form (rows of at most G blocks): lane k sums block k, then a share of the diagonal block -- the
    // longest list of the row, one contribution per incident element -- sized so that every lane handles about
    // total / G contributions; the shares are laid end to end over the lanes (prefix sum of the lanes' spare steps).
    const bool bal = FEMX_SX_BAL && nb <= G;
    int nparts = 0;   // diagonal partials to recombine (0: stored directly)
    int spare = 0, pre = 0;
    if (bal) {
      const int own = (gl < nb && gl != kd) ? t_off1 - t_off0 : 0;
      spare = max(0, (NEN * ninc + G - 1) / G - own);
    }
    if constexpr (FEMX_SX_BAL) {
      int inc = spare;
#pragma unroll
      for (int d = 1; d < G; d <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, inc, d, G);
        if (gl >= d) inc += v;
      }
      pre = inc - spare;
    }
    if (node_on && bal) {
      double acc[D2];
      if (gl < nb && gl != kd) {
        run_list(ent, t_off0, t_off1, acc);
        if (fint) add_f(acc, t_col);
        store_block(acc, gl);
      }
      const int c0 = dlo + min(pre, dn), c1 = dlo + min(pre + spare, dn);
      run_list(ent, c0, c1, acc);
      if (fint && c0 < c1) add_f(acc, (int)r);
#pragma unroll
      for (int q = 0; q < D2; ++q) dpart[gl * D2 + q] = acc[q];
      nparts = nb > 0 ? G : 0;
    } else if (node_on) {
      // tasks = the row's blocks, the diagonal block cut into chunks; lane gl takes tasks gl, gl + G, ...
      const int nchunk = max(1, min(kChunks, (dn + kChunkLen - 1) / kChunkLen));
      const int clen = (dn + nchunk - 1) / nchunk;
      for (int t = gl; t < nb + nchunk - 1; t += G) {
        // task t: block t, or chunk (t - nb + 1) of the diagonal block
        const int k = t < nb ? t : kd;
        int c0, c1, chunk = -1, col;
        if (t == gl && t < nb) { c0 = t_off0; c1 = t_off1; col = t_col; } else { c0 = off[k]; c1 = off[k + 1]; col = dm.bcol[b0 + k]; }
        if (k == kd) {
          chunk = t < nb ? 0 : t - nb + 1;
          c0 = dlo + chunk * clen;
          c1 = min(dlo + dn, c0 + clen);
        }
        double acc[D2];
        run_list(ent, c0, c1, acc);
        if (fint) add_f(acc, col);
        if (chunk < 0 || nchunk == 1) store_block(acc, k);
        else {
#pragma unroll
          for (int q = 0; q < D2; ++q) dpart[chunk * D2 + q] = acc[q];
        }
      }
      nparts = nchunk > 1 ? nchunk : 0;
    }
    __syncwarp();
    if (nparts > 0) {   // diagonal block = its partial sums in order, one entry per lane
      for (int q = gl; q < D2; q += G) {
        const int i = q / PD, j = q % PD;
        double s = dpart[q];
        for (int ch = 1; ch < nparts; ++ch) s += dpart[ch * D2 + q];
        double *row = vals + (D2 * b0 + (int64_t)i * rowlen);
        if constexpr (PEER) { if (remote) row = peer_row(pt, vals, D2 * b0 + (int64_t)i * rowlen); }
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
    load_x(nxt, Rn, xk);
    cur = nxt; nxt = nx2; R = Rn;
    __syncwarp();
  }
}

template <int PD, int G, int WARPS>
int launch_g(femx_ctx *c, const double *d_u, double *d_vals, double *d_fint) {
  constexpr int NPW = 32 / G, SLOT = PD == 3 ? FEMX_SX_SLOT3 : 10, DPART = (FEMX_SX_BAL && G > kChunks ? G : kChunks) * PD * PD;
  const PeerTab pt = c->peer.in_kernel ? peer_tab(c) : PeerTab{};
  const int mslots = c->plan.max_inc, xslots = c->plan.max_nb;
  const size_t smem = sizeof(double) * ((size_t)WARPS * NPW * (mslots * SLOT + DPART) + (size_t)WARPS * NPW * xslots * PD);
  int nsm = 148, per_sm = 1;
  cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, c->device);
  auto go = [&](auto kern) {
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, WARPS * 32, smem);
    if (per_sm < 1) per_sm = 1;
    const int64_t ngroups = ((int64_t)c->nlnds + NPW - 1) / NPW;
    int64_t grid = (ngroups + WARPS - 1) / WARPS;
    const int64_t cap = (int64_t)nsm * per_sm * 4;
    if (grid > cap) grid = cap;
    if (grid < 1) grid = 1;
    if (c->opts.reserved[2] & 32) fprintf(stderr, "femx: simplex rows kernel G %d, %d warps, smem %zu B, %d CTAs per SM, grid %lld\n", G, WARPS, smem, per_sm, (long long)grid);
    kern<<<(unsigned)grid, WARPS * 32, smem, c->stream>>>(c->dm, d_u, d_vals, d_fint, mslots, xslots, pt);
  };
  if (pt.n > 0) go(rows_simplex_kernel<PD, G, WARPS, true>); else go(rows_simplex_kernel<PD, G, WARPS, false>);
  return 1;
}

template <int PD>
int launch_pd(femx_ctx *c, const double *d_u, double *d_vals, double *d_fint, int g) {
  switch (g) {
    case 4: return launch_g<PD, 4, 4>(c, d_u, d_vals, d_fint);
    case 16: return launch_g<PD, 16, 4>(c, d_u, d_vals, d_fint);
    case 32: return launch_g<PD, 32, 8>(c, d_u, d_vals, d_fint);
    default: return launch_g<PD, 8, 4>(c, d_u, d_vals, d_fint);
  }
}

}  // namespace

bool rows_simplex_ok(const femx_ctx *c) { return c->plan.simplex && c->plan.max_inc >= 1 && c->plan.max_inc <= 32 && c->dm.sx_loc; }

int launch_rows_simplex(femx_ctx *c, const double *d_u, double *d_vals, double *d_fint) {
  // lanes per node row: 16 for tet4, 4 for tri3 (two / eight rows in flight per warp; measured on 5 M tet4 / 2 M tri3: tet4 1.22 ms at 16,
  // 1.25 at 8; tri3 0.38 ms at 4, 0.51 at 8, 0.78 at 16) unless overridden (opts.reserved[0] = 4, 8, 16, 32: tuning knob)
  const int g = c->opts.reserved[0] > 0 ? c->opts.reserved[0] : (c->pdim == 3 ? 16 : 4);
  return c->pdim == 2 ? launch_pd<2>(c, d_u, d_vals, d_fint, g) : launch_pd<3>(c, d_u, d_vals, d_fint, g);
}

}  // namespace femx
