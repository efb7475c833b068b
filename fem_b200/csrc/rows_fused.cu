// rows_fused.cu -- assembly variant 9: fused gather-ordered assembly for ANY element mix (tri3, quad4, tet4, hex8,
// cohesive), the general-mesh counterpart of the hex8 tile kernels.
//
// The two-pass path (kernels.cu) writes every element matrix to HBM and reads it back (tet4: 1.15 kB each way for
// 215 B of compulsory traffic).  Here the element row blocks never leave the SM:
//   phase A  a group of G lanes (G = power of two >= the largest number of elements around a node) owns one node
//            row; lane i computes row-block a_i of Ke of the i-th incident element (ke_row_blocks, FormElKE
//            src/m_local.F90:21-67) and its share f = K_a. u_e of the internal force into its shared-memory slot;
//   phase B  the same lanes walk the node's CSR values in storage order (lane <-> consecutive value, coalesced
//            stores) and sum, per value, the contributing (element, local node b) entries listed by the plan in
//            ascending element order -- MatSetValues(ADD_VALUES) semantics (src/main.F90:157-171) without atomics,
//            each value written exactly once.
// Plan data read per node: the incidence list and one byte per (element, b) contribution.
#include "elem_math.cuh"
#include "launch.h"

namespace femx {

namespace {

constexpr int kWarps = 8;

template <int T>
__device__ __forceinline__ void row_of(const DevMesh &dm, const int32_t *cn, int m, int a, const double *__restrict__ u, double *out, int foff) {
  constexpr int NEN = ET<T>::NEN, DIM = ET<T>::DIM, D2 = DIM * DIM;
  if constexpr (T == FEMX_COH) {
    // E = nu = 0 -> D = 0 (src/m_global.F90:250-252): explicit zero blocks that still enter the pattern
#pragma unroll
    for (int q = 0; q < NEN * D2; ++q) out[q] = 0.0;
#pragma unroll
    for (int i = 0; i < DIM; ++i) out[foff + i] = 0.0;
  } else {
    double X[NEN][DIM];
    int nd[NEN];
#pragma unroll
    for (int b = 0; b < NEN; ++b) {
      nd[b] = cn[b];
#pragma unroll
      for (int d = 0; d < DIM; ++d) X[b][d] = dm.coords[(int64_t)nd[b] * DIM + d];
    }
    const double lam = dm.lam_mu[2 * m], mu = dm.lam_mu[2 * m + 1];
    double acc[NEN][D2];
    ke_row_blocks<T>(X, a, lam, mu, acc);
#pragma unroll
    for (int b = 0; b < NEN; ++b)
#pragma unroll
      for (int q = 0; q < D2; ++q) out[b * D2 + q] = acc[b][q];
    if (u) {
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
      for (int i = 0; i < DIM; ++i) out[foff + i] = f[i];
    }
  }
}

// HEX: the mesh holds hex8 elements (keeps their 72 accumulators out of tet-only builds); PEER: rows another rank owns
// are stored into that rank's receive arena (peer.cu)
template <int PD, int G, bool HEX, bool PEER>
__global__ void __launch_bounds__(kWarps * 32)
fused_rows_kernel(DevMesh dm, const double *__restrict__ u, double *__restrict__ vals, double *__restrict__ fint, int slot, int foff,
                  const PeerTab pt) {
  extern __shared__ double sm_rows[];
  constexpr int D2 = PD * PD, NPW = 32 / G;
  constexpr int PARTS = G / D2;                  // lanes per diagonal-block value when the group is wide enough to split it
  constexpr bool BLK = G <= 8;   // few elements around a node (quad4 / tri3 / hex8): phase B maps lane <-> block
  constexpr bool SPLIT = !BLK && PARTS >= 2;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int gl = lane % G, gi = lane / G;
  double *mine = sm_rows + ((size_t)w * 32 + lane) * slot;
  const double *grp = sm_rows + ((size_t)w * 32 + gi * G) * slot;
  const int64_t ngroups = ((int64_t)dm.nlnds + NPW - 1) / NPW;
  const int64_t stride = (int64_t)gridDim.x * kWarps;
  // incidence of the first node group; the next one is fetched while the current one is being processed
  // 2-D meshes: the plan holds a flattened record per incidence (the element's nodes, type, local node, material row:
  // ri_nodes / ri_info), so the chain element -> type / connectivity -> nodes is off the device's critical path
  int64_t wg = (int64_t)blockIdx.x * kWarps + w;
  int ninc = 0, e = 0, a = 0;
  int64_t b0n = 0, rben = 0;   // next group's row metadata: first block, block count, first list entry
  int nbn = 0;
  int4 rn = make_int4(0, 0, 0, 0);
  auto fetch_inc = [&](int64_t p) {
    if constexpr (PD == 2) {
      rn = *reinterpret_cast<const int4 *>(dm.ri_nodes + 4 * p);
      e = dm.ri_info[p];   // type | a << 4 | material row << 8
    } else {
      e = dm.n2e_elem[p]; a = dm.n2e_a[p];
    }
  };
  {
    const int64_t r = wg * NPW + gi;
    if (wg < ngroups && r < dm.nlnds) {
      const int64_t p0 = dm.n2e_ptr[r];
      ninc = (int)(dm.n2e_ptr[r + 1] - p0);
      if (gl < ninc) fetch_inc(p0 + gl);
      b0n = dm.bptr[r]; nbn = (int)(dm.bptr[r + 1] - b0n); rben = dm.rbe_ptr[r];
    }
  }
  for (; wg < ngroups; wg += stride) {
    const int64_t r = wg * NPW + gi;
    const bool node_on = r < dm.nlnds;
    const int ninc_cur = ninc, e_cur = e, a_cur = a;
    const int4 rn_cur = rn;
    const int64_t b0 = b0n, rbe = rben;
    const int nb = nbn;
    // the lane's first block in phase B: its list bounds travel while phase A computes
    const uint16_t *off = dm.rb_off + b0 + r;
    int off0 = 0, off1 = 0;
    if (BLK && node_on && gl < nb) { off0 = off[gl]; off1 = off[gl + 1]; }
    {   // prefetch
      const int64_t rnext = (wg + stride) * NPW + gi;
      ninc = 0;
      if (wg + stride < ngroups && rnext < dm.nlnds) {
        const int64_t p0 = dm.n2e_ptr[rnext];
        ninc = (int)(dm.n2e_ptr[rnext + 1] - p0);
        if (gl < ninc) fetch_inc(p0 + gl);
        b0n = dm.bptr[rnext]; nbn = (int)(dm.bptr[rnext + 1] - b0n); rben = dm.rbe_ptr[rnext];
      }
    }
    // ---- phase A: one incident element per lane
    if (gl < ninc_cur) {
      if constexpr (PD == 2) {
        const int32_t cn[4] = {rn_cur.x, rn_cur.y, rn_cur.z, rn_cur.w};
        const int t = e_cur & 15, aa = (e_cur >> 4) & 15, m = e_cur >> 8;
        if (t == FEMX_QUA) row_of<FEMX_QUA>(dm, cn, m, aa, u, mine, foff);
        else if (t == FEMX_TRI) row_of<FEMX_TRI>(dm, cn, m, aa, u, mine, foff);
        else row_of<FEMX_COH>(dm, cn, m, aa, u, mine, foff);
      } else {
        const int t = dm.etype[e_cur];
        const int32_t *cn = dm.conn0 + dm.eptr[e_cur];
        const int m = dm.emat[e_cur] - 1;
        if (!HEX || t == FEMX_TET) row_of<FEMX_TET>(dm, cn, m, a_cur, u, mine, foff);
        else if constexpr (HEX) row_of<FEMX_HEX>(dm, cn, m, a_cur, u, mine, foff);
      }
    }
    __syncwarp();
    // ---- phase B: the node's CSR values in storage order
    const unsigned on_mask = __ballot_sync(0xffffffffu, node_on);
    if (node_on) {
      const int rowlen = PD * nb, nv = PD * rowlen;
      const uint16_t *ent = dm.rb_ent + rbe;
      double *vrow = vals + D2 * b0;
      double *prow[PD];   // PEER: first value of each of the node's PD rows, at its owner
      if constexpr (PEER) {
#pragma unroll
        for (int i = 0; i < PD; ++i) prow[i] = vrow + (int64_t)i * rowlen;
        if (!peer_local(pt, D2 * b0, D2 * b0 + nv)) {
#pragma unroll
          for (int i = 0; i < PD; ++i) prow[i] = peer_row(pt, vals, D2 * b0 + (int64_t)i * rowlen);
        }
      }
      if constexpr (BLK) {   // one list walk per block instead of one per value
        const bool al16 = !PEER && (reinterpret_cast<size_t>(vals) & 15) == 0;
        for (int k = gl; k < nb; k += G) {
          int c = k == gl ? off0 : off[k];
          const int c1 = k == gl ? off1 : off[k + 1];
          double acc[D2];
#pragma unroll
          for (int q = 0; q < D2; ++q) acc[q] = 0.0;
          int en = c < c1 ? ent[c] : 0;
          for (; c < c1; ++c) {
            const int en_next = c + 1 < c1 ? ent[c + 1] : 0;   // the next entry travels while this one is used
            const double *src = grp + en;
#pragma unroll
            for (int q = 0; q < D2; ++q) acc[q] += src[q];
            en = en_next;
          }
#pragma unroll
          for (int i = 0; i < PD; ++i) {
            double *row = vrow + (int64_t)i * rowlen;
            if constexpr (PEER) row = prow[i];
            if (PD == 2 && al16) {   // 2-D: 4 b0 + i * 2 nb + 2 k is even, every block row is 16-byte aligned
              *reinterpret_cast<double2 *>(row + 2 * k) = make_double2(acc[i * 2], acc[i * 2 + 1]);
            } else {
#pragma unroll
              for (int j = 0; j < PD; ++j) row[PD * k + j] = acc[i * PD + j];
            }
          }
        }
      }
      const int kd = SPLIT ? dm.bdiag[r] : -1;   // the diagonal block has one contribution per incident element: split it
      for (int v = gl; !BLK && v < nv; v += G) {
        const int i = (v >= rowlen) + (PD == 3 && v >= 2 * rowlen), rem = v - i * rowlen, k = rem / PD, j = rem - k * PD;
        if (k == kd) continue;
        const double *src = grp + i * PD + j;
        int c = off[k];
        const int c1 = off[k + 1];
        double s = 0.0;
        for (; c + 1 < c1; c += 2) {
          const double x0 = src[ent[c]], x1 = src[ent[c + 1]];
          s += x0;
          s += x1;
        }
        if (c < c1) s += src[ent[c]];
        if constexpr (PEER) (i == 0 ? prow[0] : (i == 1 ? prow[1] : prow[PD - 1]))[rem] = s;
        else vrow[v] = s;
      }
      if constexpr (SPLIT) {
        // lanes (q, part): part sums every PARTS-th contribution in ascending order, then (p0 + p1) + p2 ...
        const int q = gl % D2, part = gl / D2;
        double s = 0.0;
        if (part < PARTS) {
          const double *src = grp + q;
          const int c1 = off[kd + 1];
          for (int c = off[kd] + part; c < c1; c += PARTS) s += src[ent[c]];
        }
        double tot = __shfl_sync(on_mask, s, gi * G + q);
#pragma unroll
        for (int pp = 1; pp < PARTS; ++pp) tot += __shfl_sync(on_mask, s, gi * G + pp * D2 + q);
        if (gl < D2) {
          if constexpr (PEER) (q / PD == 0 ? prow[0] : (q / PD == 1 ? prow[1] : prow[PD - 1]))[PD * kd + q % PD] = tot;
          else vrow[(q / PD) * rowlen + PD * kd + q % PD] = tot;
        }
      }
      if (fint) {   // f_r = sum of the incident elements' shares, FP interleaved partial sums per component (fixed order)
        constexpr int FP = G >= 16 ? 4 : 1;
        const int d = gl % PD, part = gl / PD;
        double s = 0.0;
        if (part < FP)
          for (int q = part; q < ninc_cur; q += FP) s += grp[q * slot + foff + d];
        if constexpr (FP > 1) {
          double tot = __shfl_sync(on_mask, s, gi * G + d);
#pragma unroll
          for (int pp = 1; pp < FP; ++pp) tot += __shfl_sync(on_mask, s, gi * G + pp * PD + d);
          s = tot;
        }
        if (gl < PD) {
          double *fp = fint + ((int64_t)PD * (dm.nl2g[r] - 1) + gl);
          if constexpr (PEER) {
            if (!peer_local_f(pt, (int64_t)PD * r + gl)) fp = peer_f(pt, fp, (int64_t)PD * r + gl);
          }
          *fp = s;
        }
      }
    }
    __syncwarp();
  }
}

template <int PD, int G, bool HEX, bool PEER>
int launch_gp(femx_ctx *c, const double *d_u, double *d_vals, double *d_fint, int slot, int foff, const PeerTab &pt) {
  const size_t smem = (size_t)kWarps * 32 * slot * sizeof(double);
  cudaFuncSetAttribute(fused_rows_kernel<PD, G, HEX, PEER>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  int nsm = 148, per_sm = 1;
  cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, c->device);
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fused_rows_kernel<PD, G, HEX, PEER>, kWarps * 32, smem);
  if (per_sm < 1) per_sm = 1;
  const int64_t ngroups = ((int64_t)c->nlnds + 32 / G - 1) / (32 / G);
  int64_t grid = (ngroups + kWarps - 1) / kWarps;
  const int64_t cap = (int64_t)nsm * per_sm * 4;   // a few CTAs per resident slot, grid-stride over the rest
  if (grid > cap) grid = cap;
  if (grid < 1) grid = 1;
  fused_rows_kernel<PD, G, HEX, PEER><<<(unsigned)grid, kWarps * 32, smem, c->stream>>>(c->dm, d_u, d_vals, d_fint, slot, foff, pt);
  return 1;
}

template <int PD, int G, bool HEX>
int launch_g(femx_ctx *c, const double *d_u, double *d_vals, double *d_fint, int slot, int foff) {
  const PeerTab pt = c->peer.in_kernel ? peer_tab(c) : PeerTab{};
  return pt.n > 0 ? launch_gp<PD, G, HEX, true>(c, d_u, d_vals, d_fint, slot, foff, pt)
                  : launch_gp<PD, G, HEX, false>(c, d_u, d_vals, d_fint, slot, foff, pt);
}

template <int PD, bool HEX>
int launch_pd(femx_ctx *c, const double *d_u, double *d_vals, double *d_fint, int g, int slot, int foff) {
  switch (g) {
    case 4: return launch_g<PD, 4, HEX>(c, d_u, d_vals, d_fint, slot, foff);
    case 8: return launch_g<PD, 8, HEX>(c, d_u, d_vals, d_fint, slot, foff);
    case 16: return launch_g<PD, 16, HEX>(c, d_u, d_vals, d_fint, slot, foff);
    default: return launch_g<PD, 32, HEX>(c, d_u, d_vals, d_fint, slot, foff);
  }
}

}  // namespace

bool rows_fused_ok(const femx_ctx *c) { return c->plan.max_inc >= 1 && c->plan.max_inc <= 32; }

int launch_rows_fused(femx_ctx *c, const double *d_u, double *d_vals, double *d_fint) {
  int g = 4;
  while (g < c->plan.max_inc) g *= 2;
  int maxnen = 3;
  for (const auto &b : c->buckets) maxnen = maxnen > type_nen(b.type) ? maxnen : type_nen(b.type);
  const int pd = c->pdim, slot = c->plan.rb_slot, foff = c->plan.rb_foff;   // the plan's entries are offsets into these slots
  if (pd == 2) return launch_pd<2, false>(c, d_u, d_vals, d_fint, g, slot, foff);
  return maxnen == 8 ? launch_pd<3, true>(c, d_u, d_vals, d_fint, g, slot, foff) : launch_pd<3, false>(c, d_u, d_vals, d_fint, g, slot, foff);
}

}  // namespace femx
