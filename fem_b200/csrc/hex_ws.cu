// hex_ws.cu -- assembly variant 3: warp-specialised fused hex8 K (+ f_int) kernel.
//
// Same arithmetic, tile program and summation order as hex_fused.cu (variant 2); what changes is WHO
// does what, so that the latency-bound parts overlap the FP64 loop instead of alternating with it:
//   warps 0-3  "consumers" (128 threads, 240 registers each after setmaxnreg.inc): the 16-contribution
//              FMA loop into 9 register blocks, G' -> K, f partials, transpose dump to shared memory;
//   warps 4-15 "helpers"   (384 threads, 88 registers): cp.async prefetch ring of node layers,
//              geometry of the NEXT element layer into the other gradient buffer, and the global
//              stores of the PREVIOUS step's dump.
// Hand-over through named barriers (bar.arrive / bar.sync): full[2] / empty[2] for the double-buffered
// gradients, dump_full / dump_empty for the transpose area.  One CTA per SM (222 KB shared memory),
// tile 8x4 node positions, z-chunks as in variant 2.
#include "elem_math.cuh"
#include "launch.h"

namespace femx {

namespace {

constexpr int TX = 8, TY = 4, NP = TX * TY;
constexpr int NCONS = 4 * NP;          // 128 consumer threads
constexpr int NHELP = 384;             // helper threads (12 warps: one geometry item each per layer)
constexpr int NTHREADS = NCONS + NHELP;
constexpr int EX = TX + 1, EY = TY + 1, EL = EX * EY;
constexpr int UX = TX + 2, UY = TY + 2, UL = UX * UY;
constexpr int kGeoStride = 194;        // doubles per element (8*24 + 2 pad; 16-byte stride 97, odd)
constexpr int kDumpStride = 82;        // doubles per storing thread (81 + 1 pad; 16-byte stride 41, odd)
constexpr int NWRITER = NP * 3;

enum { BAR_FULL0 = 1, BAR_FULL1 = 2, BAR_EMPTY0 = 3, BAR_EMPTY1 = 4, BAR_DUMP_FULL = 5, BAR_DUMP_EMPTY = 6, BAR_HELP = 7 };

struct DumpHdr {       // where one storing thread's 81 values go
  double *row0;        // &vals[slot of (row i = 0, block position 0)], nullptr = nothing to store
  int stride;          // doubles between the node's scalar rows: 3 * nb
  unsigned char k[12]; // block position per canonical slot of the group (255 = absent)
};

struct Smem {
  double geo[2][EL * kGeoStride];
  double dump[NWRITER * kDumpStride];
  double xlay[4][UL][3];
  double ulay[4][UL][3];
  double esc[4][EL];
  double fcarry[NP][3];
  HexRowMeta rowmeta[4][NP];
  DumpHdr hdr[NWRITER];
  int nid[8][UL];
};

__device__ __forceinline__ void bar_sync(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }
__device__ __forceinline__ void bar_arrive(int id, int n) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(n) : "memory"); }
__device__ __forceinline__ void cp_async8(void *smem, const void *gmem) {
  const unsigned sa = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(sa), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async4(void *smem, const void *gmem) {
  const unsigned sa = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sa), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }
__device__ __forceinline__ int hs_index(int sx, int sy) { return sy ? (sx ? 2 : 3) : (sx ? 1 : 0); }

// ------------------------------------------------------------------------------------------ helpers
struct Tile {
  int i0, j0, l0, l1;
};

__device__ __forceinline__ long long cell_of(const DevHex &hx, const Tile &T, int t, int lay, bool &ok) {
  const int ni = T.i0 - 1 + t % UX, nj = T.j0 - 1 + t / UX;
  ok = lay >= 0 && lay < hx.NK && ni >= 0 && nj >= 0 && ni < hx.NI && nj < hx.NJ;
  return ((long long)lay * hx.NJ + nj) * hx.NI + ni;
}

template <bool WITH_F>
__device__ __forceinline__ void fetch_layer(Smem &S, const DevHex &hx, const Tile &T, const double *u, int lay, int h) {
  const int buf = lay & 3;
  for (int t = h; t < UL; t += NHELP) {
    bool ok;
    const long long c = cell_of(hx, T, t, lay, ok);
    double *xb = S.xlay[buf][t];
    if (ok) {
#pragma unroll
      for (int d = 0; d < 3; ++d) cp_async8(xb + d, hx.pos_x + c * 3 + d);
    } else {
      xb[0] = 0.0; xb[1] = 0.0; xb[2] = 0.0;
    }
    if (WITH_F) {
      double *ub = S.ulay[buf][t];
      const int r = S.nid[lay & 7][t];
      if (r >= 0) {
#pragma unroll
        for (int d = 0; d < 3; ++d) cp_async8(ub + d, u + (long long)r * 3 + d);
      } else {
        ub[0] = 0.0; ub[1] = 0.0; ub[2] = 0.0;
      }
    }
    const int ux = t % UX, uy = t / UX;
    if (ux >= 1 && ux <= TX && uy >= 1 && uy <= TY) {
      HexRowMeta *rm = &S.rowmeta[buf][(uy - 1) * TX + (ux - 1)];
      if (ok) {
        const double *src = reinterpret_cast<const double *>(hx.pos_meta + c);
#pragma unroll
        for (int d = 0; d < 6; ++d) cp_async8(reinterpret_cast<double *>(rm) + d, src + d);
      } else {
        rm->nb = 0;
      }
    }
  }
}

__device__ __forceinline__ void fetch_nid(Smem &S, const DevHex &hx, const Tile &T, int lay, int h) {
  for (int t = h; t < UL; t += NHELP) {
    bool ok;
    const long long c = cell_of(hx, T, t, lay, ok);
    if (ok) cp_async4(&S.nid[lay & 7][t], hx.node_at + c);
    else S.nid[lay & 7][t] = -1;
  }
}

__device__ __forceinline__ void fetch_esc(Smem &S, const DevHex &hx, const Tile &T, int lay, int h) {
  for (int t = h; t < EL; t += NHELP) {
    const int ei = T.i0 - 1 + t % EX, ej = T.j0 - 1 + t / EX;
    if (lay >= 0 && lay < hx.NK - 1 && ei >= 0 && ej >= 0 && ei < hx.NI && ej < hx.NJ)
      cp_async8(&S.esc[lay & 3][t], hx.cellmu + ((long long)lay * hx.NJ + ej) * hx.NI + ei);
    else
      S.esc[lay & 3][t] = 0.0;
  }
}

// scaled physical gradients of element layer m into geo[buf]: one (element, ip) item per thread pass
__device__ __forceinline__ void geometry(Smem &S, int m, int buf, int h) {
  const double *xlo = &S.xlay[m & 3][0][0], *xhi = &S.xlay[(m + 1) & 3][0][0];
  const double *escm = S.esc[m & 3];
  for (int item = h; item < EL * 8; item += NHELP) {
    const int es = item % EL, ip = item / EL;
    const double smu = escm[es];
    if (smu == 0.0) continue;
    const int ex = es % EX, ey = es / EX;
    // integration point ip sits beside node ip (SamPtsHex, src/m_elems.F90:500-507)
    const bool ppx = (ip == 1) | (ip == 2) | (ip == 5) | (ip == 6), ppy = (ip == 2) | (ip == 3) | (ip == 6) | (ip == 7);
    const double e = ppx ? FEMX_G : -FEMX_G, n = ppy ? FEMX_G : -FEMX_G, s = ip >= 4 ? FEMX_G : -FEMX_G;
    const double fx[2] = {1.0 - e, 1.0 + e}, fy[2] = {1.0 - n, 1.0 + n}, fz[2] = {1.0 - s, 1.0 + s};
    double J[3][3];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) J[i][j] = 0.0;
#pragma unroll
    for (int b = 0; b < 8; ++b) {
      const int bx = (b == 1 || b == 2 || b == 5 || b == 6) ? 1 : 0, by = ((b & 3) >= 2) ? 1 : 0, bz = b >> 2;
      const double *xs = (bz ? xhi : xlo) + ((ey + by) * UX + (ex + bx)) * 3;
      const double X0 = xs[0], X1 = xs[1], X2 = xs[2];
      // ShapeFuncdHex, src/m_elems.F90:574-579
      const double d0 = (bx ? 0.125 : -0.125) * fy[by] * fz[bz];
      const double d1 = (by ? 0.125 : -0.125) * fx[bx] * fz[bz];
      const double d2 = (bz ? 0.125 : -0.125) * fx[bx] * fy[by];
      J[0][0] = fma(d0, X0, J[0][0]); J[0][1] = fma(d0, X1, J[0][1]); J[0][2] = fma(d0, X2, J[0][2]);
      J[1][0] = fma(d1, X0, J[1][0]); J[1][1] = fma(d1, X1, J[1][1]); J[1][2] = fma(d1, X2, J[1][2]);
      J[2][0] = fma(d2, X0, J[2][0]); J[2][1] = fma(d2, X1, J[2][1]); J[2][2] = fma(d2, X2, J[2][2]);
    }
    double adj[3][3];   // inverse * det
    adj[0][0] = J[1][1] * J[2][2] - J[1][2] * J[2][1];
    adj[0][1] = J[0][2] * J[2][1] - J[0][1] * J[2][2];
    adj[0][2] = J[0][1] * J[1][2] - J[0][2] * J[1][1];
    adj[1][0] = J[1][2] * J[2][0] - J[1][0] * J[2][2];
    adj[1][1] = J[0][0] * J[2][2] - J[0][2] * J[2][0];
    adj[1][2] = J[0][2] * J[1][0] - J[0][0] * J[1][2];
    adj[2][0] = J[1][0] * J[2][1] - J[1][1] * J[2][0];
    adj[2][1] = J[0][1] * J[2][0] - J[0][0] * J[2][1];
    adj[2][2] = J[0][0] * J[1][1] - J[0][1] * J[1][0];
    const double det = J[0][0] * adj[0][0] + J[0][1] * adj[1][0] + J[0][2] * adj[2][0];
    const double sc = smu * sqrt(1.0 / det);   // hex weights are 1 (src/m_elems.F90:508)
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) adj[i][j] *= sc;
    double *g = &S.geo[buf][es * kGeoStride + ip * 24];
#pragma unroll
    for (int b = 0; b < 8; b += 2) {
      double v[2][3];
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        const int bb = b + q;
        const int bx = (bb == 1 || bb == 2 || bb == 5 || bb == 6) ? 1 : 0, by = ((bb & 3) >= 2) ? 1 : 0, bz = bb >> 2;
        const double d0 = (bx ? 0.125 : -0.125) * fy[by] * fz[bz];
        const double d1 = (by ? 0.125 : -0.125) * fx[bx] * fz[bz];
        const double d2 = (bz ? 0.125 : -0.125) * fx[bx] * fy[by];
#pragma unroll
        for (int i = 0; i < 3; ++i) v[q][i] = adj[i][0] * d0 + adj[i][1] * d1 + adj[i][2] * d2;
      }
#pragma unroll
      for (int i = 0; i < 3; ++i) *reinterpret_cast<double2 *>(g + i * 8 + b) = make_double2(v[0][i], v[1][i]);
    }
  }
}

// global stores of one dump: helper warp hw moves the 81 values of 12 storing threads, 32 lanes per pass.
// Branch-free (predicated stores) so that the shared-memory reads of four storing threads overlap.
__device__ __forceinline__ void store_dump(Smem &S, int hw, int lane, const int (&lane_i)[3], const int (&lane_s)[3],
                                           const int (&lane_j)[3]) {
#pragma unroll 4
  for (int q = 0; q < NWRITER / 12; ++q) {
    const int w = hw * (NWRITER / 12) + q;
    const DumpHdr &hd = S.hdr[w];
    double *row0 = hd.row0;
    const int stride = hd.stride;
    const double *src = &S.dump[w * kDumpStride];
    int k[3];
    double v[3];
#pragma unroll
    for (int it = 0; it < 3; ++it) {
      k[it] = hd.k[lane_s[it]];
      v[it] = src[min(it * 32 + lane, 80)];
    }
#pragma unroll
    for (int it = 0; it < 3; ++it) {
      const bool on = row0 != nullptr && (it * 32 + lane) < 81 && k[it] != 255;
      double *dst = on ? row0 + (lane_i[it] * stride + 3 * k[it] + lane_j[it]) : nullptr;
      if (on) *dst = v[it];
    }
  }
}

}  // namespace

template <bool WITH_F>
__global__ void __launch_bounds__(NTHREADS, 1)
hex_ws_kernel(DevHex hx, const double *__restrict__ u, double *__restrict__ vals, double *__restrict__ fint, int LZ) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Smem &S = *reinterpret_cast<Smem *>(smem_raw);
  const int tid = threadIdx.x;
  Tile T;
  T.i0 = blockIdx.x * TX; T.j0 = blockIdx.y * TY;
  T.l0 = blockIdx.z * LZ; T.l1 = min(T.l0 + LZ, hx.NK) - 1;
  const int nsteps = T.l1 - T.l0 + 2;   // element layers l0-1 .. l1

  if (tid >= NCONS) {
    // =========================================================================== helper warps
    asm volatile("setmaxnreg.dec.sync.aligned.u32 88;");
    const int h = tid - NCONS, lane = h & 31, hw = h >> 5;
    int lane_i[3], lane_s[3], lane_j[3];
#pragma unroll
    for (int it = 0; it < 3; ++it) {
      const int v = min(it * 32 + lane, 80);
      lane_i[it] = v / 27; lane_s[it] = (v % 27) / 3; lane_j[it] = v % 3;
    }
    if (WITH_F && h < NP * 3) (&S.fcarry[0][0])[h] = 0.0;
    fetch_nid(S, hx, T, T.l0 - 1, h); fetch_nid(S, hx, T, T.l0, h); fetch_nid(S, hx, T, T.l0 + 1, h);
    fetch_esc(S, hx, T, T.l0 - 1, h);
    cp_async_wait_all();
    bar_sync(BAR_HELP, NHELP);
    fetch_layer<WITH_F>(S, hx, T, u, T.l0 - 1, h); fetch_layer<WITH_F>(S, hx, T, u, T.l0, h);
    cp_async_wait_all();
    bar_sync(BAR_HELP, NHELP);
    for (int it = 0; it < nsteps; ++it) {
      const int m = T.l0 - 1 + it;
      if (it >= 2) {
        bar_sync(BAR_DUMP_FULL, NTHREADS);        // dump of step it-2 is complete
        store_dump(S, hw, lane, lane_i, lane_s, lane_j);
        __threadfence_block();
        bar_arrive(BAR_DUMP_EMPTY, NTHREADS);
        bar_sync((it & 1) ? BAR_EMPTY1 : BAR_EMPTY0, NTHREADS);   // consumers are done with step it-2
      }
      // prefetch for the next iteration (layer m+2 data, ids of m+3, cell scale of m+1); the ring slots
      // they overwrite belonged to step it-2, which the consumers have left
      fetch_nid(S, hx, T, m + 3, h);
      fetch_layer<WITH_F>(S, hx, T, u, m + 2, h);
      fetch_esc(S, hx, T, m + 1, h);
      geometry(S, m, it & 1, h);
      __threadfence_block();
      bar_arrive((it & 1) ? BAR_FULL1 : BAR_FULL0, NTHREADS);
      cp_async_wait_all();
      bar_sync(BAR_HELP, NHELP);
    }
    // drain: dumps of the last two steps
    if (nsteps >= 2) {
      bar_sync(BAR_DUMP_FULL, NTHREADS);
      store_dump(S, hw, lane, lane_i, lane_s, lane_j);
      bar_arrive(BAR_DUMP_EMPTY, NTHREADS);
    }
    bar_sync(BAR_DUMP_FULL, NTHREADS);
    store_dump(S, hw, lane, lane_i, lane_s, lane_j);
    return;
  }

  // ============================================================================= consumer warps
  asm volatile("setmaxnreg.inc.sync.aligned.u32 240;");
  const int lane = tid & 31;
  const int p = tid >> 2, role = tid & 3;
  const int px = p % TX, py = p / TX;
  const double rho = hx.rho;
  double acc[9][9];
#pragma unroll
  for (int s = 0; s < 9; ++s)
#pragma unroll
    for (int q = 0; q < 9; ++q) acc[s][q] = 0.0;

  for (int it = 0; it < nsteps; ++it) {
    const int m = T.l0 - 1 + it;
    const int par = (m + 2) & 1;
    const bool isA = (role == 1 + par), isB = (role == 2 - par);
    const bool upper = role == 0 || isB;                 // works on node layer m+1
    const int lay = upper ? m + 1 : m;
    const int atop = upper ? 4 : 0;                      // our node sits on the top face of layer-m elements
    const int btop = (role == 3 || isB) ? 4 : 0;
    if (!isA) {
#pragma unroll
      for (int s = 0; s < 9; ++s)
#pragma unroll
        for (int q = 0; q < 9; ++q) acc[s][q] = 0.0;
    }
    bar_sync((it & 1) ? BAR_FULL1 : BAR_FULL0, NTHREADS);      // gradients of layer m (and its node data) are in
    const HexRowMeta *rm = &S.rowmeta[lay & 3][p];
    const bool active = rm->nb > 0 && lay >= T.l0 && lay <= T.l1;
    const double *geo = S.geo[it & 1];
    const double *escm = S.esc[m & 3];
    if (active) {
#pragma unroll
      for (int oy = 0; oy < 2; ++oy)
#pragma unroll
        for (int ox = 0; ox < 2; ++ox) {
          const int es = (py + oy) * EX + (px + ox);
          if (escm[es] == 0.0) continue;
          const double *g = geo + es * kGeoStride;
          const int a = hs_index(1 - ox, 1 - oy) + atop;
#pragma unroll 2
          for (int ip = 0; ip < 8; ++ip) {
            const double *gp = g + ip * 24;
            double ga[3], gb[3][4];
#pragma unroll
            for (int i = 0; i < 3; ++i) {
              ga[i] = gp[i * 8 + a];
              const double2 v0 = *reinterpret_cast<const double2 *>(gp + i * 8 + btop);
              const double2 v1 = *reinterpret_cast<const double2 *>(gp + i * 8 + btop + 2);
              gb[i][0] = v0.x; gb[i][1] = v0.y; gb[i][2] = v1.x; gb[i][3] = v1.y;
            }
#pragma unroll
            for (int bb = 0; bb < 4; ++bb) {
              const int bx = (bb == 1 || bb == 2) ? 1 : 0, by = bb >= 2 ? 1 : 0;
              const int s = (by + oy) * 3 + (bx + ox);
#pragma unroll
              for (int i = 0; i < 3; ++i)
#pragma unroll
                for (int j = 0; j < 3; ++j) acc[s][i * 3 + j] = fma(ga[i], gb[j][bb], acc[s][i * 3 + j]);
            }
          }
        }
    }
    const bool writer = active && !isB;
    const int grp = role == 0 ? 0 : (role == 3 ? 2 : 1);       // dz = -1, 0, +1
    double fpart[3] = {0.0, 0.0, 0.0};
    if (writer) {
      // G' -> K per block: K = rho G' + G'^T + tr(G') I
#pragma unroll
      for (int s = 0; s < 9; ++s) {
        const double tr = acc[s][0] + acc[s][4] + acc[s][8];
        double k9[9];
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
          for (int j = 0; j < 3; ++j) k9[i * 3 + j] = fma(rho, acc[s][i * 3 + j], acc[s][j * 3 + i]) + (i == j ? tr : 0.0);
#pragma unroll
        for (int q = 0; q < 9; ++q) acc[s][q] = k9[q];
      }
      if (WITH_F) {
        const double *ub = &S.ulay[(lay + grp - 1) & 3][0][0];
        double fp[3][3];   // three independent chains per component (fixed order), summed at the end
#pragma unroll
        for (int sy = 0; sy < 3; ++sy) {
#pragma unroll
          for (int i = 0; i < 3; ++i) fp[sy][i] = 0.0;
#pragma unroll
          for (int sx = 0; sx < 3; ++sx) {
            const int s = sy * 3 + sx;
            const double *uc = ub + ((py + sy) * UX + (px + sx)) * 3;
            const double u0 = uc[0], u1 = uc[1], u2 = uc[2];
#pragma unroll
            for (int i = 0; i < 3; ++i) fp[sy][i] += acc[s][i * 3] * u0 + acc[s][i * 3 + 1] * u1 + acc[s][i * 3 + 2] * u2;
          }
        }
#pragma unroll
        for (int i = 0; i < 3; ++i) fpart[i] = (fp[0][i] + fp[1][i]) + fp[2][i];
      }
    }
    // header of this thread's dump (read before the layer ring slot can be recycled)
    DumpHdr hd;
    hd.row0 = writer ? vals + rm->base : nullptr;
    hd.stride = 3 * rm->nb;
#pragma unroll
    for (int s = 0; s < 9; ++s) hd.k[s] = writer ? rm->k[grp * 9 + s] : (unsigned char)255;
    if (WITH_F) {
      // f of the lower node = carry(dz=-1, previous step) + dz=0 (finisher) + dz=+1 (role 3)
      const int src3 = (lane & ~3) + 3;
      double f3[3];
#pragma unroll
      for (int i = 0; i < 3; ++i) f3[i] = __shfl_sync(0xffffffffu, fpart[i], src3);
      if (isA && active) {
#pragma unroll
        for (int i = 0; i < 3; ++i) fint[(long long)rm->grow + i] = (S.fcarry[p][i] + fpart[i]) + f3[i];
      }
      __syncwarp();
      if (role == 0) {
#pragma unroll
        for (int i = 0; i < 3; ++i) S.fcarry[p][i] = active ? fpart[i] : 0.0;
      }
    }
    __threadfence_block();
    bar_arrive((it & 1) ? BAR_EMPTY1 : BAR_EMPTY0, NTHREADS);  // geo[it&1] and this step's layer data are free
    if (it >= 1) bar_sync(BAR_DUMP_EMPTY, NTHREADS);            // the previous dump has left
    {
      const int w = p * 3 + grp;
      if (!isB) {
        S.hdr[w] = hd;
        if (writer) {
          double2 *dst = reinterpret_cast<double2 *>(&S.dump[w * kDumpStride]);
#pragma unroll
          for (int q = 0; q < 40; ++q) {
            const int v0 = 2 * q, v1 = 2 * q + 1;      // v = i*27 + slot*3 + j
            dst[q] = make_double2(acc[(v0 % 27) / 3][(v0 / 27) * 3 + v0 % 3], acc[(v1 % 27) / 3][(v1 / 27) * 3 + v1 % 3]);
          }
          S.dump[w * kDumpStride + 80] = acc[8][8];
        }
      }
    }
    __threadfence_block();
    bar_arrive(BAR_DUMP_FULL, NTHREADS);
  }
}

int launch_hex_ws(femx_ctx *c, const double *d_u, double *d_vals, double *d_fint) {
  const DevHex &hx = c->dhex;
  const int want = c->opts.reserved[1] > 0 ? c->opts.reserved[1] : 25;   // node layers per z-chunk
  const int nchunk = (hx.NK + want / 2) / want > 0 ? (hx.NK + want / 2) / want : 1;
  const int LZ = (hx.NK + nchunk - 1) / nchunk;
  dim3 grid((hx.NI + TX - 1) / TX, (hx.NJ + TY - 1) / TY, (hx.NK + LZ - 1) / LZ);
  const size_t smem = sizeof(Smem);
  if (d_u) {
    cudaFuncSetAttribute(hex_ws_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    hex_ws_kernel<true><<<grid, NTHREADS, smem, c->stream>>>(hx, d_u, d_vals, d_fint, LZ);
  } else {
    cudaFuncSetAttribute(hex_ws_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    hex_ws_kernel<false><<<grid, NTHREADS, smem, c->stream>>>(hx, d_u, d_vals, d_fint, LZ);
  }
  return 1;
}

}  // namespace femx
