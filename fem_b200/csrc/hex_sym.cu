// hex_sym.cu -- assembly variant 4: symmetric pair-block fused hex8 K (+ f_int) kernel.
//
// Variants 2/3 give every node row its own threads, which recompute K_cr = K_rc^T on both sides and
// re-read every scaled gradient ~10 times from shared memory (2.4 FMA per loaded double: the LSU
// wavefront pipe, not the FP64 pipe, bounds them).  Here the unit of FP64 work is the ELEMENT:
//   * four threads per element accumulate, over the 8 integration points, the 36 unique node-pair
//     blocks G'_ab = sum_ip g~_a g~_b^T (a <= b) in registers -- 10 pairs of the bottom/top face
//     (78 accumulators: off-diagonal 3x3, diagonal symmetric 6) and 2 x 8 bottom x top pairs
//     (72 accumulators) -- 5 FMA per loaded double, every operand an LDS.128;
//   * face pairs are summed over the two elements below/above the face in registers (the two face
//     warps alternate "start"/"finish" from layer to layer), turned into K_ab = rho G' + G'^T + tr I
//     and parked in a per-element shared-memory scratch (234 doubles);
//   * a gather pass, lane <-> consecutive CSR value, sums the <= 4 element contributions of every
//     value of a node row in fixed (ascending cell) order through a 243-entry offset table, takes the
//     transposed entry for c < r, forms f_r += K_rc u_c on the way and stores coalesced 216-byte runs.
// One CTA (256 threads, 1 per SM) owns 7 x 7 node positions, i.e. the 8 x 8 cells touching them, and
// marches through the node layers of a z-chunk.  No atomics, fixed summation order, each CSR value
// stored exactly once.  Same arithmetic as FormElKE (src/m_local.F90:21-67) in the (lambda, mu)
// block form of elem_math.cuh.
#include "elem_math.cuh"
#include "launch.h"

namespace femx {

namespace {

constexpr int NX = 7, NY = 7, NP = NX * NY;   // owned node positions of a tile
constexpr int EX = NX + 1, EY = NY + 1, EL = EX * EY;   // cells per layer (all cells touching the tile)
constexpr int UX = NX + 2, UY = NY + 2, UL = UX * UY;   // node positions per layer incl. halo
constexpr int NT = 4 * EL;                    // 256 threads: 4 per cell
constexpr int NW = NT / 32;
constexpr int GS = 194;                       // doubles per cell in the gradient buffer (8 ip x 24 + 2; 16-byte stride odd)
constexpr int SS = 234;                       // doubles per cell of block scratch: 10 face blocks + 16 bottom x top blocks
constexpr int BT_OFF = 90;
static_assert(EL == 64 && NT == 256, "thread mapping assumes 64 cells per layer");

struct Smem {
  union {
    double geo[EL * GS];   // scaled gradients [cell][ip][i][node]
    double scr[EL * SS];   // K blocks [cell][block][i][j] (written after the FMA phase has left geo)
  } a;
  double xlay[4][UL][3];
  double ulay[4][UL][3];
  double esc[4][EL];
  double fcarry[NP][3];
  HexRowMeta rowmeta[4][NP];
  uint2 tab[9][32];        // gather table: [i*3+g][lane] -> four 16-bit scratch offsets (0xffff = none)
  int nid[8][UL];
};

__device__ __forceinline__ void cp_async8(void *smem, const void *gmem) {
  const unsigned sa = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(sa), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async4(void *smem, const void *gmem) {
  const unsigned sa = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sa), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

__host__ __device__ constexpr int face_index(int lx, int ly) { return ly ? (lx ? 2 : 3) : (lx ? 1 : 0); }
__host__ __device__ constexpr int pair_q(int a, int b) { return a * 4 - a * (a - 1) / 2 + (b - a); }   // a <= b, 0..9
__host__ __device__ constexpr int offdiag_p(int a, int b) { return a == 0 ? b - 1 : (a == 1 ? b + 1 : 5); }   // a < b, 0..5
__host__ __device__ constexpr int sym_s(int i, int j) { return i == 0 ? j : (i == 1 ? j + 2 : 5); }     // i <= j, 0..5

struct Tile {
  int i0, j0, l0, l1;
};

__device__ __forceinline__ long long cell_of(const DevHex &hx, const Tile &T, int t, int lay, bool &ok) {
  const int ni = T.i0 - 1 + t % UX, nj = T.j0 - 1 + t / UX;
  ok = lay >= 0 && lay < hx.NK && ni >= 0 && nj >= 0 && ni < hx.NI && nj < hx.NJ;
  return ((long long)lay * hx.NJ + nj) * hx.NI + ni;
}

template <bool WITH_F>
__device__ __forceinline__ void fetch_layer(Smem &S, const DevHex &hx, const Tile &T, const double *u, int lay, int tid) {
  const int buf = lay & 3;
  for (int t = tid; t < UL; t += NT) {
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
    if (ux >= 1 && ux <= NX && uy >= 1 && uy <= NY) {
      HexRowMeta *rm = &S.rowmeta[buf][(uy - 1) * NX + (ux - 1)];
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

__device__ __forceinline__ void fetch_nid(Smem &S, const DevHex &hx, const Tile &T, int lay, int tid) {
  for (int t = tid; t < UL; t += NT) {
    bool ok;
    const long long c = cell_of(hx, T, t, lay, ok);
    if (ok) cp_async4(&S.nid[lay & 7][t], hx.node_at + c);
    else S.nid[lay & 7][t] = -1;
  }
}

__device__ __forceinline__ void fetch_esc(Smem &S, const DevHex &hx, const Tile &T, int lay, int tid) {
  for (int t = tid; t < EL; t += NT) {
    const int ei = T.i0 - 1 + t % EX, ej = T.j0 - 1 + t / EX;
    if (lay >= 0 && lay < hx.NK - 1 && ei >= 0 && ej >= 0 && ei < hx.NI && ej < hx.NJ)
      cp_async8(&S.esc[lay & 3][t], hx.cellmu + ((long long)lay * hx.NJ + ej) * hx.NI + ei);
    else
      S.esc[lay & 3][t] = 0.0;
  }
}

// gather table entry of lane l (value (kk, j) of a 27-value run) for scalar row i and group g = dz + 1
__device__ __forceinline__ uint2 table_entry(int i, int g, int l) {
  unsigned o[4] = {0xffffu, 0xffffu, 0xffffu, 0xffffu};
  if (l < 27) {
    const int kk = l / 3, j = l % 3, dy = kk / 3 - 1, dx = kk % 3 - 1;
    int n = 0;
    for (int oy = 0; oy < 2; ++oy)
      for (int ox = 0; ox < 2; ++ox) {
        const int lx = 1 - ox, ly = 1 - oy, cx = lx + dx, cy = ly + dy;   // row / column node inside cell (px+ox, py+oy)
        if (cx < 0 || cx > 1 || cy < 0 || cy > 1) continue;
        const int fa = face_index(lx, ly), fb = face_index(cx, cy);
        int off;
        if (g == 1) off = fa <= fb ? pair_q(fa, fb) * 9 + i * 3 + j : pair_q(fb, fa) * 9 + j * 3 + i;
        else if (g == 2) off = BT_OFF + (fa * 4 + fb) * 9 + i * 3 + j;     // row node on the bottom face
        else off = BT_OFF + (fb * 4 + fa) * 9 + j * 3 + i;                 // row node on the top face: transposed block
        o[n++] = (unsigned)((oy * EX + ox) * SS + off);
      }
  }
  return make_uint2(o[0] | (o[1] << 16), o[2] | (o[3] << 16));
}

// scaled physical gradients g~_b(ip) = adj(J)^T grad N_b * sqrt(mu w / detJ) of cell layer m (FormdNdetJ,
// src/m_local.F90:175-189, with the closed-form adjugate instead of dgetrf/dgetri)
__device__ __forceinline__ void geometry(Smem &S, int m, int tid) {
  const double *xlo = &S.xlay[m & 3][0][0], *xhi = &S.xlay[(m + 1) & 3][0][0];
  const int es = tid & (EL - 1);
  const double smu = S.esc[m & 3][es];
  if (smu == 0.0) return;
  const int ex = es % EX, ey = es / EX;
#pragma unroll 1
  for (int q = 0; q < 2; ++q) {
    const int ip = (tid >> 6) + 4 * q;
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
    double *g = &S.a.geo[es * GS + ip * 24];
#pragma unroll
    for (int b = 0; b < 8; b += 2) {
      double v[2][3];
#pragma unroll
      for (int qq = 0; qq < 2; ++qq) {
        const int bb = b + qq;
        const int bx = (bb == 1 || bb == 2 || bb == 5 || bb == 6) ? 1 : 0, by = ((bb & 3) >= 2) ? 1 : 0, bz = bb >> 2;
        const double d0 = (bx ? 0.125 : -0.125) * fy[by] * fz[bz];
        const double d1 = (by ? 0.125 : -0.125) * fx[bx] * fz[bz];
        const double d2 = (bz ? 0.125 : -0.125) * fx[bx] * fy[by];
#pragma unroll
        for (int i = 0; i < 3; ++i) v[qq][i] = adj[i][0] * d0 + adj[i][1] * d1 + adj[i][2] * d2;
      }
#pragma unroll
      for (int i = 0; i < 3; ++i) *reinterpret_cast<double2 *>(g + i * 8 + b) = make_double2(v[0][i], v[1][i]);
    }
  }
}

// ---------------------------------------------------------------------------------- FMA phase
__device__ __forceinline__ void load_face(const double *gp, double (&v)[3][4]) {
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    const double2 a = *reinterpret_cast<const double2 *>(gp + i * 8);
    const double2 b = *reinterpret_cast<const double2 *>(gp + i * 8 + 2);
    v[i][0] = a.x; v[i][1] = a.y; v[i][2] = b.x; v[i][3] = b.y;
  }
}
// 10 pair blocks among the four nodes of a face: 6 off-diagonal (9 entries), 4 diagonal (6 entries)
__device__ __forceinline__ void fma_face(const double (&v)[3][4], double (&acc)[78]) {
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = a + 1; b < 4; ++b)
#pragma unroll
      for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) {
          double &c = acc[offdiag_p(a, b) * 9 + i * 3 + j];
          c = fma(v[i][a], v[j][b], c);
        }
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = i; j < 3; ++j) {
        double &c = acc[54 + a * 6 + sym_s(i, j)];
        c = fma(v[i][a], v[j][a], c);
      }
}
__device__ __forceinline__ void load_bt(const double *gp, int a0, double (&va)[3][2], double (&vb)[3][4]) {
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    const double2 a = *reinterpret_cast<const double2 *>(gp + i * 8 + a0);
    const double2 b = *reinterpret_cast<const double2 *>(gp + i * 8 + 4);
    const double2 c = *reinterpret_cast<const double2 *>(gp + i * 8 + 6);
    va[i][0] = a.x; va[i][1] = a.y;
    vb[i][0] = b.x; vb[i][1] = b.y; vb[i][2] = c.x; vb[i][3] = c.y;
  }
}
// 2 bottom nodes x 4 top nodes
__device__ __forceinline__ void fma_bt(const double (&va)[3][2], const double (&vb)[3][4], double (&acc)[78]) {
#pragma unroll
  for (int a = 0; a < 2; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b)
#pragma unroll
      for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) {
          double &c = acc[(a * 4 + b) * 9 + i * 3 + j];
          c = fma(va[i][a], vb[j][b], c);
        }
}
// G' -> K in place: K = rho G' + G'^T + tr(G') I  (elem_math.cuh block form of B_a^T D B_b)
__device__ __forceinline__ void to_k_block(double *g, double rho) {
  const double tr = g[0] + g[4] + g[8];
  double k9[9];
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) k9[i * 3 + j] = fma(rho, g[i * 3 + j], g[j * 3 + i]) + (i == j ? tr : 0.0);
#pragma unroll
  for (int q = 0; q < 9; ++q) g[q] = k9[q];
}
__device__ __forceinline__ void to_k_sym(double *g, double rho1) {   // symmetric G' (6 entries), rho1 = rho + 1
  const double tr = g[0] + g[3] + g[5];
#pragma unroll
  for (int s = 0; s < 6; ++s) g[s] *= rho1;
  g[0] += tr; g[3] += tr; g[5] += tr;
}
// value at flat offset `off` (0..89) of a face's scratch area
__device__ __forceinline__ double face_out(const double (&acc)[78], int off) {
  const int q = off / 9, r = off % 9, i = r / 3, j = r % 3;
  const int a = q < 4 ? 0 : (q < 7 ? 1 : (q < 9 ? 2 : 3));
  const int b = a + (q - pair_q(a, a));
  if (a == b) return acc[54 + a * 6 + sym_s(i < j ? i : j, i < j ? j : i)];
  return acc[offdiag_p(a, b) * 9 + r];
}

// ---------------------------------------------------------------------------------- gather phase
template <int I, int G, bool WITH_F>
__device__ __forceinline__ double gather_pass(const double *eb, const uint2 t, const HexRowMeta *rm, bool act, int kk, int j,
                                              double *__restrict__ vals, const double *ub) {
  const unsigned o0 = t.x & 0xffffu, o1 = t.x >> 16, o2 = t.y & 0xffffu, o3 = t.y >> 16;
  double v = 0.0;
  if (o0 != 0xffffu) v = eb[o0];
  if (o1 != 0xffffu) v += eb[o1];
  if (o2 != 0xffffu) v += eb[o2];
  if (o3 != 0xffffu) v += eb[o3];
  const int k = rm->k[G * 9 + kk];
  if (act && o0 != 0xffffu && k != 255) vals[rm->base + (long long)(I * 3) * rm->nb + 3 * k + j] = v;
  if (WITH_F) return v * ub[0];
  return 0.0;
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int d = 16; d >= 1; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
  return v;
}

}  // namespace

template <bool WITH_F>
__global__ void __launch_bounds__(NT, 1)
hex_sym_kernel(DevHex hx, const double *__restrict__ u, double *__restrict__ vals, double *__restrict__ fint, int LZ) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Smem &S = *reinterpret_cast<Smem *>(smem_raw);
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  Tile T;
  T.i0 = blockIdx.x * NX; T.j0 = blockIdx.y * NY;
  T.l0 = blockIdx.z * LZ; T.l1 = min(T.l0 + LZ, hx.NK) - 1;
  const int nsteps = T.l1 - T.l0 + 2;   // cell layers l0-1 .. l1
  const double rho = hx.rho;

  // FMA-phase role: warps 0,1 / 2,3 = the two alternating face roles, 4,5 = bottom nodes 0,1 x top, 6,7 = bottom nodes 2,3 x top
  const int type = w >> 1;
  const int es = (w & 1) * 32 + lane;
  // gather-phase lane constants
  const int gl = lane < 27 ? lane : 26;
  const int kk = gl / 3, jj = gl % 3;
  const int uoff = ((kk / 3) * UX + (kk % 3)) * 3 + jj;   // halo'd node (px + dx + 1, py + dy + 1), component j

  if (tid < 243) S.tab[tid / 27][tid % 27] = table_entry((tid / 27) / 3, (tid / 27) % 3, tid % 27);
  if (tid < 9 * 5) S.tab[tid / 5][27 + tid % 5] = make_uint2(0xffffffffu, 0xffffffffu);
  if (WITH_F && tid < NP * 3) (&S.fcarry[0][0])[tid] = 0.0;
  fetch_nid(S, hx, T, T.l0 - 1, tid); fetch_nid(S, hx, T, T.l0, tid); fetch_nid(S, hx, T, T.l0 + 1, tid);
  fetch_esc(S, hx, T, T.l0 - 1, tid);
  cp_async_wait_all();
  __syncthreads();
  fetch_layer<WITH_F>(S, hx, T, u, T.l0 - 1, tid); fetch_layer<WITH_F>(S, hx, T, u, T.l0, tid);

  double acc[78];
#pragma unroll
  for (int q = 0; q < 78; ++q) acc[q] = 0.0;

  for (int it = 0; it < nsteps; ++it) {
    const int m = T.l0 - 1 + it;
    cp_async_wait_all();
    __syncthreads();              // layer data of m, m+1 is in; the previous gather has left the scratch
    geometry(S, m, tid);
    __syncthreads();
    // prefetch for the next step: ids of m+3, node layer m+2, cell scale of m+1 (ring slots of step it-2)
    fetch_nid(S, hx, T, m + 3, tid);
    fetch_layer<WITH_F>(S, hx, T, u, m + 2, tid);
    fetch_esc(S, hx, T, m + 1, tid);

    // ------------------------------------------------------------------ FMA phase
    const bool cell_on = S.esc[m & 3][es] != 0.0;
    const double *g = &S.a.geo[es * GS];
    const bool finisher = type == (it & 1);        // face role: finisher adds the bottom face of layer m to the carried top face of m-1
    if (type < 2) {
      if (!finisher) {
#pragma unroll
        for (int q = 0; q < 78; ++q) acc[q] = 0.0;
      }
      const bool need = finisher ? it > 0 : it + 1 < nsteps;   // first finish / last start feed no owned row
      if (cell_on && need) {
        const double *gf = g + (finisher ? 0 : 4);
        double v0[3][4], v1[3][4];
        load_face(gf, v0);
#pragma unroll 1
        for (int ip = 0; ip < 8; ip += 2) {
          load_face(gf + (ip + 1) * 24, v1);
          fma_face(v0, acc);
          if (ip + 2 < 8) load_face(gf + (ip + 2) * 24, v0);
          fma_face(v1, acc);
        }
      }
    } else {
#pragma unroll
      for (int q = 0; q < 78; ++q) acc[q] = 0.0;
      if (cell_on) {
        const int a0 = (type - 2) * 2;
        double va0[3][2], vb0[3][4], va1[3][2], vb1[3][4];
        load_bt(g, a0, va0, vb0);
#pragma unroll 1
        for (int ip = 0; ip < 8; ip += 2) {
          load_bt(g + (ip + 1) * 24, a0, va1, vb1);
          fma_bt(va0, vb0, acc);
          if (ip + 2 < 8) load_bt(g + (ip + 2) * 24, a0, va0, vb0);
          fma_bt(va1, vb1, acc);
        }
      }
    }
    __syncthreads();              // everybody has left the gradient buffer
    // ------------------------------------------------------------------ G' -> K, park in the scratch
    {
      double *sc = &S.a.scr[es * SS];
      if (type < 2) {
        if (finisher) {
#pragma unroll
          for (int p = 0; p < 6; ++p) to_k_block(&acc[p * 9], rho);
#pragma unroll
          for (int a = 0; a < 4; ++a) to_k_sym(&acc[54 + a * 6], rho + 1.0);
#pragma unroll
          for (int q = 0; q < 45; ++q)
            *reinterpret_cast<double2 *>(sc + 2 * q) = make_double2(face_out(acc, 2 * q), face_out(acc, 2 * q + 1));
        }
      } else {
#pragma unroll
        for (int p = 0; p < 8; ++p) to_k_block(&acc[p * 9], rho);
        double *dst = sc + BT_OFF + (type - 2) * 72;
#pragma unroll
        for (int q = 0; q < 36; ++q) *reinterpret_cast<double2 *>(dst + 2 * q) = make_double2(acc[2 * q], acc[2 * q + 1]);
      }
    }
    __syncthreads();
    // ------------------------------------------------------------------ gather: rows of node layer m (dz = 0, +1) and m+1 (dz = -1)
    {
      uint2 tb[9];
#pragma unroll
      for (int q = 0; q < 9; ++q) tb[q] = S.tab[q][lane];
      const bool lay_lo = m >= T.l0 && m <= T.l1, lay_hi = m + 1 >= T.l0 && m + 1 <= T.l1;
      for (int n = w; n < NP; n += NW) {
        const int px = n % NX, py = n / NX;
        const double *eb = &S.a.scr[(py * EX + px) * SS];
        const HexRowMeta *rlo = &S.rowmeta[m & 3][n], *rhi = &S.rowmeta[(m + 1) & 3][n];
        const bool act_lo = lay_lo && rlo->nb > 0, act_hi = lay_hi && rhi->nb > 0;
        const double *ulo = &S.ulay[m & 3][0][0] + (py * UX + px) * 3 + uoff;
        const double *uhi = &S.ulay[(m + 1) & 3][0][0] + (py * UX + px) * 3 + uoff;
        double f_lo[3], f_hi[3];
        // group dz = 0 / +1 of the lower node: columns on layers m / m+1; group dz = -1 of the upper node: columns on layer m
        f_lo[0] = gather_pass<0, 1, WITH_F>(eb, tb[1], rlo, act_lo, kk, jj, vals, ulo) + gather_pass<0, 2, WITH_F>(eb, tb[2], rlo, act_lo, kk, jj, vals, uhi);
        f_lo[1] = gather_pass<1, 1, WITH_F>(eb, tb[4], rlo, act_lo, kk, jj, vals, ulo) + gather_pass<1, 2, WITH_F>(eb, tb[5], rlo, act_lo, kk, jj, vals, uhi);
        f_lo[2] = gather_pass<2, 1, WITH_F>(eb, tb[7], rlo, act_lo, kk, jj, vals, ulo) + gather_pass<2, 2, WITH_F>(eb, tb[8], rlo, act_lo, kk, jj, vals, uhi);
        f_hi[0] = gather_pass<0, 0, WITH_F>(eb, tb[0], rhi, act_hi, kk, jj, vals, ulo);
        f_hi[1] = gather_pass<1, 0, WITH_F>(eb, tb[3], rhi, act_hi, kk, jj, vals, ulo);
        f_hi[2] = gather_pass<2, 0, WITH_F>(eb, tb[6], rhi, act_hi, kk, jj, vals, ulo);
        if (WITH_F) {
#pragma unroll
          for (int i = 0; i < 3; ++i) { f_lo[i] = warp_sum(f_lo[i]); f_hi[i] = warp_sum(f_hi[i]); }
          if (lane < 3) {
            const double flo = lane == 0 ? f_lo[0] : (lane == 1 ? f_lo[1] : f_lo[2]);
            const double fhi = lane == 0 ? f_hi[0] : (lane == 1 ? f_hi[1] : f_hi[2]);
            // f of the lower node = (dz = -1 part carried from the previous step) + (dz = 0, +1 parts)
            if (act_lo) fint[(long long)rlo->grow + lane] = S.fcarry[n][lane] + flo;
            S.fcarry[n][lane] = act_hi ? fhi : 0.0;
          }
        }
      }
    }
  }
}

int launch_hex_sym(femx_ctx *c, const double *d_u, double *d_vals, double *d_fint) {
  const DevHex &hx = c->dhex;
  const int tiles = ((hx.NI + NX - 1) / NX) * ((hx.NJ + NY - 1) / NY);
  int nsm = 148;
  cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, c->device);
  // z-chunk length: fewest (waves x steps per chunk); every chunk pays one extra cell layer plus the prologue
  int LZ = c->opts.reserved[1] > 0 ? c->opts.reserved[1] : 0;
  if (LZ <= 0) {
    double best = 1e300;
    for (int nch = 1; nch <= hx.NK; ++nch) {
      const int lz = (hx.NK + nch - 1) / nch, chunks = (hx.NK + lz - 1) / lz;
      const long long ctas = (long long)tiles * chunks;
      const double waves = (double)((ctas + nsm - 1) / nsm);
      const double cost = waves * (lz + 2.0);
      if (cost < best) { best = cost; LZ = lz; }
    }
  }
  if (LZ > hx.NK) LZ = hx.NK;
  dim3 grid((hx.NI + NX - 1) / NX, (hx.NJ + NY - 1) / NY, (hx.NK + LZ - 1) / LZ);
  const size_t smem = sizeof(Smem);
  if (d_u) {
    cudaFuncSetAttribute(hex_sym_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    hex_sym_kernel<true><<<grid, NT, smem, c->stream>>>(hx, d_u, d_vals, d_fint, LZ);
  } else {
    cudaFuncSetAttribute(hex_sym_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    hex_sym_kernel<false><<<grid, NT, smem, c->stream>>>(hx, d_u, d_vals, d_fint, LZ);
  }
  return 1;
}

}  // namespace femx
