// hex_mma.cu -- assembly variant 20: warp-specialised, DMMA-based fused hex8 K (+ f_int) tile kernel.
//
// Same arithmetic as FormElKE (src/m_local.F90:21-67) in the (lambda, mu) block form of elem_math.cuh and the same
// MatSetValues(ADD) semantics (src/main.F90:157-166: every CSR value written once, no atomics, fixed order), but the
// phases of variant 12 (hex_sym2.cu: geometry | FMA | G' -> K | gather behind four CTA-wide barriers per layer, all of
// them loading the LSU pipe) become two groups of warps that run against each other through a ring of cell rows:
//
//   * 8 MMA warps, one per cell column of the tile.  A warp marches along y through its column.  For the next four
//     cells it computes the scaled physical gradients g~_b(ip) = adj(J)^T grad N_b sqrt(mu w / detJ) (lane = cell x
//     integration point, closed-form adjugate instead of dgetrf/dgetri) into a private scratch in the DMMA fragment
//     layout, then, cell by cell, 18 DMMA.8x8x4 accumulate over the 8 integration points six 8 x 8 node-pair tiles:
//     D_i = sum g^i_a g^i_b (i = 0..2) and O_ij = sum rho g^i_a g^j_b + g^j_a g^i_b = K^{ij}_ab (i < j, rho = lambda /
//     mu).  One 64-bit register per lane is both the A and the B fragment of a gradient component, so a cell costs six
//     operand registers and three 128-bit shared loads -- the FMA phase of variant 12 needed 48 operand wavefronts.
//     The tile rows are indexed by (z bit, x bit, PARITY of the node's y position): consecutive cells of the column
//     share their common face's rows, so node pairs on that face are summed across the two cells in the accumulators
//     themselves (no shared-memory carry in y); pairs on the top face are carried to the next cell layer through
//     72 doubles of shared memory per cell.  Complete pairs (36 of 64 per tile) go to the ring as
//     K^{ii} = (rho + 1) D_i + (D_0 + D_1 + D_2), K^{ij} = O_ij; K^{ji}_ab is read as K^{ij}_ba.
//   * 8 gather warps, one node at a time, lane <-> consecutive CSR value: the nine 27-value runs of a node (3 scalar
//     rows x dz in {-1, 0, +1}) are sums of at most TWO ring entries (the cells left and right of the node; y and z
//     are already summed), stored as coalesced 216-byte runs; f_r = sum_c K_rc u_c from the same registers.
//   * producer / consumer hand-over per cell row with mbarriers (full: 8 MMA warps, empty: 2 x 7 node items), node
//     layers (coordinates, u, row metadata, cell scale) prefetched two layers ahead with cp.async by the gather warps
//     and published through an mbarrier (cp.async.mbarrier.arrive.noinc).  No CTA-wide barrier inside the march.
// One CTA per SM (512 threads, ~216 KB shared memory) owns 7 x NYT node positions and marches through the node layers
// of a z-chunk.  DESIGN.md section 4.1.
#include <cstdio>

#include "elem_math.cuh"
#include "launch.h"

namespace femx {

namespace {

#ifndef FEMX_MMA_NYT
#define FEMX_MMA_NYT 13
#endif
// FEMX_MMA_PROF = 1 (diagnostic build, hex_mma_prof.cu = variant 21): lane 0 of every warp accumulates clock64 intervals per phase
#ifndef FEMX_MMA_PROF
#define FEMX_MMA_PROF 0
#endif
#ifndef FEMX_MMA_NGW
#define FEMX_MMA_NGW 8
#endif
#ifndef FEMX_MMA_KERNEL
#define FEMX_MMA_KERNEL hex_mma_kernel
#define FEMX_MMA_LAUNCH launch_hex_mma
#endif
constexpr int NXT = 7, EXC = NXT + 1;            // owned node columns, cell columns (= MMA warps)
constexpr int NYT = FEMX_MMA_NYT, EYC = NYT + 1; // owned node rows, cell rows
constexpr int UX = NXT + 2, UY = NYT + 2, UL = UX * UY;   // node positions of a layer incl. halo
constexpr int NP = NXT * NYT;
constexpr int NMW = 8, NGW = FEMX_MMA_NGW, NT = 32 * (NMW + NGW);   // MMA and gather warps
#ifndef FEMX_MMA_RING
#define FEMX_MMA_RING 5
#endif
constexpr int RING = FEMX_MMA_RING;             // cell rows in flight between the MMA and the gather warps (3 / 4 / 5: 1.333 / 1.327 / 1.297 ms; 6 do not fit)
constexpr int LR = 3, NR = 4;                    // node-layer ring, node-id ring
constexpr int TP = 36, CP = 6 * TP, SP = 1729;   // doubles per tile / cell / ring slot (annealed against bank conflicts, tools/hex_mma_anneal.py)
constexpr int GP = 192;                          // doubles per cell of scaled gradients [i][row][ip & 3][ip >> 2]
constexpr int ZT = 12, ZC = 6 * ZT;              // z-carry entries per tile / per cell
constexpr int XP = UX * 3 + 1;                   // doubles per row of the coordinate layer
#ifndef FEMX_MMA_PF_ROW
#define FEMX_MMA_PF_ROW 6
#endif
constexpr int PF_ROW = FEMX_MMA_PF_ROW;                        // gather warps prefetch the next node layer when they reach this node row
static_assert(EXC == NMW, "one MMA warp per cell column");
static_assert(sizeof(HexRowMeta) % 16 == 0, "bulk copies move multiples of 16 bytes");
#ifndef FEMX_MMA_REG_M
#define FEMX_MMA_REG_M 160
#define FEMX_MMA_REG_G 96
#endif
// setmaxnreg moves registers inside the CTA's pool only: what the gather warps release (launch allocation - REG_G each) must cover
// what the MMA warps take (REG_M - launch allocation each), or setmaxnreg.inc never returns; and setmaxnreg is a warpgroup
// instruction: the gather warps must come in groups of four (7 gather warps, one per node column, trap in the protocol's bounded
// waits).  Validated splits: 8 gather warps 160 / 96 (launch 128), 12 gather warps 128 / 72 or 136 / 64 (launch 96).
static_assert(NGW % 4 == 0 && NMW % 4 == 0, "setmaxnreg works on warpgroups of four warps");
static_assert(32 * (NMW * FEMX_MMA_REG_M + NGW * FEMX_MMA_REG_G) <= 65536, "register split exceeds the register file");
static_assert(NYT > PF_ROW + 1 && EYC > RING, "tile too short for the prefetch point / the ring");

// position of entry (row m, column n) of a tile inside its 36-entry block, by the parity P of the cell row's +y nodes
// (complete entries only: not both on the +y face, not both on the top face), permuted so that the gather's loads and
// the MMA warps' stores meet few bank conflicts
__constant__ unsigned char c_perm[2][TP] = {
    {21, 6, 14, 12, 27, 29, 25, 30, 33, 26, 7, 11, 23, 17, 18, 16, 3, 13, 20, 15, 34, 9, 0, 24, 8, 5, 22, 4, 35, 1, 10, 19, 2, 28, 31, 32},
    {17, 4, 21, 10, 7, 1, 20, 18, 6, 25, 12, 27, 32, 5, 15, 26, 23, 13, 0, 14, 2, 31, 8, 35, 34, 11, 16, 19, 33, 30, 9, 22, 24, 28, 3, 29}};

struct Smem {
  double ring[RING][SP];          // K entries of one cell row: [cell column][tile][36]
  alignas(16) double zc[EXC][EYC][ZC];        // top-face pair sums of the previous cell layer (raw accumulators)
  double gsc[NMW][4][GP];         // per MMA warp: scaled gradients of its next four cells
  double xlay[LR][UY * XP];
  double ulay[LR][UL][3];
  double esc[LR][EYC * EXC];
  alignas(16) HexRowMeta meta[LR][NP];   // rows arrive by bulk copy: 16-byte aligned
  int nid[NR][UL];
  double fcarry[NP][3];
  unsigned long long bar_full[RING], bar_empty[RING], bar_lay[LR];
};

__device__ __forceinline__ unsigned saddr(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void cp_async8z(void *smem, const void *gmem, bool ok) {   // 8 bytes, zero-filled when !ok
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" ::"r"(saddr(smem)), "l"(gmem), "r"(ok ? 8 : 0) : "memory");
}
__device__ __forceinline__ void cp_async4(void *smem, const void *gmem) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(saddr(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void mbar_init(unsigned long long *b, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(saddr(b)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned long long *b) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(saddr(b)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_cp_async(unsigned long long *b) {   // arrives when this thread's cp.async are done
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(saddr(b)) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *b, unsigned bytes) {   // bytes a bulk copy will deliver in this phase
  asm volatile("mbarrier.expect_tx.relaxed.cta.shared::cta.b64 [%0], %1;" ::"r"(saddr(b)), "r"(bytes) : "memory");
}
// TMA engine, 1-D: size and both addresses multiples of 16 bytes; completion is counted in bytes on the mbarrier
__device__ __forceinline__ void bulk_g2s(void *smem, const void *gmem, unsigned bytes, unsigned long long *b) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(saddr(smem)), "l"(gmem), "r"(bytes), "r"(saddr(b)) : "memory");
}
__device__ __forceinline__ bool mbar_try(unsigned long long *b, unsigned parity) {
  unsigned ok;
  asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
               : "=r"(ok) : "r"(saddr(b)), "r"(parity) : "memory");
  return ok != 0;
}
// a protocol error must not hang the GPU: give up after 2^22 failed polls (try_wait suspends the thread for a hardware-defined
// interval each time, so this is a fraction of a second) and trap.  Inline on purpose: a call in the wait path would spill
// the accumulators around it.
__device__ __forceinline__ void mbar_wait(unsigned long long *b, unsigned parity) {
  if (mbar_try(b, parity)) return;
  unsigned n = 0;
  while (!mbar_try(b, parity)) {
#ifndef FEMX_MMA_NS
#define FEMX_MMA_NS 40
#endif
    __nanosleep(FEMX_MMA_NS);            // polling goes through the shared-memory pipe the working warps need
    if (++n > (1u << 22)) __trap();
  }
}
__device__ __forceinline__ void dmma(double &c0, double &c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
__device__ __forceinline__ double lds_pred(unsigned addr, bool on) {
  double v;
  asm volatile("{\n .reg .pred p;\n setp.ne.s32 p, %2, 0;\n mov.f64 %0, 0d0000000000000000;\n @p ld.shared.f64 %0, [%1];\n}"
               : "=d"(v) : "r"(addr), "r"((int)on) : "memory");
  return v;
}

__host__ __device__ constexpr int sym_t(int i, int j) { return i == 0 ? j : (i == 1 ? j + 2 : 5); }   // i <= j: 00 01 02 11 12 22

__device__ __forceinline__ int emit_idx(int P, int m, int n) {
  const int s = n & 1, q = n >> 1;
  if (s != P) return q < 2 ? q * 8 + m : 16 + (q - 2) * 4 + m;          // column on the -y face: rows 0..7 (0..3 under a top column)
  const int r2 = m >> 1;                                                // column on the +y face: rows on the -y face only
  return q < 2 ? 24 + q * 4 + r2 : 32 + (q - 2) * 2 + r2;
}
__device__ __forceinline__ int emit_off(int P, int t, int m, int n) { return t * TP + c_perm[P][emit_idx(P, m, n)]; }

// value (dy, dx, j) of a 27-value run that gather lane l serves: lanes 0..8 hold the dx = 0 neighbours (two contributing
// cells; their second load touches one half-warp only), lanes 9..26 the dx = -1 / +1 neighbours (one cell)
__device__ __forceinline__ bool lane_value(int l, int &dy, int &dx, int &j) {
  if (l < 9) { dy = l / 3 - 1; dx = 0; j = l % 3; return true; }
  if (l < 27) { const int k = l - 9; dx = k < 9 ? -1 : 1; dy = (k / 3) % 3 - 1; j = k % 3; return true; }
  dy = 1; dx = 1; j = 2;
  return false;
}
// byte offset (relative to the ring entry block of cell column px of the lane's cell row) of contribution c of the value
// lane l serves in group g = dz + 1, scalar row i, for a node whose tile-local y position has parity par; 0xffff = none
__device__ __forceinline__ unsigned gather_off(int par, int g, int c, int i, int l) {
  int dy, dx, j;
  if (!lane_value(l, dy, dx, j)) return 0xffffu;
  const int bzr = g == 0, bzc = g == 2;
  int n = 0;
  for (int ox = 0; ox < 2; ++ox) {
    const int bxr = 1 - ox, bxc = bxr + dx;      // the node is the +x node of the cell on its left, the -x node of the one on its right
    if (bxc < 0 || bxc > 1) continue;
    if (n++ != c) continue;
    // dy = -1: cell row below the node (its +y nodes have the node's parity); dy = 0, +1: the cell row above
    const int P = dy < 0 ? par : par ^ 1, pc = par ^ (dy & 1);
    const int m = 4 * bzr + 2 * bxr + par, nn = 4 * bzc + 2 * bxc + pc;
    const int off = i <= j ? emit_off(P, sym_t(i, j), m, nn) : emit_off(P, sym_t(j, i), nn, m);
    return (unsigned)(ox * CP + off) * 8u;
  }
  return 0xffffu;
}

struct Tile {
  int i0, j0, l0, l1;
};

#if FEMX_MMA_PROF
// MMA warps: 0 wait layer, 1 geometry, 2 carry + z-in, 3 operands + DMMA, 4 wait ring slot, 5 emit + z-out + arrive
// gather warps: 8 wait layer, 9 prefetch, 10 wait cell row, 11 metadata + loads, 12 stores, 13 f, 14 arrive
__device__ unsigned long long g_prof[24];
#define PROF_DECL long long tprev = clock64(); long long pacc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#define PROF_MARK(i) { const long long now_ = clock64(); pacc[(i) & 7] += now_ - tprev; tprev = now_; }
#define PROF_FLUSH(base) if (lane == 0) { for (int i_ = 0; i_ < 8; ++i_) atomicAdd(&g_prof[(base) + i_], (unsigned long long)pacc[i_]); }
#else
#define PROF_DECL long long tprev = 0; long long pacc[1] = {0}; (void)tprev; (void)pacc;
#define PROF_MARK(i)
#define PROF_FLUSH(base)
#endif

// ------------------------------------------------------------------------------------------------ node-layer prefetch
// (issued by the 256 threads of the gather warps; gt = thread index among them)
__device__ __forceinline__ long long pos_of(const DevHex &hx, const Tile &T, int t, int lay, bool &ok) {
  const int ni = T.i0 - 1 + t % UX, nj = T.j0 - 1 + t / UX;
  ok = lay >= 0 && lay < hx.NK && ni >= 0 && nj >= 0 && ni < hx.NI && nj < hx.NJ;
  return ok ? ((long long)lay * hx.NJ + nj) * hx.NI + ni : 0;
}
__device__ __forceinline__ void fetch_nid(Smem &S, const DevHex &hx, const Tile &T, int rl, int gt) {
  const int lay = T.l0 - 1 + rl;
  for (int t = gt; t < UL; t += 32 * NGW) {
    bool ok;
    const long long c = pos_of(hx, T, t, lay, ok);
    if (ok) cp_async4(&S.nid[rl & (NR - 1)][t], hx.node_at + c);
    else S.nid[rl & (NR - 1)][t] = -1;   // read back by this thread only
  }
}
template <bool WITH_F>
__device__ __forceinline__ void fetch_layer(Smem &S, const DevHex &hx, const Tile &T, const double *u, int rl, int gt) {
  const int lay = T.l0 - 1 + rl, sl = rl % LR;
  for (int t = gt; t < UL; t += 32 * NGW) {
    bool ok;
    const long long c = pos_of(hx, T, t, lay, ok);
    double *xb = &S.xlay[sl][(t / UX) * XP + (t % UX) * 3];
#pragma unroll
    for (int d = 0; d < 3; ++d) cp_async8z(xb + d, hx.pos_x + c * 3 + d, ok);
    if (WITH_F) {
      const int r = S.nid[rl & (NR - 1)][t];
      const double *src = u + (long long)(r >= 0 ? r : 0) * 3;
#pragma unroll
      for (int d = 0; d < 3; ++d) cp_async8z(&S.ulay[sl][t][d], src + d, r >= 0);
    }
  }
  // row metadata (48 bytes per node, the tile's 7 nodes of a row are contiguous in pos_meta): one bulk copy by the TMA engine per
  // node row, issued by one thread per row and counted in bytes on the layer barrier; positions outside the grid are zero-filled
  // (nb = 0: no node here) with cp.async as before
  for (int row = gt; row < NYT; row += 32 * NGW) {
    const int nj = T.j0 + row;
    const int nvalid = (lay >= 0 && lay < hx.NK && nj < hx.NJ) ? max(0, min(NXT, hx.NI - T.i0)) : 0;
    HexRowMeta *dst = &S.meta[sl][row * NXT];
    if (nvalid > 0) {
      mbar_expect_tx(&S.bar_lay[sl], (unsigned)(nvalid * sizeof(HexRowMeta)));
      bulk_g2s(dst, hx.pos_meta + ((long long)lay * hx.NJ + nj) * hx.NI + T.i0, (unsigned)(nvalid * sizeof(HexRowMeta)), &S.bar_lay[sl]);
    }
    for (int q = nvalid; q < NXT; ++q) {
      double *rm = reinterpret_cast<double *>(dst + q);
#pragma unroll
      for (int d = 0; d < 6; ++d) cp_async8z(rm + d, hx.pos_meta, false);
    }
  }
  for (int t = gt; t < EYC * EXC; t += 32 * NGW) {
    const int ei = T.i0 - 1 + t % EXC, ej = T.j0 - 1 + t / EXC;
    const bool ok = lay >= 0 && lay < hx.NK - 1 && ei >= 0 && ej >= 0 && ei < hx.NI && ej < hx.NJ;
    cp_async8z(&S.esc[sl][t], hx.cellmu + (ok ? ((long long)lay * hx.NJ + ej) * hx.NI + ei : 0), ok);
  }
  mbar_arrive_cp_async(&S.bar_lay[sl]);
}

// ------------------------------------------------------------------------------------------------ MMA warps
// scaled physical gradients of cells (cx, cy0 .. cy0 + 3): lane = (cell cq = lane >> 3, integration point ip = lane & 7)
// (FormdNdetJ, src/m_local.F90:175-189).  Tile row of node b = 4 bz + 2 bx + parity of its y position.
__device__ __forceinline__ void geometry4(const double *xlo, const double *xhi, const double *esc, double *gw, int cx, int cy0, int lane) {
  const int cq = lane >> 3, ip = lane & 7, cy = cy0 + cq;
  if (cy >= EYC) return;
  const double smu = esc[cy * EXC + cx];
  if (smu == 0.0) return;
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
    const double *xs = (bz ? xhi : xlo) + (cy + by) * XP + (cx + bx) * 3;
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
  const double sc = smu * rsqrt(det);   // hex weights are 1 (src/m_elems.F90:508)
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) adj[i][j] *= sc;
  double *g = gw + cq * GP + (ip & 3) * 2 + (ip >> 2);
  const int pc = cq & 1;   // cy0 is a multiple of 4: parity of the cell's -y node row
#pragma unroll
  for (int b = 0; b < 8; ++b) {
    const int bx = (b == 1 || b == 2 || b == 5 || b == 6) ? 1 : 0, by = ((b & 3) >= 2) ? 1 : 0, bz = b >> 2;
    const double d0 = (bx ? 0.125 : -0.125) * fy[by] * fz[bz];
    const double d1 = (by ? 0.125 : -0.125) * fx[bx] * fz[bz];
    const double d2 = (bz ? 0.125 : -0.125) * fx[bx] * fy[by];
    const int row = 4 * bz + 2 * bx + (by ^ pc);
#pragma unroll
    for (int i = 0; i < 3; ++i) g[(i * 8 + row) * 8] = adj[i][0] * d0 + adj[i][1] * d1 + adj[i][2] * d2;
  }
}

struct MmaLane {           // per-lane constants of an MMA warp
  int m, q;                // tile row, column pair of the lane's two accumulators per tile
  int offA[2], offB[2];    // ring offsets (doubles, within a tile) of the -y-column / +y-column accumulator, by row parity P
};

// one cell of an MMA warp's column.  P = parity of the cell's +y nodes (cy + 1) & 1; acc[t][s]: tile t, column 2 q + s.
template <int P>
__device__ __forceinline__ void mma_cell(Smem &S, double (&acc)[6][2], const MmaLane &L, const double *gcell, int cx, unsigned rc, int cy,
                                         bool on, bool zin, double rho, int lane, long long &tprev, long long *pacc, int dbg) {
  const bool rowmy = (L.m & 1) != P;                          // row node on the -y face of this cell
  const bool keep = cy > 0 && (L.m & 1) == 1 - P;             // pairs on the face shared with the previous cell stay in the accumulators
  double *zcell = &S.zc[cx][cy][0];
#pragma unroll
  for (int t = 0; t < 6; ++t) {
    acc[t][P] = 0.0;
    if (!keep) acc[t][1 - P] = 0.0;
  }
  if (zin && L.m < 4 && L.q < 2 && !(dbg & 4)) {              // bottom x bottom pairs start from the top x top sums of the cell below
    const double2 *zA = reinterpret_cast<const double2 *>(zcell + (L.m * 2 + L.q) * 6), *zB = reinterpret_cast<const double2 *>(zcell + (8 + (L.m >> 1) * 2 + L.q) * 6);
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const double2 a = zA[k];
      acc[2 * k][1 - P] += a.x; acc[2 * k + 1][1 - P] += a.y;
      if (rowmy) {
        const double2 b = zB[k];
        acc[2 * k][P] += b.x; acc[2 * k + 1][P] += b.y;
      }
    }
  }
  PROF_MARK(2)
  if (on && !(dbg & 2)) {
    double g[3][2], rg[2][2];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const double2 w = *reinterpret_cast<const double2 *>(gcell + i * 64 + 2 * lane);
      g[i][0] = w.x; g[i][1] = w.y;
    }
#pragma unroll
    for (int s = 0; s < 2; ++s) { rg[0][s] = rho * g[0][s]; rg[1][s] = rho * g[1][s]; }
#pragma unroll
    for (int s = 0; s < 2; ++s) {
      dmma(acc[0][0], acc[0][1], g[0][s], g[0][s]);
      dmma(acc[3][0], acc[3][1], g[1][s], g[1][s]);
      dmma(acc[5][0], acc[5][1], g[2][s], g[2][s]);
      dmma(acc[1][0], acc[1][1], rg[0][s], g[1][s]);
      dmma(acc[2][0], acc[2][1], rg[0][s], g[2][s]);
      dmma(acc[4][0], acc[4][1], rg[1][s], g[2][s]);
      dmma(acc[1][0], acc[1][1], g[1][s], g[0][s]);
      dmma(acc[2][0], acc[2][1], g[2][s], g[0][s]);
      dmma(acc[4][0], acc[4][1], g[2][s], g[1][s]);
    }
  }
  const bool toptop = L.m >= 4 && L.q >= 2;
#if FEMX_MMA_PROF
  if (acc[0][0] + acc[5][1] == 1.2345e300) tprev = 0;   // wait for the accumulators before reading the clock
#endif
  PROF_MARK(3)
  const unsigned slot = rc % RING;
  if (rc >= RING) mbar_wait(&S.bar_empty[slot], (rc / RING - 1) & 1);
  PROF_MARK(4)
  if (!toptop && !(dbg & 8)) {
    // K^{ii} = (rho + 1) D_i + tr, K^{ij} = O_ij  (elem_math.cuh block form of B_a^T D B_b)
    const double rho1 = rho + 1.0;
    double *rcell = &S.ring[slot][cx * CP];
    {
      const double tr = acc[0][1 - P] + acc[3][1 - P] + acc[5][1 - P];
      double *r = rcell + L.offA[P];
      r[0 * TP] = fma(rho1, acc[0][1 - P], tr); r[1 * TP] = acc[1][1 - P]; r[2 * TP] = acc[2][1 - P];
      r[3 * TP] = fma(rho1, acc[3][1 - P], tr); r[4 * TP] = acc[4][1 - P]; r[5 * TP] = fma(rho1, acc[5][1 - P], tr);
    }
    if (rowmy) {
      const double tr = acc[0][P] + acc[3][P] + acc[5][P];
      double *r = rcell + L.offB[P];
      r[0 * TP] = fma(rho1, acc[0][P], tr); r[1 * TP] = acc[1][P]; r[2 * TP] = acc[2][P];
      r[3 * TP] = fma(rho1, acc[3][P], tr); r[4 * TP] = acc[4][P]; r[5 * TP] = fma(rho1, acc[5][P], tr);
    }
  } else if (toptop && !(dbg & 4)) {                          // top x top pairs wait for the cell above
    double2 *zA = reinterpret_cast<double2 *>(zcell + ((L.m - 4) * 2 + (L.q - 2)) * 6), *zB = reinterpret_cast<double2 *>(zcell + (8 + ((L.m - 4) >> 1) * 2 + (L.q - 2)) * 6);
#pragma unroll
    for (int t = 0; t < 6; t += 2) {
      zA[t >> 1] = make_double2(acc[t][1 - P], acc[t + 1][1 - P]);
      if (rowmy) zB[t >> 1] = make_double2(acc[t][P], acc[t + 1][P]);
    }
  }
  __syncwarp();
  if (lane == 0) mbar_arrive(&S.bar_full[slot]);
  PROF_MARK(5)
}

}  // namespace

template <bool WITH_F>
__global__ void __launch_bounds__(NT, 1)
FEMX_MMA_KERNEL(DevHex hx, const double *__restrict__ u, double *__restrict__ vals, double *__restrict__ fint, int LZ, int dbg_in,
                const int *__restrict__ cta_list) {
#if FEMX_MMA_PROF
  const int dbg = dbg_in;   // diagnostic: bit 0 skip geometry, 1 DMMA, 2 z-carry, 3 ring stores, 4 gather loads, 5 value stores, 6 f
#else
  constexpr int dbg = 0;
#endif
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Smem &S = *reinterpret_cast<Smem *>(smem_raw);
  const int tid = threadIdx.x, lane = tid & 31;
  const int wu = __shfl_sync(0xffffffffu, tid >> 5, 0);   // warp index in a uniform register
  Tile T;
  int bx = blockIdx.x, by = blockIdx.y, bz = blockIdx.z;
  if (cta_list) {   // a subset of the grid (peer exchange: interface tiles first, on their own stream): linear ids, x fastest
    const int gx = (hx.NI + NXT - 1) / NXT, gy = (hx.NJ + NYT - 1) / NYT, id = cta_list[blockIdx.x];
    bx = id % gx; by = (id / gx) % gy; bz = id / (gx * gy);
  }
  T.i0 = bx * NXT; T.j0 = by * NYT;
  T.l0 = bz * LZ; T.l1 = min(T.l0 + LZ, hx.NK) - 1;
  const int nsteps = T.l1 - T.l0 + 2;   // cell layers l0-1 .. l1; node layers rl = 0 .. nsteps (layer l0 - 1 + rl)
  const double rho = hx.rho;

  if (tid == 0) {
    for (int r = 0; r < RING; ++r) { mbar_init(&S.bar_full[r], NMW); mbar_init(&S.bar_empty[r], 2 * NXT); }
    for (int r = 0; r < LR; ++r) mbar_init(&S.bar_lay[r], 32 * NGW);
  }
  __syncthreads();

  if (wu >= NMW) {
    // ============================================================================================ gather warps
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(FEMX_MMA_REG_G));      // the MMA warps take the registers (geometry + accumulators)      // the MMA warps take the registers (geometry + accumulators + pending stores)
    const int gw = wu - NMW, gt = tid - 32 * NMW;
    if (WITH_F && lane < 3)
      for (int item = gw; item < NP; item += NGW) S.fcarry[item][lane] = 0.0;   // read back by this warp only
    // prologue: node ids of layers 0..2, then layers 0 and 1
    fetch_nid(S, hx, T, 0, gt); fetch_nid(S, hx, T, 1, gt); fetch_nid(S, hx, T, 2, gt);
    cp_async_commit();
    cp_async_wait<0>();
    fetch_layer<WITH_F>(S, hx, T, u, 0, gt);
    fetch_layer<WITH_F>(S, hx, T, u, 1, gt);
    cp_async_commit();
    // lane constants
    int dy, dx, jj;
    const bool lact = lane_value(lane, dy, dx, jj);
    const int kk = (dy + 1) * 3 + dx + 1;
    const bool two = lact && dx == 0;
    const int uoff = ((dy + 1) * UX + dx + 1) * 3 + jj;
    unsigned tab[2][9];   // [parity][(g, c, i) pairs]: two 16-bit byte offsets per register
#pragma unroll
    for (int par = 0; par < 2; ++par)
#pragma unroll
      for (int e = 0; e < 9; ++e) {
        const int e0 = 2 * e, e1 = 2 * e + 1;
        tab[par][e] = gather_off(par, e0 / 6, (e0 / 3) & 1, e0 % 3, lane) | gather_off(par, e1 / 6, (e1 / 3) & 1, e1 % 3, lane) << 16;
      }
    const bool upper = (lane & 16) != 0;
    PROF_DECL
    for (int it = 0; it < nsteps; ++it) {
      const int m = T.l0 - 1 + it;
      const bool lay_lo = m >= T.l0 && m <= T.l1, lay_hi = m + 1 >= T.l0 && m + 1 <= T.l1;
      mbar_wait(&S.bar_lay[(it + 1) % LR], ((it + 1) / LR) & 1);
      if (it == 0) mbar_wait(&S.bar_lay[0], 0);
      PROF_MARK(8)
      bool prefetched = false;
      for (int item = gw; item < NP; item += NGW) {
        const int jn = item / NXT, px = item - jn * NXT;
        if (!prefetched && jn >= PF_ROW) {
          // every gather warp is past node row PF_ROW - 1 of this step only when the whole previous step has been consumed
          // (ring depth): the slots of node layer it - 1 are free
          prefetched = true;
          if (it + 3 <= nsteps) fetch_nid(S, hx, T, it + 3, gt);
          cp_async_commit();
          cp_async_wait<1>();     // this thread's node ids of layer it + 2 (issued one step ago) have landed
          if (it + 2 <= nsteps) fetch_layer<WITH_F>(S, hx, T, u, it + 2, gt);
          cp_async_commit();
          PROF_MARK(9)
        }
        const int par = (jn + 1) & 1;
        const unsigned rc = (unsigned)it * EYC + jn + 1;      // sequence number of the cell row above the node row
        const unsigned slotB = rc % RING, slotA = (rc - 1) % RING;
        mbar_wait(&S.bar_full[slotB], (rc / RING) & 1);
        PROF_MARK(10)
        const unsigned eb = saddr(&S.ring[dy < 0 ? slotA : slotB][px * CP]);
        double ulo = 0.0, uhi = 0.0;
        if (WITH_F) {
          ulo = (&S.ulay[it % LR][0][0])[(jn * UX + px) * 3 + uoff];
          uhi = (&S.ulay[(it + 1) % LR][0][0])[(jn * UX + px) * 3 + uoff];
        }
        const HexRowMeta *rlo = &S.meta[it % LR][item], *rhi = &S.meta[(it + 1) % LR][item];
        const bool act_lo = lay_lo && rlo->nb > 0, act_hi = lay_hi && rhi->nb > 0;
        // group dz = 0 / +1 of the lower node: columns on layers m / m+1; group dz = -1 of the upper node: columns on layer m
        const int k0 = rhi->k[kk], k1 = rlo->k[9 + kk], k2 = rlo->k[18 + kk];
        const bool st = !(dbg & 32);
        const bool on0 = st && act_hi && lact && k0 != 255, on1 = st && act_lo && lact && k1 != 255, on2 = st && act_lo && lact && k2 != 255;
        const unsigned q0 = (unsigned)rhi->base + 3u * k0 + jj, q1 = (unsigned)rlo->base + 3u * k1 + jj, q2 = (unsigned)rlo->base + 3u * k2 + jj;
        const int rs_lo = 3 * rlo->nb, rs_hi = 3 * rhi->nb;
        double v[9];   // v[i * 3 + g]
#pragma unroll
        for (int g = 0; g < 3; ++g)
#pragma unroll
          for (int i = 0; i < 3; ++i) {
            const int e0 = (g * 2 + 0) * 3 + i, e1 = (g * 2 + 1) * 3 + i;
            const unsigned w0 = par ? tab[1][e0 >> 1] : tab[0][e0 >> 1], w1 = par ? tab[1][e1 >> 1] : tab[0][e1 >> 1];
            const unsigned o0 = (e0 & 1) ? w0 >> 16 : w0 & 0xffffu, o1 = (e1 & 1) ? w1 >> 16 : w1 & 0xffffu;
            v[i * 3 + g] = lds_pred(eb + o0, lact && !(dbg & 16)) + lds_pred(eb + o1, two && !(dbg & 16));
          }
#if FEMX_MMA_PROF
        if (v[0] + v[4] + v[8] == 1.2345e300) tprev = 0;
#endif
        PROF_MARK(11)
#pragma unroll
        for (int i = 0; i < 3; ++i) {
          if (on0) vals[q0 + (unsigned)(i * rs_hi)] = v[i * 3 + 0];
          if (on1) vals[q1 + (unsigned)(i * rs_lo)] = v[i * 3 + 1];
          if (on2) vals[q2 + (unsigned)(i * rs_lo)] = v[i * 3 + 2];
        }
        PROF_MARK(12)
        if (WITH_F && !(dbg & 64)) {
          // f_r = sum_c K_rc u_c: lane products, butterfly-summed (fixed order); lanes 0-15 end up with the three sums of
          // the lower node, lanes 16-31 with those of the upper node
          double keep[3];
#pragma unroll
          for (int i = 0; i < 3; ++i) {
            const double flo = fma(v[i * 3 + 1], ulo, v[i * 3 + 2] * uhi), fhi = v[i * 3 + 0] * ulo;
            keep[i] = (upper ? fhi : flo) + __shfl_xor_sync(0xffffffffu, upper ? flo : fhi, 16);
          }
          const bool b3 = (lane & 8) != 0, b2 = (lane & 4) != 0;
          const double r0 = __shfl_xor_sync(0xffffffffu, b3 ? keep[0] : keep[2], 8);
          const double r1 = __shfl_xor_sync(0xffffffffu, keep[1], 8);
          const double x0 = (b3 ? keep[2] : keep[0]) + r0, x1 = b3 ? 0.0 : keep[1] + r1;
          double mine = (b2 ? x1 : x0) + __shfl_xor_sync(0xffffffffu, b2 ? x0 : x1, 4);
          mine += __shfl_xor_sync(0xffffffffu, mine, 2);
          mine += __shfl_xor_sync(0xffffffffu, mine, 1);
          const int li = (lane & 15) >> 2;
          const bool wr = (lane & 3) == 0 && li < 3;
          // f of the lower node = (dz = -1 part carried from the previous step) + (dz = 0, +1 parts)
          if (wr && !upper && act_lo) fint[(long long)rlo->grow + li] = S.fcarry[item][li] + mine;
          __syncwarp();
          if (wr && upper) S.fcarry[item][li] = act_hi ? mine : 0.0;
        }
        PROF_MARK(13)
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(&S.bar_empty[slotA]);
          mbar_arrive(&S.bar_empty[slotB]);
          if (jn == 0) mbar_arrive(&S.bar_empty[slotA]);          // stands in for the node row below the tile
          if (jn == NYT - 1) mbar_arrive(&S.bar_empty[slotB]);    // ... and the one above
        }
        PROF_MARK(14)
      }
    }
    cp_async_wait<0>();
    PROF_FLUSH(8)
  } else {
    // ============================================================================================ MMA warps
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(FEMX_MMA_REG_M));
    const int cx = wu;
    MmaLane L;
    L.m = lane >> 2; L.q = lane & 3;
#pragma unroll
    for (int P = 0; P < 2; ++P) {
      const bool tt = L.m >= 4 && L.q >= 2;
      L.offA[P] = tt ? 0 : c_perm[P][emit_idx(P, L.m, 2 * L.q + 1 - P)];
      L.offB[P] = (tt || (L.m & 1) == P) ? 0 : c_perm[P][emit_idx(P, L.m, 2 * L.q + P)];
    }
    double acc[6][2];
#pragma unroll
    for (int t = 0; t < 6; ++t) acc[t][0] = acc[t][1] = 0.0;
    double *gw = &S.gsc[wu][0][0];
    PROF_DECL
    for (int it = 0; it < nsteps; ++it) {
      mbar_wait(&S.bar_lay[(it + 1) % LR], ((it + 1) / LR) & 1);
      if (it == 0) mbar_wait(&S.bar_lay[0], 0);
      PROF_MARK(0)
      const double *xlo = &S.xlay[it % LR][0], *xhi = &S.xlay[(it + 1) % LR][0], *esc = &S.esc[it % LR][0];
      for (int cy0 = 0; cy0 < EYC; cy0 += 4) {
        __syncwarp();
        if (!(dbg & 1)) geometry4(xlo, xhi, esc, gw, cx, cy0, lane);
        __syncwarp();
        PROF_MARK(1)
#pragma unroll
        for (int cq = 0; cq < 4; ++cq) {
          const int cy = cy0 + cq;
          if (cy < EYC) {
            const unsigned rc = (unsigned)it * EYC + cy;
            const bool on = esc[cy * EXC + cx] != 0.0;
            // cy0 is a multiple of 4: the parity of the cell's +y nodes is that of cq + 1
            if (cq & 1) mma_cell<0>(S, acc, L, gw + cq * GP, cx, rc, cy, on, it > 0, rho, lane, tprev, pacc, dbg);
            else mma_cell<1>(S, acc, L, gw + cq * GP, cx, rc, cy, on, it > 0, rho, lane, tprev, pacc, dbg);
          }
        }
      }
    }
    PROF_FLUSH(0)
  }
}

// z-chunk length: fewest (waves x steps per chunk); every chunk pays one extra cell layer
static int mma_lz(const femx_ctx *c, int nsm) {
  const DevHex &hx = c->dhex;
  const int tiles = ((hx.NI + NXT - 1) / NXT) * ((hx.NJ + NYT - 1) / NYT);
  int LZ = c->opts.reserved[1] > 0 ? c->opts.reserved[1] : 0;
  if (LZ <= 0) {
    double best = 1e300;
    for (int nch = 1; nch <= hx.NK; ++nch) {
      const int lz = (hx.NK + nch - 1) / nch, chunks = (hx.NK + lz - 1) / lz;
      const long long ctas = (long long)tiles * chunks;
      const double waves = (double)((ctas + nsm - 1) / nsm);
      const double cost = waves * (lz + 1.5);
      if (cost < best) { best = cost; LZ = lz; }
    }
  }
  return LZ > hx.NK ? hx.NK : LZ;
}

#if !FEMX_MMA_PROF
// grid of the tile kernel: tiles in x, tiles in y, z-chunks, node columns / rows per tile, layers per chunk
void hex_mma_grid(const femx_ctx *c, int *gx, int *gy, int *gz, int *nxt, int *nyt, int *lz) {
  int nsm = 148;
  cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, c->device);
  const DevHex &hx = c->dhex;
  *lz = mma_lz(c, nsm);
  *gx = (hx.NI + NXT - 1) / NXT; *gy = (hx.NJ + NYT - 1) / NYT; *gz = (hx.NK + *lz - 1) / *lz;
  *nxt = NXT; *nyt = NYT;
}
// n CTAs of the grid, named by d_list (linear ids), on stream st
int launch_hex_mma_part(femx_ctx *c, const double *d_u, double *d_vals, double *d_fint, const int *d_list, int n, cudaStream_t st) {
  int nsm = 148;
  cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, c->device);
  const int LZ = mma_lz(c, nsm);
  const size_t smem = sizeof(Smem);
  auto go = [&](auto kern) {
    cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    kern<<<n, NT, smem, st>>>(c->dhex, d_u, d_vals, d_fint, LZ, c->opts.reserved[2], d_list);
  };
  if (d_u) go(FEMX_MMA_KERNEL<true>); else go(FEMX_MMA_KERNEL<false>);
  return 1;
}
#endif

int FEMX_MMA_LAUNCH(femx_ctx *c, const double *d_u, double *d_vals, double *d_fint) {
  const DevHex &hx = c->dhex;
  int nsm = 148;
  cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, c->device);
  const int LZ = mma_lz(c, nsm);
  dim3 grid((hx.NI + NXT - 1) / NXT, (hx.NJ + NYT - 1) / NYT, (hx.NK + LZ - 1) / LZ);
  const size_t smem = sizeof(Smem);
  static_assert(sizeof(Smem) <= 232448, "shared memory exceeds what a CTA may opt into");
  auto go = [&](auto kern) {
    cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    kern<<<grid, NT, smem, c->stream>>>(hx, d_u, d_vals, d_fint, LZ, c->opts.reserved[2], nullptr);
  };
  if (d_u) go(FEMX_MMA_KERNEL<true>); else go(FEMX_MMA_KERNEL<false>);
#if FEMX_MMA_PROF
  {
    static const char *nm[24] = {"M wait layer", "M geometry", "M carry + z-in", "M operands + DMMA", "M wait slot", "M emit + z-out", "-", "-",
                                 "G wait layer", "G prefetch", "G wait row", "G meta + loads", "G stores", "G f", "G arrive", "-",
                                 "-", "-", "-", "-", "-", "-", "-", "-"};
    unsigned long long h[24], z[24] = {};
    cudaStreamSynchronize(c->stream);
    cudaMemcpyFromSymbol(h, g_prof, sizeof(h));
    cudaMemcpyToSymbol(g_prof, z, sizeof(z));
    const double rows = (double)grid.x * grid.y * grid.z * (LZ + 1) * EYC;
    fprintf(stderr, "femx hex_mma prof (cycles per warp per cell row; %.0f cell rows, LZ %d):\n ", rows, LZ);
    for (int i = 0; i < 24; ++i)
      if (nm[i][0] != '-') fprintf(stderr, " %s %.0f |", nm[i], (double)h[i] / rows / (i < 16 ? 8.0 : 4.0));
    fprintf(stderr, "\n");
  }
#endif
  return 1;
}

}  // namespace femx
