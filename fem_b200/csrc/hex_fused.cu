// hex_fused.cu -- assembly variant 2: fused, gather-ordered hex8 K (+ f_int) kernel.
//
// One CTA owns a TX x TY column of logical node positions and marches through node layers.
// Per step one ELEMENT layer is resident in shared memory as scaled physical gradients
//   gt_b(ip) = adj(J)^T dN_b * sqrt(mu w / detJ)            [8 ip][3][8 nodes] fp64 per element
// so that  sum_ip gt_a gt_b^T = mu * sum_ip w detJ (grad N_a)(grad N_b)^T = G'_ab  and
//   K_ab = rho G'_ab + G'_ab^T + tr(G'_ab) I,   rho = lambda/mu      (== B_a^T D B_b, FormElKE
// src/m_local.F90:61-65 with D of src/m_local.F90:214-228).
// Four threads serve each node position; every thread sums 16 (corner element, node b)
// contributions into 9 register-resident 3x3 blocks (no atomics, fixed order):
//   role 0  upper node (layer m+1), nodes b of the bottom face  -> its dz=-1 blocks, stored now
//   role 3  lower node (layer m),   nodes b of the top face     -> its dz=+1 blocks, stored now
//   roles 1/2 alternate: "start" the dz=0 blocks of the upper node with this layer's elements,
//            and next step "finish" them with the layer above, then store.
// Every CSR value is written exactly once; rows leave through a per-warp shared-memory
// transpose so that global stores are contiguous 216-byte runs.
#include "elem_math.cuh"
#include "launch.h"

namespace femx {

namespace {

constexpr int kGeoStride = 194;   // doubles per element: 8*24 + 2 pad -> 16B-unit stride 97 (odd)
constexpr int kDumpStride = 82;   // doubles per storing thread: 81 + 1 pad -> 16B-unit stride 41 (odd)

using RowMeta = HexRowMeta;   // per node position: where its row lives in the CSR value array

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

template <int TX, int TY, bool WITH_F>
__global__ void __launch_bounds__(4 * TX * TY, (TX * TY <= 16) ? 4 : ((TX * TY <= 32) ? 2 : 1))
hex_fused_kernel(DevMesh dm, DevHex hx, const double *__restrict__ u, double *__restrict__ vals,
                 double *__restrict__ fint, int LZ, int dbg) {
  constexpr int NP = TX * TY, NT = 4 * NP, NW = NT / 32;
  constexpr int EX = TX + 1, EY = TY + 1, EL = EX * EY;
  constexpr int UX = TX + 2, UY = TY + 2, UL = UX * UY;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  constexpr int GEO = (EL * kGeoStride > NP * 3 * kDumpStride) ? EL * kGeoStride : NP * 3 * kDumpStride;
  double *geo = reinterpret_cast<double *>(smem_raw);                    // [EL][kGeoStride]; after the FMA
                                                                         // loop reused as [NP*3][kDumpStride]
  // node-layer cache, 3 layers deep (two in use, one being prefetched with cp.async)
  double *xlay = geo + GEO;                                              // [3][UL][3] node coordinates
  double *ulay = xlay + 3 * UL * 3;                                      // [3][UL][3] displacements (WITH_F)
  double *fcarry = ulay + (WITH_F ? 3 * UL * 3 : 0);                     // [NP][3]              (WITH_F)
  double *esc = fcarry + (WITH_F ? NP * 3 : 0);                          // [2][EL] sqrt(mu) of the cell, 0 = absent
  RowMeta *rowmeta = reinterpret_cast<RowMeta *>(esc + 2 * EL);          // [3][NP]
  int *nid = reinterpret_cast<int *>(rowmeta + 3 * NP);                  // [4][UL] node id, -1 = absent

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int p = tid >> 2, role = tid & 3;
  const int px = p % TX, py = p / TX;
  const int i0 = blockIdx.x * TX, j0 = blockIdx.y * TY;
  const int l0 = blockIdx.z * LZ, l1 = min(l0 + LZ, hx.NK) - 1;
  const double rho = hx.rho;
  int lane_i[3], lane_s[3], lane_j[3];     // value v = it*32 + lane of an 81-value dump -> (row i, slot, j)
#pragma unroll
  for (int it = 0; it < 3; ++it) {
    const int v = min(it * 32 + lane, 80);
    lane_i[it] = v / 27; lane_s[it] = (v % 27) / 3; lane_j[it] = v % 3;
  }

  double acc[9][9];
#pragma unroll
  for (int s = 0; s < 9; ++s)
#pragma unroll
    for (int q = 0; q < 9; ++q) acc[s][q] = 0.0;

  // ---- asynchronous loaders (global -> shared without registers); a layer is identified by lay,
  //      ring slots are (lay+3)%3 for data and (lay+4)&3 for node ids
  auto cell_of = [&](int t, int lay, bool &ok) -> long long {
    const int ni = i0 - 1 + t % UX, nj = j0 - 1 + t / UX;
    ok = lay >= 0 && lay < hx.NK && ni >= 0 && nj >= 0 && ni < hx.NI && nj < hx.NJ;
    return ((long long)lay * hx.NJ + nj) * hx.NI + ni;
  };
  auto fetch_nid = [&](int lay) {
    int *dst = nid + ((lay + 4) & 3) * UL;
    for (int t = tid; t < UL; t += NT) {
      bool ok;
      const long long c = cell_of(t, lay, ok);
      if (ok) cp_async4(dst + t, hx.node_at + c);
      else dst[t] = -1;
    }
  };
  auto fetch_layer = [&](int lay) {   // needs nid of that layer resident (WITH_F)
    const int buf = (lay + 3) % 3;
    for (int t = tid; t < UL; t += NT) {
      bool ok;
      const long long c = cell_of(t, lay, ok);
      double *xb = xlay + (buf * UL + t) * 3;
      if (ok) {
#pragma unroll
        for (int d = 0; d < 3; ++d) cp_async8(xb + d, hx.pos_x + c * 3 + d);
      } else {
        xb[0] = 0.0; xb[1] = 0.0; xb[2] = 0.0;
      }
      if (WITH_F) {
        double *ub = ulay + (buf * UL + t) * 3;
        const int r = nid[((lay + 4) & 3) * UL + t];
        if (r >= 0) {
#pragma unroll
          for (int d = 0; d < 3; ++d) cp_async8(ub + d, u + (long long)r * 3 + d);
        } else {
          ub[0] = 0.0; ub[1] = 0.0; ub[2] = 0.0;
        }
      }
      const int ux = t % UX, uy = t / UX;
      if (ux >= 1 && ux <= TX && uy >= 1 && uy <= TY) {
        RowMeta *rm = rowmeta + buf * NP + (uy - 1) * TX + (ux - 1);
        if (ok) {
          const double *src = reinterpret_cast<const double *>(hx.pos_meta + c);
#pragma unroll
          for (int d = 0; d < 6; ++d) cp_async8(reinterpret_cast<double *>(rm) + d, src + d);
        } else {
          rm->nb = 0;
        }
      }
    }
  };
  auto fetch_esc = [&](int lay) {     // element layer
    double *dst = esc + (lay & 1) * EL;
    for (int t = tid; t < EL; t += NT) {
      const int ei = i0 - 1 + t % EX, ej = j0 - 1 + t / EX;
      if (lay >= 0 && lay < hx.NK - 1 && ei >= 0 && ej >= 0 && ei < hx.NI && ej < hx.NJ)
        cp_async8(dst + t, hx.cellmu + ((long long)lay * hx.NJ + ej) * hx.NI + ei);
      else
        dst[t] = 0.0;
    }
  };
  if (WITH_F && tid < NP * 3) fcarry[tid] = 0.0;
  fetch_nid(l0 - 1); fetch_nid(l0); fetch_nid(l0 + 1);
  fetch_esc(l0 - 1);
  cp_async_wait_all();
  __syncthreads();
  fetch_layer(l0 - 1); fetch_layer(l0);
  cp_async_wait_all();
  __syncthreads();

  for (int m = l0 - 1; m <= l1; ++m) {
    // ------------------------------------------------------------ phase A: geometry of element layer m
    // prefetch what the NEXT steps need; it lands while this step computes
    if (!(dbg & 8)) {
      fetch_nid(m + 3);
      fetch_layer(m + 2);
      fetch_esc(m + 1);
    }
    const double *escm = esc + (m & 1) * EL;
    {
      const double *xlo = xlay + ((m + 3) % 3) * UL * 3, *xhi = xlay + ((m + 4) % 3) * UL * 3;
      for (int item = tid; item < EL * 8; item += NT) {
        if ((dbg & 2) && m > l0) break;
        const int es = item % EL, ip = item / EL;
        const double smu = escm[es];
        if (smu == 0.0) continue;
        const int ex = es % EX, ey = es / EX;
        double X[8][3];
#pragma unroll
        for (int b = 0; b < 8; ++b) {
          const int bx = (b == 1 || b == 2 || b == 5 || b == 6) ? 1 : 0, by = ((b & 3) >= 2) ? 1 : 0;
          const double *xs = (b >= 4 ? xhi : xlo) + ((ey + by) * UX + (ex + bx)) * 3;
          X[b][0] = xs[0]; X[b][1] = xs[1]; X[b][2] = xs[2];
        }
        double dN[3][8];
        parent_grad<FEMX_HEX>(ip, dN);
        double J[3][3];
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
          for (int j = 0; j < 3; ++j) {
            double s = 0.0;
#pragma unroll
            for (int a = 0; a < 8; ++a) s += dN[i][a] * X[a][j];
            J[i][j] = s;
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
        double *g = geo + es * kGeoStride + ip * 24;
#pragma unroll
        for (int i = 0; i < 3; ++i) {
          double row[8];
#pragma unroll
          for (int b = 0; b < 8; ++b) row[b] = adj[i][0] * dN[0][b] + adj[i][1] * dN[1][b] + adj[i][2] * dN[2][b];
#pragma unroll
          for (int b = 0; b < 8; b += 2) *reinterpret_cast<double2 *>(g + i * 8 + b) = make_double2(row[b], row[b + 1]);
        }
      }
    }
    __syncthreads();

    // ------------------------------------------------------------ phase B: 16 contributions per thread
    const int par = (m + 2) & 1;
    const bool isA = (role == 1 + par), isB = (role == 2 - par);
    const bool upper = role == 0 || isB;                 // works on node layer m+1
    const int lay = upper ? m + 1 : m;
    const RowMeta *rm = rowmeta + ((lay + 3) % 3) * NP + p;
    const bool active = rm->nb > 0 && lay >= l0 && lay <= l1;   // node exists and its layer belongs to this chunk
    const int atop = upper ? 4 : 0;                      // our node sits on the top face of layer-m elements
    const int btop = (role == 3 || isB) ? 4 : 0;
    if (!isA) {
#pragma unroll
      for (int s = 0; s < 9; ++s)
#pragma unroll
        for (int q = 0; q < 9; ++q) acc[s][q] = 0.0;
    }
    if (active && !(dbg & 4)) {
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
    // ------------------------------------------------------------ store the groups completed by this step
    const bool writer = active && !isB;
    const int grp = role == 0 ? 0 : (role == 3 ? 2 : 1);       // dz = -1, 0, +1
    const unsigned wmask = __ballot_sync(0xffffffffu, writer);
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
        // partial f = sum over this group's 9 neighbours K_rc u_c; neighbour (dx,dy) of plane lay+dz
        const int nl = lay + grp - 1;
        const double *ub = ulay + ((nl + 3) % 3) * UL * 3;
#pragma unroll
        for (int s = 0; s < 9; ++s) {
          const double *uc = ub + ((py + s / 3) * UX + (px + s % 3)) * 3;
          const double u0 = uc[0], u1 = uc[1], u2 = uc[2];
#pragma unroll
          for (int i = 0; i < 3; ++i) fpart[i] += acc[s][i * 3] * u0 + acc[s][i * 3 + 1] * u1 + acc[s][i * 3 + 2] * u2;
        }
      }
    }
    if (WITH_F) {
      // f of the lower node = carry(dz=-1, previous step) + dz=0 (finisher) + dz=+1 (role 3)
      const int src3 = (lane & ~3) + 3;
      double f3[3];
#pragma unroll
      for (int i = 0; i < 3; ++i) f3[i] = __shfl_sync(0xffffffffu, fpart[i], src3);
      if (isA && active) {
#pragma unroll
        for (int i = 0; i < 3; ++i) fint[(long long)rm->grow + i] = (fcarry[p * 3 + i] + fpart[i]) + f3[i];
      }
      __syncwarp();
      if (role == 0) {
#pragma unroll
        for (int i = 0; i < 3; ++i) fcarry[p * 3 + i] = active ? fpart[i] : 0.0;
      }
    }
    __syncwarp();
    // every warp is done reading the gradients: the buffer becomes the transpose area
    __syncthreads();
    if (writer) {
      double2 *dst = reinterpret_cast<double2 *>(geo + (p * 3 + grp) * kDumpStride);
#pragma unroll
      for (int q = 0; q < 40; ++q) {
        const int v0 = 2 * q, v1 = 2 * q + 1;      // v = i*27 + slot*3 + j
        dst[q] = make_double2(acc[(v0 % 27) / 3][(v0 / 27) * 3 + v0 % 3], acc[(v1 % 27) / 3][(v1 / 27) * 3 + v1 % 3]);
      }
      geo[(p * 3 + grp) * kDumpStride + 80] = acc[8][8];
    }
    __syncwarp();
    // 24 storing threads per warp (8 positions x 3 groups); all 32 lanes move the 81 values of one
    // of them per pass: contiguous 216-byte runs when the node's neighbours are numbered canonically
#pragma unroll
    for (int pw = 0; pw < 8; ++pw)
#pragma unroll
      for (int g3 = 0; g3 < 3; ++g3) {
        const int t = 4 * pw + (g3 == 0 ? 0 : (g3 == 1 ? 1 + par : 3));
        const bool on = (wmask >> t) & 1u;
        const RowMeta *wr = rowmeta + (((g3 == 0 ? m + 4 : m + 3)) % 3) * NP + warp * 8 + pw;
        const double *src = geo + ((warp * 8 + pw) * 3 + g3) * kDumpStride;
        double *row0 = vals + wr->base;
        const int stride = 3 * wr->nb;
#pragma unroll
        for (int it = 0; it < 3; ++it) {
          const int v = it * 32 + lane;
          const int ii = lane_i[it], sl = lane_s[it], jj = lane_j[it];
          const int k = wr->k[g3 * 9 + sl];
          if (on && v < 81 && k != 255 && !(dbg & 1)) row0[ii * stride + 3 * k + jj] = src[v];
        }
      }
    cp_async_wait_all();
    __syncthreads();
  }
}

template <int TX, int TY>
size_t fused_smem_bytes(bool with_f) {
  constexpr int NP = TX * TY, EL = (TX + 1) * (TY + 1), UL = (TX + 2) * (TY + 2);
  constexpr int GEO = (EL * kGeoStride > NP * 3 * kDumpStride) ? EL * kGeoStride : NP * 3 * kDumpStride;
  size_t b = sizeof(double) * (GEO + 3 * UL * 3 + 2 * EL);
  if (with_f) b += sizeof(double) * (3 * UL * 3 + NP * 3);
  b += sizeof(RowMeta) * 3 * NP + sizeof(int) * 4 * UL;
  return b;
}

template <int TX, int TY>
int launch_tile(femx_ctx *c, const double *d_u, double *d_vals, double *d_fint, int LZ) {
  const DevHex &hx = c->dhex;
  dim3 grid((hx.NI + TX - 1) / TX, (hx.NJ + TY - 1) / TY, (hx.NK + LZ - 1) / LZ);
  const bool wf = d_u != nullptr;
  const size_t smem = fused_smem_bytes<TX, TY>(wf);
  if (wf) {
    cudaFuncSetAttribute(hex_fused_kernel<TX, TY, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    hex_fused_kernel<TX, TY, true><<<grid, 4 * TX * TY, smem, c->stream>>>(c->dm, hx, d_u, d_vals, d_fint, LZ, c->opts.reserved[2]);
  } else {
    cudaFuncSetAttribute(hex_fused_kernel<TX, TY, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    hex_fused_kernel<TX, TY, false><<<grid, 4 * TX * TY, smem, c->stream>>>(c->dm, hx, d_u, d_vals, d_fint, LZ, c->opts.reserved[2]);
  }
  return 1;
}

}  // namespace

int launch_hex_fused(femx_ctx *c, const double *d_u, double *d_vals, double *d_fint) {
  const int tile = c->opts.reserved[0] ? c->opts.reserved[0] : 3;   // tuning knob; 3 = 4x4 positions, 4 CTAs/SM
  const int want = c->opts.reserved[1] > 0 ? c->opts.reserved[1] : 17;   // node layers per z-chunk
  const int nchunk = (c->dhex.NK + want / 2) / want > 0 ? (c->dhex.NK + want / 2) / want : 1;
  const int LZ = (c->dhex.NK + nchunk - 1) / nchunk;
  switch (tile) {
    case 1: return launch_tile<8, 8>(c, d_u, d_vals, d_fint, LZ);
    case 2: return launch_tile<8, 6>(c, d_u, d_vals, d_fint, LZ);
    case 3: return launch_tile<4, 4>(c, d_u, d_vals, d_fint, LZ);
    case 4: return launch_tile<8, 2>(c, d_u, d_vals, d_fint, LZ);
    case 5: return launch_tile<8, 4>(c, d_u, d_vals, d_fint, LZ);
    default: return launch_tile<4, 4>(c, d_u, d_vals, d_fint, LZ);
  }
}

}  // namespace femx
