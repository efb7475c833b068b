// plan_gpu.cu -- the one-time integer pass on the device (default; plan.cpp / plan_hex.cpp stay as the host cross-check,
// opts.reserved[4] = 1).
//
// Replaces what PETSc derives implicitly from the reference's MatSetValues(ADD_VALUES) calls (src/main.F90:153-171: a
// preallocation guess, then a sorted insert per entry): row r of the assembled AIJ holds, in ascending order, every
// global dof coupled to r through some element, explicit zeros included.  Same outputs as plan.cpp, bit for bit
// (tests/test_gpu_parity.py::test_device_plan_equals_host_plan), built with sorts instead of per-row loops:
//   * node -> (element, local node) incidence: stable radix sort of the connectivity by node (ascending element order
//     falls out of the stability);
//   * node-block CSR: the (row node, column node) pair of every (element, a, b), radix-sorted as one 64-bit key and
//     made unique -- bcol is the low word, bptr a lower bound per node;
//   * element -> block position: binary search of node b in the row of node a;
//   * per-row contribution lists of the fused rows kernels (rb_off / rb_ent / sx_ent / sx_inc / ri_*): one thread per node,
//     same loops as the host pass;
//   * logical-grid proof for the hex8 tile kernels (plan_hex.cpp's breadth-first walk becomes a label propagation: an
//     element with coordinates gives them to its nodes, a node with coordinates to its elements, until nothing changes;
//     then every (element, node) incidence is verified -- a consistent labelling is unique per component, so the result
//     equals the walk's), the dense logical maps, the 27-stencil block positions, the per-position coordinates and row
//     metadata.
// CUB (radix sort, scan, select) does the heavy lifting: this pass is not the hot path.
#include <cub/cub.cuh>
#include <thrust/iterator/counting_iterator.h>
#include <thrust/iterator/transform_iterator.h>

#include <chrono>
#include <cstdlib>

#include "api_util.h"

namespace femx {

namespace {

#define PCK(call)                                                                                              \
  do {                                                                                                         \
    cudaError_t _e = (call);                                                                                   \
    if (_e != cudaSuccess) {                                                                                   \
      char _b[512];                                                                                            \
      snprintf(_b, sizeof _b, "%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(_e));            \
      c->err = _b;                                                                                             \
      return FEMX_ERR_CUDA;                                                                                    \
    }                                                                                                          \
  } while (0)

struct Arena {   // the pass's temporaries: one allocation, bump pointer, freed when the pass returns
  char *base = nullptr;
  size_t cap = 0, used = 0;
  cudaError_t reserve(size_t bytes) { cap = bytes; return cudaMalloc((void **)&base, bytes); }
  template <class T> T *take(size_t n) {
    const size_t b = (sizeof(T) * (n ? n : 1) + 255) & ~(size_t)255;
    if (used + b > cap) return nullptr;
    T *q = (T *)(base + used);
    used += b;
    return q;
  }
  ~Arena() { if (base) cudaFree(base); }
};
#define TAKE(ptr, T, n)                                                        \
  do {                                                                         \
    ptr = A.take<T>(n);                                                        \
    if (!ptr) { c->err = "plan scratch arena exhausted"; return FEMX_ERR_ALLOC; } \
  } while (0)

constexpr int BS = 256;
inline int grid_for(int64_t n) { return (int)((n + BS - 1) / BS); }
inline int bits_for(int64_t n) { int b = 1; while ((1ll << b) < n) ++b; return b; }

// ---------------------------------------------------------------------------------------------- element table
struct Nen2 { __host__ __device__ int64_t operator()(int t) const { const int n = type_nen(t); return (int64_t)n * n; } };
struct Nen2D { int d2; __host__ __device__ int64_t operator()(int t) const { const int n = type_nen(t); return (int64_t)n * n * d2; } };
struct NenD { int d; __host__ __device__ int64_t operator()(int t) const { return (int64_t)type_nen(t) * d; } };
struct TypeIs { const int32_t *etype; int t; __device__ bool operator()(int e) const { return etype[e] == t; } };
struct IsNonlinear { const int32_t *enlmat; __device__ bool operator()(int e) const { return enlmat[e] != 0; } };
// flags[0] bit 3: nlMat != 0 on a non-cohesive element, bit 4: material id out of range, bit 5: a non-simplex element;
// flags[8 + t]: elements of type t; flags[6]: elements with nlMat != 0; flags[7]: largest element node count
__global__ void elem_table_kernel(int nels, int nmts, const int32_t *__restrict__ etype, const int32_t *__restrict__ emat, const int32_t *__restrict__ enlmat,
                                  int *__restrict__ flags) {
  __shared__ int cnt[8];
  if (threadIdx.x < 8) cnt[threadIdx.x] = 0;
  __syncthreads();
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e < nels) {
    const int t = etype[e];
    int bits = 0;
    if (enlmat[e] != 0) { atomicAdd(&cnt[6], 1); if (t != FEMX_COH) bits |= 8; }
    if (t != FEMX_COH && !(emat[e] >= 1 && emat[e] <= nmts)) bits |= 16;
    if (t != FEMX_TRI && t != FEMX_TET) bits |= 32;
    if (bits) atomicOr(&flags[0], bits);
    atomicAdd(&cnt[t], 1);
  }
  __syncthreads();
  if (threadIdx.x >= 1 && threadIdx.x <= 6 && cnt[threadIdx.x]) atomicAdd(&flags[threadIdx.x == 6 ? 6 : 8 + threadIdx.x], cnt[threadIdx.x]);
}

// ---------------------------------------------------------------------------------------------- incidence
// conn0 (0-based local nodes), sort keys (node) and payload (position in the connectivity), element of every position
__global__ void conn_kernel(int nels, const int32_t *__restrict__ eptr, const int32_t *__restrict__ eind, int nlnds, int32_t *__restrict__ conn0,
                            uint32_t *__restrict__ key, uint32_t *__restrict__ pos, int32_t *__restrict__ posel, int *__restrict__ flags) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= nels) return;
  for (int i = eptr[e]; i < eptr[e + 1]; ++i) {
    int v = eind[i] - 1;
    if (v < 0 || v >= nlnds) { atomicOr(flags, 1); v = 0; }
    conn0[i] = v; key[i] = (uint32_t)v; pos[i] = (uint32_t)i; posel[i] = e;
  }
}
__global__ void lower_bound_u32_kernel(int n, const uint32_t *__restrict__ sorted, int64_t m, int64_t *__restrict__ out) {   // out[r] = first index with sorted >= r, r = 0..n
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r > n) return;
  int64_t lo = 0, hi = m;
  while (lo < hi) { const int64_t mid = (lo + hi) >> 1; if (sorted[mid] < (uint32_t)r) lo = mid + 1; else hi = mid; }
  out[r] = lo;
}
__global__ void n2e_kernel(int64_t nconn, const uint32_t *__restrict__ pos_sorted, const int32_t *__restrict__ posel, const int32_t *__restrict__ eptr,
                           int32_t *__restrict__ n2e_elem, uint8_t *__restrict__ n2e_a) {
  const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= nconn) return;
  const uint32_t i = pos_sorted[p];
  const int e = posel[i];
  n2e_elem[p] = e; n2e_a[p] = (uint8_t)(i - (uint32_t)eptr[e]);
}
// ---------------------------------------------------------------------------------------------- block rows
__global__ void pair_kernel(int64_t nconn, const int32_t *__restrict__ posel, const int32_t *__restrict__ eptr, const int64_t *__restrict__ eb_ptr,
                            const int32_t *__restrict__ conn0, uint64_t *__restrict__ key) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nconn) return;
  const int e = posel[i], e0 = eptr[e], nen = eptr[e + 1] - e0, a = (int)(i - e0);
  const uint64_t r = (uint64_t)(uint32_t)conn0[i] << 32;
  uint64_t *out = key + eb_ptr[e] + (int64_t)a * nen;
  for (int b = 0; b < nen; ++b) out[b] = r | (uint32_t)conn0[e0 + b];
}
__global__ void bcol_kernel(int64_t nblk, const uint64_t *__restrict__ ukey, int32_t *__restrict__ bcol) {
  const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k < nblk) bcol[k] = (int32_t)(uint32_t)ukey[k];
}
__global__ void bptr_kernel(int nlnds, const uint64_t *__restrict__ ukey, int64_t nblk, int64_t *__restrict__ bptr) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r > nlnds) return;
  const uint64_t want = (uint64_t)(uint32_t)r << 32;
  int64_t lo = 0, hi = nblk;
  while (lo < hi) { const int64_t mid = (lo + hi) >> 1; if (ukey[mid] < want) lo = mid + 1; else hi = mid; }
  bptr[r] = lo;
}
__device__ __forceinline__ int row_find(const int32_t *__restrict__ b0, int nb, int32_t want) {   // std::lower_bound
  int lo = 0, hi = nb;
  while (lo < hi) { const int mid = (lo + hi) >> 1; if (b0[mid] < want) lo = mid + 1; else hi = mid; }
  return lo;
}
// bdiag, widest row, incidence statistics (largest number of elements around a node, list entries per node)
__global__ void node_stats_kernel(int nlnds, const int64_t *__restrict__ bptr, const int32_t *__restrict__ bcol, const int64_t *__restrict__ n2e_ptr,
                                  const int32_t *__restrict__ n2e_elem, const int32_t *__restrict__ eptr, const int32_t *__restrict__ etype, int coh_counts,
                                  int32_t *__restrict__ bdiag, int64_t *__restrict__ rbe_cnt, int32_t *__restrict__ node_count, int *__restrict__ stats) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= nlnds) return;
  const int nb = (int)(bptr[r + 1] - bptr[r]);
  bdiag[r] = row_find(bcol + bptr[r], nb, r);
  int64_t n = 0;
  int cnt = 0;
  for (int64_t p = n2e_ptr[r]; p < n2e_ptr[r + 1]; ++p) {
    const int e = n2e_elem[p];
    n += eptr[e + 1] - eptr[e];
    if (etype[e] != FEMX_COH || coh_counts) ++cnt;
  }
  rbe_cnt[r] = n;
  node_count[r] = cnt;
  atomicMax(&stats[0], nb);
  atomicMax(&stats[1], (int)(n2e_ptr[r + 1] - n2e_ptr[r]));
}
// simplex rows kernel: the nodes of every element around node r are columns of r's block row -- name them by their position
// there (one byte each), so that the kernel fetches the coordinates of a row's column nodes once instead of once per incidence
__global__ void sx_loc_kernel(int nlnds, int nen, const int64_t *__restrict__ n2e_ptr, const int64_t *__restrict__ bptr, const int32_t *__restrict__ bcol,
                              const int32_t *__restrict__ sx_inc, uint32_t *__restrict__ sx_loc, int *__restrict__ max_nb) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= nlnds) return;
  const int32_t *b0 = bcol + bptr[r];
  const int nb = (int)(bptr[r + 1] - bptr[r]);
  for (int64_t p = n2e_ptr[r]; p < n2e_ptr[r + 1]; ++p) {
    uint32_t w = 0;
    for (int b = 0; b < nen; ++b) w |= (uint32_t)row_find(b0, nb, sx_inc[4 * p + b]) << (8 * b);
    sx_loc[p] = w;
  }
  atomicMax(max_nb, nb);
}
__global__ void bpos_kernel(int64_t nconn, const int32_t *__restrict__ posel, const int32_t *__restrict__ eptr, const int64_t *__restrict__ eb_ptr,
                            const int32_t *__restrict__ conn0, const int64_t *__restrict__ bptr, const int32_t *__restrict__ bcol, uint16_t *__restrict__ bpos) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nconn) return;
  const int e = posel[i], e0 = eptr[e], nen = eptr[e + 1] - e0, a = (int)(i - e0), r = conn0[i];
  const int32_t *b0 = bcol + bptr[r];
  const int nb = (int)(bptr[r + 1] - bptr[r]);
  uint16_t *out = bpos + eb_ptr[e] + (int64_t)a * nen;
  for (int b = 0; b < nen; ++b) out[b] = (uint16_t)row_find(b0, nb, conn0[e0 + b]);
}
// ---------------------------------------------------------------------------------------------- row gather lists
// (plan.cpp "row gather lists": invert bpos per node; needs at most 32 elements around a node, hence rows of <= 225 blocks)
constexpr int kMaxRowBlocks = 32 * (kMaxNen - 1) + 1;
__global__ void row_lists_kernel(int nlnds, int pd, int slot, int simplex, const int64_t *__restrict__ bptr, const int64_t *__restrict__ n2e_ptr,
                                 const int32_t *__restrict__ n2e_elem, const uint8_t *__restrict__ n2e_a, const int32_t *__restrict__ eptr,
                                 const int64_t *__restrict__ eb_ptr, const uint16_t *__restrict__ bpos, const int64_t *__restrict__ rbe_ptr,
                                 uint16_t *__restrict__ rb_off, uint16_t *__restrict__ rb_ent, uint16_t *__restrict__ sx_ent) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= nlnds) return;
  const int nb = (int)(bptr[r + 1] - bptr[r]), d2 = pd * pd;
  uint16_t cntk[kMaxRowBlocks + 1];
  for (int k = 0; k <= nb; ++k) cntk[k] = 0;
  for (int64_t p = n2e_ptr[r]; p < n2e_ptr[r + 1]; ++p) {
    const int e = n2e_elem[p], a = n2e_a[p], nen = eptr[e + 1] - eptr[e];
    const uint16_t *bp = bpos + eb_ptr[e] + a * nen;
    for (int b = 0; b < nen; ++b) ++cntk[bp[b] + 1];
  }
  for (int k = 0; k < nb; ++k) cntk[k + 1] += cntk[k];
  uint16_t *off = rb_off + bptr[r] + r;
  for (int k = 0; k <= nb; ++k) off[k] = cntk[k];
  uint16_t *ent = rb_ent + rbe_ptr[r];
  int idx = 0;
  for (int64_t p = n2e_ptr[r]; p < n2e_ptr[r + 1]; ++p, ++idx) {   // ascending element order
    const int e = n2e_elem[p], a = n2e_a[p], nen = eptr[e + 1] - eptr[e];
    const uint16_t *bp = bpos + eb_ptr[e] + a * nen;
    for (int b = 0; b < nen; ++b) {
      const int cc = cntk[bp[b]]++;
      ent[cc] = (uint16_t)(idx * slot + b * d2);
      if (simplex) sx_ent[rbe_ptr[r] + cc] = (uint16_t)((idx << 4) | (a << 2) | b);
    }
  }
}
// flattened incidence records: simplex rows kernel (sx_inc, sx_mat) and 2-D fused rows kernel (ri_nodes, ri_info)
__global__ void incidence_records_kernel(int64_t ninc, const int32_t *__restrict__ n2e_elem, const uint8_t *__restrict__ n2e_a,
                                         const int32_t *__restrict__ eptr, const int32_t *__restrict__ conn0, const int32_t *__restrict__ etype,
                                         const int32_t *__restrict__ emat, int32_t *__restrict__ sx_inc, int32_t *__restrict__ sx_mat,
                                         int32_t *__restrict__ ri_nodes, int32_t *__restrict__ ri_info) {
  const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= ninc) return;
  const int e = n2e_elem[p], e0 = eptr[e], nen = eptr[e + 1] - e0;
  int nd[4] = {0, 0, 0, 0};
  for (int b = 0; b < nen && b < 4; ++b) nd[b] = conn0[e0 + b];
  if (sx_inc) {
    for (int b = 0; b < 4; ++b) sx_inc[4 * p + b] = nd[b];
    sx_mat[p] = emat[e] - 1;
  }
  if (ri_nodes) {
    for (int b = 0; b < 4; ++b) ri_nodes[4 * p + b] = nd[b];
    const int m = etype[e] == FEMX_COH ? 0 : emat[e] - 1;
    ri_info[p] = etype[e] | ((int)n2e_a[p] << 4) | (m << 8);
  }
}
// pure tri3 / tet4 meshes: any element with detJ <= 0 (femx_build_plan's simplex admission test)
__global__ void simplex_detj_kernel(int nels, int pd, const int32_t *__restrict__ eptr, const int32_t *__restrict__ conn0, const double *__restrict__ coords,
                                    int *__restrict__ bad) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= nels) return;
  const int32_t *nd = conn0 + eptr[e];
  double d[3][3];
  for (int b = 1; b <= pd; ++b)
    for (int k = 0; k < pd; ++k) d[b - 1][k] = coords[(size_t)nd[b] * pd + k] - coords[(size_t)nd[0] * pd + k];
  // same expression, same rounding as the host test (no contraction)
  const double det = pd == 2 ? __dsub_rn(__dmul_rn(d[0][0], d[1][1]), __dmul_rn(d[0][1], d[1][0]))
                             : __dadd_rn(__dsub_rn(__dmul_rn(d[0][0], __dsub_rn(__dmul_rn(d[1][1], d[2][2]), __dmul_rn(d[1][2], d[2][1]))),
                                                   __dmul_rn(d[0][1], __dsub_rn(__dmul_rn(d[1][0], d[2][2]), __dmul_rn(d[1][2], d[2][0])))),
                                         __dmul_rn(d[0][2], __dsub_rn(__dmul_rn(d[1][0], d[2][1]), __dmul_rn(d[1][1], d[2][0]))));
  if (!(det > 0.0)) atomicOr(bad, 1);
}
// cohesive overlay: block id bptr[r] + bpos of every (cohesive element, a, b)
__global__ void coh_blocks_kernel(int ncoh, const int32_t *__restrict__ coh_elems, const int32_t *__restrict__ eptr, const int32_t *__restrict__ conn0,
                                  const int64_t *__restrict__ eb_ptr, const uint16_t *__restrict__ bpos, const int64_t *__restrict__ bptr,
                                  int32_t *__restrict__ rows, int64_t *__restrict__ blk) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= ncoh * 16) return;
  const int ci = t >> 4, a = (t >> 2) & 3, b = t & 3, e = coh_elems[ci], r = conn0[eptr[e] + a];
  rows[ci * 4 + a] = r;
  blk[t] = bptr[r] + bpos[eb_ptr[e] + a * 4 + b];
}

// ---------------------------------------------------------------------------------------------- hex8 logical grid
__device__ const int kSx[8] = {0, 1, 1, 0, 0, 1, 1, 0};
__device__ const int kSy[8] = {0, 0, 1, 1, 0, 0, 1, 1};
__device__ const int kSz[8] = {0, 0, 0, 0, 1, 1, 1, 1};
constexpr int UNSET = INT32_MIN;

// one propagation sweep: an element without coordinates takes them from any labelled node; a labelled element labels its
// unlabelled nodes.  Racing writers of one node write the same triple when the mesh is consistent; when it is not, the
// verification below fails, whatever got written.
__global__ void hex_sweep_kernel(int nels, const int32_t *__restrict__ conn0, int32_t *__restrict__ ec, int32_t *__restrict__ nc, int *__restrict__ changed) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= nels) return;
  const int32_t *cn = conn0 + 8 * (size_t)e;
  volatile int32_t *vn = nc;
  int x = ec[3 * (size_t)e], y = ec[3 * (size_t)e + 1], z = ec[3 * (size_t)e + 2];
  if (x != UNSET && (y == UNSET || z == UNSET)) x = UNSET;
  const bool was_set = x != UNSET;
  // an unlabelled element polls its nodes a few times per launch: labels written by CTAs running at the same time travel
  // several elements per sweep (how far is timing-dependent, the fixed point is not)
  for (int rep = 0; rep < 8 && x == UNSET; ++rep)
    for (int a = 0; a < 8; ++a) {
      const int r = cn[a];
      const int nx = vn[3 * (size_t)r], ny = vn[3 * (size_t)r + 1], nz = vn[3 * (size_t)r + 2];
      if (nx != UNSET && ny != UNSET && nz != UNSET) { x = nx - kSx[a]; y = ny - kSy[a]; z = nz - kSz[a]; break; }
    }
  if (x == UNSET) return;
  bool any = !was_set;
  if (!was_set) { ec[3 * (size_t)e + 1] = y; ec[3 * (size_t)e + 2] = z; ec[3 * (size_t)e] = x; }
  for (int b = 0; b < 8; ++b) {
    const int r = cn[b];
    if (vn[3 * (size_t)r] == UNSET || vn[3 * (size_t)r + 1] == UNSET || vn[3 * (size_t)r + 2] == UNSET) {
      vn[3 * (size_t)r + 1] = y + kSy[b]; vn[3 * (size_t)r + 2] = z + kSz[b]; vn[3 * (size_t)r] = x + kSx[b];
      any = true;
    }
  }
  if (any) *changed = 1;
}
// after convergence: first element still without coordinates (next component's seed), largest i so far, bounding box, consistency
__global__ void hex_scan_kernel(int nels, int nlnds, const int32_t *__restrict__ conn0, const int32_t *__restrict__ ec, const int32_t *__restrict__ nc,
                                int *__restrict__ out /* [0] first unset element, [1] max i, [2..4] min, [5..7] max, [8] inconsistent, [9] unset node */) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < nels) {
    const int x = ec[3 * (size_t)t], y = ec[3 * (size_t)t + 1], z = ec[3 * (size_t)t + 2];
    if (x == UNSET) atomicMin(&out[0], t);
    else
      for (int b = 0; b < 8; ++b) {
        const int r = conn0[8 * (size_t)t + b];
        if (nc[3 * (size_t)r] != x + kSx[b] || nc[3 * (size_t)r + 1] != y + kSy[b] || nc[3 * (size_t)r + 2] != z + kSz[b]) out[8] = 1;
      }
  }
  if (t < nlnds) {
    const int x = nc[3 * (size_t)t], y = nc[3 * (size_t)t + 1], z = nc[3 * (size_t)t + 2];
    if (x == UNSET) out[9] = 1;
    else {
      atomicMax(&out[1], x);
      atomicMin(&out[2], x); atomicMin(&out[3], y); atomicMin(&out[4], z);
      atomicMax(&out[5], x); atomicMax(&out[6], y); atomicMax(&out[7], z);
    }
  }
}
__global__ void hex_seed_kernel(int e, int off, const int32_t *__restrict__ conn0, int32_t *__restrict__ ec) {
  ec[3 * (size_t)e] = off; ec[3 * (size_t)e + 1] = 0; ec[3 * (size_t)e + 2] = 0;
  (void)conn0;
}
__global__ void hex_place_kernel(int nels, int nlnds, int NI, int NJ, int mi, int mj, int mk, const int32_t *__restrict__ ec, const int32_t *__restrict__ nc,
                                 int32_t *__restrict__ node_at, int32_t *__restrict__ elem_at, int *__restrict__ err) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < nlnds) {
    const int64_t q = ((int64_t)(nc[3 * (size_t)t + 2] - mk) * NJ + (nc[3 * (size_t)t + 1] - mj)) * NI + (nc[3 * (size_t)t] - mi);
    if (atomicCAS(&node_at[q], -1, t) != -1) atomicOr(err, 1);
  }
  if (t < nels) {
    const int64_t q = ((int64_t)(ec[3 * (size_t)t + 2] - mk) * NJ + (ec[3 * (size_t)t + 1] - mj)) * NI + (ec[3 * (size_t)t] - mi);
    if (atomicCAS(&elem_at[q], -1, t) != -1) atomicOr(err, 2);
  }
}
// canonical neighbour -> block position (every block of the row must be reached this way), per-position coordinates and row
// metadata, padded coordinates
__global__ void hex_rows_kernel(int nlnds, int NI, int NJ, int NK, int mi, int mj, int mk, const int32_t *__restrict__ nc, const int32_t *__restrict__ node_at,
                                const int64_t *__restrict__ bptr, const int32_t *__restrict__ bcol, const int32_t *__restrict__ nl2g,
                                const double *__restrict__ coords, uint8_t *__restrict__ kmap, double *__restrict__ coords4, double *__restrict__ pos_x,
                                HexRowMeta *__restrict__ pos_meta, int *__restrict__ err) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= nlnds) return;
  const int i = nc[3 * (size_t)r] - mi, j = nc[3 * (size_t)r + 1] - mj, k = nc[3 * (size_t)r + 2] - mk;
  const int32_t *b0 = bcol + bptr[r];
  const int nb = (int)(bptr[r + 1] - bptr[r]);
  HexRowMeta mm;
  mm.base = 9 * bptr[r]; mm.nb = nb; mm.grow = 3 * (nl2g[r] - 1);
  for (int t = 0; t < 28; ++t) mm.k[t] = 255;
  int found = 0;
  for (int dz = -1; dz <= 1; ++dz)
    for (int dy = -1; dy <= 1; ++dy)
      for (int dx = -1; dx <= 1; ++dx) {
        const int x = i + dx, y = j + dy, z = k + dz;
        if (x < 0 || y < 0 || z < 0 || x >= NI || y >= NJ || z >= NK) continue;
        const int cn = node_at[((int64_t)z * NJ + y) * NI + x];
        if (cn < 0) continue;
        const int p = row_find(b0, nb, cn);
        if (p == nb || b0[p] != cn) continue;   // present on the grid but not coupled through an element
        mm.k[(dz + 1) * 9 + (dy + 1) * 3 + (dx + 1)] = (unsigned char)p;
        ++found;
      }
  if (found != nb || nb > 27) atomicOr(err, 4);
  for (int t = 0; t < 32; ++t) kmap[(size_t)r * 32 + t] = t < 27 ? mm.k[t] : 255;
  const int64_t q = ((int64_t)k * NJ + j) * NI + i;
  for (int d = 0; d < 3; ++d) {
    const double v = coords[(size_t)r * 3 + d];
    coords4[(size_t)r * 4 + d] = v;
    pos_x[q * 3 + d] = v;
  }
  coords4[(size_t)r * 4 + 3] = 0.0;
  pos_meta[q] = mm;
}
__global__ void hex_meta_init_kernel(int64_t n, HexRowMeta *__restrict__ pos_meta) {
  const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= n) return;
  HexRowMeta mm;
  mm.base = 0; mm.nb = 0; mm.grow = 0;
  for (int t = 0; t < 28; ++t) mm.k[t] = 255;
  pos_meta[q] = mm;
}
__global__ void fill_i32_kernel(int64_t n, int32_t *__restrict__ p, int32_t v) {
  const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (q < n) p[q] = v;
}

template <class T>
int tracked(femx_ctx *c, T **p, int64_t n) { return dev_alloc(c, p, n); }

// the hex8 logical-grid plan (plan_hex.cpp on the device).  Leaves c->hex.ok / why and c->dhex.
int hex_plan_device(femx_ctx *c, cudaStream_t st, Arena &A, bool all_hex) {
  HexPlan &H = c->hex;
  DevMesh &dm = c->dm;
  H = HexPlan();
  const int nels = c->nels, nn = c->nlnds;
  if (c->pdim != 3 || nels == 0) { H.why = "not a 3-D mesh"; return FEMX_OK; }
  if (!all_hex) { H.why = "mesh holds non-hex8 elements"; return FEMX_OK; }
  int32_t *ec = nullptr, *nc = nullptr;
  int *d_out = nullptr;
  TAKE(ec, int32_t, 3 * (size_t)nels); TAKE(nc, int32_t, 3 * (size_t)nn); TAKE(d_out, int, 16);
  fill_i32_kernel<<<grid_for(3 * (int64_t)nels), BS, 0, st>>>(3 * (int64_t)nels, ec, UNSET);
  fill_i32_kernel<<<grid_for(3 * (int64_t)nn), BS, 0, st>>>(3 * (int64_t)nn, nc, UNSET);
  int h_out[16];
  int seed = 0, offset = 0;
  const int nmax = std::max(nels, nn);
  for (int comp = 0;; ++comp) {
    if (comp > 64) { H.why = "more than 64 disconnected components"; return FEMX_OK; }
    hex_seed_kernel<<<1, 1, 0, st>>>(seed, offset, dm.conn0, ec);
    for (int64_t sweeps = 0;;) {
      PCK(cudaMemsetAsync(d_out, 0, sizeof(int), st));
      for (int s = 0; s < 8; ++s) hex_sweep_kernel<<<grid_for(nels), BS, 0, st>>>(nels, dm.conn0, ec, nc, d_out);
      sweeps += 8;
      PCK(cudaMemcpyAsync(h_out, d_out, sizeof(int), cudaMemcpyDeviceToHost, st));
      PCK(cudaStreamSynchronize(st));
      if (!h_out[0]) break;
      if (sweeps > 4ll * nels + 64) { H.why = "label propagation did not converge"; return FEMX_OK; }
    }
    const int init[16] = {INT32_MAX, INT32_MIN, INT32_MAX, INT32_MAX, INT32_MAX, INT32_MIN, INT32_MIN, INT32_MIN, 0, 0, 0, 0, 0, 0, 0, 0};
    PCK(cudaMemcpyAsync(d_out, init, sizeof(init), cudaMemcpyHostToDevice, st));
    hex_scan_kernel<<<grid_for(nmax), BS, 0, st>>>(nels, nn, dm.conn0, ec, nc, d_out);
    PCK(cudaMemcpyAsync(h_out, d_out, sizeof(h_out), cudaMemcpyDeviceToHost, st));
    PCK(cudaStreamSynchronize(st));
    if (h_out[8]) { H.why = "elements are not consistently oriented on a logical grid"; return FEMX_OK; }
    if (h_out[0] == INT32_MAX) break;          // every element labelled
    seed = h_out[0];
    offset = h_out[1] + 3;                     // keep disconnected components apart
  }
  if (h_out[9]) { H.why = "a node belongs to no element"; return FEMX_OK; }
  const int mi = h_out[2], mj = h_out[3], mk = h_out[4];
  const int64_t NI = (int64_t)h_out[5] - mi + 1, NJ = (int64_t)h_out[6] - mj + 1, NK = (int64_t)h_out[7] - mk + 1;
  if (NI * NJ * NK > 4 * (int64_t)nn + 4096) { H.why = "logical bounding grid is too sparse"; return FEMX_OK; }
  const int64_t npos = NI * NJ * NK;
  DevHex &hx = c->dhex;
  int rc;
  if ((rc = tracked(c, &hx.node_at, npos)) || (rc = tracked(c, &hx.elem_at, npos)) || (rc = tracked(c, &hx.kmap, 32 * (int64_t)nn)) ||
      (rc = tracked(c, &hx.coords4, 4 * (int64_t)nn)) || (rc = tracked(c, &hx.pos_x, 3 * npos)) || (rc = tracked(c, &hx.pos_meta, npos)))
    return rc;
  fill_i32_kernel<<<grid_for(npos), BS, 0, st>>>(npos, hx.node_at, -1);
  fill_i32_kernel<<<grid_for(npos), BS, 0, st>>>(npos, hx.elem_at, -1);
  PCK(cudaMemsetAsync(hx.pos_x, 0, sizeof(double) * 3 * (size_t)npos, st));
  hex_meta_init_kernel<<<grid_for(npos), BS, 0, st>>>(npos, hx.pos_meta);
  PCK(cudaMemsetAsync(d_out, 0, sizeof(int), st));
  hex_place_kernel<<<grid_for(nmax), BS, 0, st>>>(nels, nn, (int)NI, (int)NJ, mi, mj, mk, ec, nc, hx.node_at, hx.elem_at, d_out);
  hex_rows_kernel<<<grid_for(nn), BS, 0, st>>>(nn, (int)NI, (int)NJ, (int)NK, mi, mj, mk, nc, hx.node_at, dm.bptr, dm.bcol, dm.nl2g, dm.coords, hx.kmap,
                                               hx.coords4, hx.pos_x, hx.pos_meta, d_out);
  PCK(cudaMemcpyAsync(h_out, d_out, sizeof(int), cudaMemcpyDeviceToHost, st));
  PCK(cudaStreamSynchronize(st));
  PCK(cudaGetLastError());
  if (h_out[0] & 1) { H.why = "two nodes share one logical position"; return FEMX_OK; }
  if (h_out[0] & 2) { H.why = "two elements share one logical cell"; return FEMX_OK; }
  if (h_out[0] & 4) { H.why = "a node couples to nodes outside its 27-point stencil"; return FEMX_OK; }
  H.NI = (int)NI; H.NJ = (int)NJ; H.NK = (int)NK;
  hx.NI = H.NI; hx.NJ = H.NJ; hx.NK = H.NK;
  H.elem_at.resize((size_t)npos);   // refresh_materials maps cells to materials on the host
  PCK(cudaMemcpyAsync(H.elem_at.data(), hx.elem_at, sizeof(int32_t) * (size_t)npos, cudaMemcpyDeviceToHost, st));
  PCK(cudaStreamSynchronize(st));
  H.ok = true;
  return FEMX_OK;
}

}  // namespace

// Everything femx_build_plan derives from the connectivity, on the device.  Fills c->dm (mesh + plan arrays), c->plan's
// scalars and small host arrays (bptr, bdiag, cohesive overlay), c->buckets, c->hex / c->dhex.  The large host arrays
// (bcol, bpos, eb_ptr) are fetched on demand by plan_host_blocks / plan_host_bpos.
int build_plan_device(femx_ctx *c) {
  const auto t0 = std::chrono::steady_clock::now();
  cudaStream_t st = c->stream;
  const bool timing = getenv("FEMX_PLAN_TIMING") != nullptr;   // stage times to stderr (synchronises between stages)
  auto tprev = t0;
  auto stage = [&](const char *name) {
    if (!timing) return;
    cudaStreamSynchronize(st);
    const auto now = std::chrono::steady_clock::now();
    fprintf(stderr, "femx plan: %-28s %8.3f ms\n", name, std::chrono::duration<double, std::milli>(now - tprev).count());
    tprev = now;
  };
  Plan &P = c->plan;
  P = Plan();
  const int pdim = c->pdim, nels = c->nels, nlnds = c->nlnds, d2 = pdim * pdim;
  P.pdim = pdim; P.nels = nels; P.nlnds = nlnds; P.nnds = c->nnds;
  const int64_t nconn = c->eptr[nels];
  // femx_set_mesh has validated type codes, node counts, dimensions and node ids; what depends on later calls (materials,
  // nlMat) is checked below on the device
  const int64_t npairs_max = (int64_t)kMaxNen * nconn;
  DevMesh &dm = c->dm;
  dm.pdim = pdim; dm.nels = nels; dm.nlnds = nlnds; dm.nnds = c->nnds; dm.cp = pdim == 2 ? 3 : 6;
  int rc;
  Arena A;
  {
    const size_t bytes = 24 * (size_t)nconn + 16 * (size_t)npairs_max + 12 * ((size_t)nels + nlnds) + 16 * ((size_t)nlnds + 1) + 24 * ((size_t)nels + 1) +
                         ((size_t)96 << 20);
    cudaError_t e = A.reserve(bytes);
    if (e != cudaSuccess) { c->err = std::string("cudaMalloc (plan scratch) failed: ") + cudaGetErrorString(e); return FEMX_ERR_ALLOC; }
  }
#define UP(dst, src) if ((rc = upload(c, &dst, src))) return rc
  UP(dm.eptr, c->eptr); UP(dm.emat, c->emat); UP(dm.enlmat, c->enlmat); UP(dm.nl2g, c->nl2g); UP(dm.coords, c->coords);
  if ((rc = tracked(c, &dm.etype, (int64_t)nels + 1))) return rc;   // one readable sentinel behind the last element (scan inputs)
  PCK(cudaMemsetAsync(dm.etype + nels, 0, sizeof(int32_t), st));
  if (nels) PCK(cudaMemcpyAsync(dm.etype, c->etype.data(), sizeof(int32_t) * (size_t)nels, cudaMemcpyHostToDevice, st));
  int32_t *d_eind = nullptr, *posel = nullptr;
  uint32_t *key = nullptr, *pos = nullptr, *key2 = nullptr, *pos2 = nullptr;
  int *d_flags = nullptr;
  void *cubtmp = nullptr;
  const size_t cubtmp_bytes = (size_t)64 << 20;   // CUB temporaries (histograms, tile states): O(tiles), far below this
  TAKE(d_eind, int32_t, nconn); TAKE(posel, int32_t, nconn);
  TAKE(key, uint32_t, nconn); TAKE(pos, uint32_t, nconn); TAKE(key2, uint32_t, nconn); TAKE(pos2, uint32_t, nconn);
  TAKE(d_flags, int, 32); TAKE(cubtmp, char, cubtmp_bytes);
  PCK(cudaMemsetAsync(d_flags, 0, 32 * sizeof(int), st));
  if (nconn) PCK(cudaMemcpyAsync(d_eind, c->eind.data(), sizeof(int32_t) * (size_t)nconn, cudaMemcpyHostToDevice, st));
  stage("mesh uploads");
  // ---- element table: packed offsets (scans over the type codes), type counts, checks
  if ((rc = tracked(c, &dm.eb_ptr, (int64_t)nels + 1)) || (rc = tracked(c, &dm.ke_off, (int64_t)nels + 1)) || (rc = tracked(c, &dm.fe_off, (int64_t)nels + 1))) return rc;
  if (nels) elem_table_kernel<<<grid_for(nels), BS, 0, st>>>(nels, c->nmts, dm.etype, dm.emat, dm.enlmat, d_flags);
  {
    auto scan = [&](auto op, int64_t *out) -> cudaError_t {
      auto in = thrust::make_transform_iterator((const int32_t *)dm.etype, op);
      size_t tb = 0;
      cudaError_t e = cub::DeviceScan::ExclusiveSum(nullptr, tb, in, out, nels + 1, st);
      if (e != cudaSuccess) return e;
      if (tb > cubtmp_bytes) return cudaErrorMemoryAllocation;
      return cub::DeviceScan::ExclusiveSum(cubtmp, tb, in, out, nels + 1, st);
    };
    PCK(scan(Nen2{}, dm.eb_ptr)); PCK(scan(Nen2D{d2}, dm.ke_off)); PCK(scan(NenD{pdim}, dm.fe_off));
  }
  int h_flags[32];
  int64_t tot[3] = {0, 0, 0};
  PCK(cudaMemcpyAsync(h_flags, d_flags, sizeof(h_flags), cudaMemcpyDeviceToHost, st));
  PCK(cudaMemcpyAsync(&tot[0], dm.eb_ptr + nels, sizeof(int64_t), cudaMemcpyDeviceToHost, st));
  PCK(cudaMemcpyAsync(&tot[1], dm.ke_off + nels, sizeof(int64_t), cudaMemcpyDeviceToHost, st));
  PCK(cudaMemcpyAsync(&tot[2], dm.fe_off + nels, sizeof(int64_t), cudaMemcpyDeviceToHost, st));
  PCK(cudaStreamSynchronize(st));
  if (h_flags[0] & 8) { c->err = "nlMat != 0 on a non-cohesive element"; return FEMX_ERR_UNSUPPORTED; }
  if (h_flags[0] & 16) { c->err = "element material id out of range (femx_set_materials first)"; return FEMX_ERR_ARG; }
  const int64_t npairs = tot[0];
  P.eb_ptr.clear();                       // fetched with bpos (plan_host_bpos)
  P.ke_off.assign(nels + 1, 0); P.fe_off.assign(nels + 1, 0);   // only the totals are read on the host
  P.ke_off[nels] = tot[1]; P.fe_off[nels] = tot[2];
  c->plan_npairs = npairs;
  const bool simplex = nels > 0 && !(h_flags[0] & 32);
  const int ncoh = h_flags[6];
  int maxnen = 3;
  for (int t = FEMX_TRI; t <= FEMX_COH; ++t) if (h_flags[8 + t]) maxnen = std::max(maxnen, type_nen(t));
  stage("element table");
  if ((rc = tracked(c, &dm.conn0, nconn)) || (rc = tracked(c, &dm.n2e_ptr, (int64_t)nlnds + 1)) || (rc = tracked(c, &dm.n2e_elem, nconn)) ||
      (rc = tracked(c, &dm.n2e_a, nconn)))
    return rc;
  // ---- incidence
  if (nels) conn_kernel<<<grid_for(nels), BS, 0, st>>>(nels, dm.eptr, d_eind, nlnds, dm.conn0, key, pos, posel, d_flags + 16);
  {
    cub::DoubleBuffer<uint32_t> dk(key, key2), dv(pos, pos2);
    size_t tb = 0;
    PCK(cub::DeviceRadixSort::SortPairs(nullptr, tb, dk, dv, nconn, 0, bits_for(std::max(nlnds, 2)), st));
    if (tb > cubtmp_bytes) { c->err = "CUB temporary storage exceeds the plan arena"; return FEMX_ERR_ALLOC; }
    PCK(cub::DeviceRadixSort::SortPairs(cubtmp, tb, dk, dv, nconn, 0, bits_for(std::max(nlnds, 2)), st));
    lower_bound_u32_kernel<<<grid_for((int64_t)nlnds + 1), BS, 0, st>>>(nlnds, dk.Current(), nconn, dm.n2e_ptr);
    if (nconn) n2e_kernel<<<grid_for(nconn), BS, 0, st>>>(nconn, dv.Current(), posel, dm.eptr, dm.n2e_elem, dm.n2e_a);
  }
  stage("incidence (sort by node)");
  // ---- block rows
  if ((rc = tracked(c, &dm.bptr, (int64_t)nlnds + 1))) return rc;
  int64_t nblk = 0;
  {
    uint64_t *pk = nullptr, *pk2 = nullptr;
    int64_t *d_nsel = nullptr;
    TAKE(pk, uint64_t, npairs); TAKE(pk2, uint64_t, npairs); TAKE(d_nsel, int64_t, 1);
    if (nconn) pair_kernel<<<grid_for(nconn), BS, 0, st>>>(nconn, posel, dm.eptr, dm.eb_ptr, dm.conn0, pk);
    const int endbit = 32 + bits_for(std::max(nlnds, 2));
    cub::DoubleBuffer<uint64_t> dk(pk, pk2);
    size_t tb = 0, tb2 = 0;
    PCK(cub::DeviceRadixSort::SortKeys(nullptr, tb, dk, npairs, 0, endbit, st));
    PCK(cub::DeviceSelect::Unique(nullptr, tb2, pk, pk2, d_nsel, npairs, st));
    if (std::max(tb, tb2) > cubtmp_bytes) { c->err = "CUB temporary storage exceeds the plan arena"; return FEMX_ERR_ALLOC; }
    PCK(cub::DeviceRadixSort::SortKeys(cubtmp, tb, dk, npairs, 0, endbit, st));
    uint64_t *sorted = dk.Current(), *uniq = dk.Alternate();
    PCK(cub::DeviceSelect::Unique(cubtmp, tb2, sorted, uniq, d_nsel, npairs, st));
    PCK(cudaMemcpyAsync(&nblk, d_nsel, sizeof(int64_t), cudaMemcpyDeviceToHost, st));
    PCK(cudaStreamSynchronize(st));
    if ((rc = tracked(c, &dm.bcol, nblk))) return rc;
    if (nblk) bcol_kernel<<<grid_for(nblk), BS, 0, st>>>(nblk, uniq, dm.bcol);
    bptr_kernel<<<grid_for((int64_t)nlnds + 1), BS, 0, st>>>(nlnds, uniq, nblk, dm.bptr);
  }
  P.nblk = nblk;
  c->nnz = (int64_t)pdim * pdim * nblk;
  stage("block rows (sort + unique)");
  // ---- per-node statistics, bdiag, stress multiplicity, list sizes
  int64_t *rbe_cnt = nullptr;
  TAKE(rbe_cnt, int64_t, (size_t)nlnds + 1);
  PCK(cudaMemsetAsync(rbe_cnt, 0, sizeof(int64_t) * ((size_t)nlnds + 1), st));
  if ((rc = tracked(c, &dm.bdiag, nlnds)) || (rc = tracked(c, &dm.node_count, nlnds)) || (rc = tracked(c, &dm.rbe_ptr, (int64_t)nlnds + 1))) return rc;
  if (nlnds)
    node_stats_kernel<<<grid_for(nlnds), BS, 0, st>>>(nlnds, dm.bptr, dm.bcol, dm.n2e_ptr, dm.n2e_elem, dm.eptr, dm.etype,
                                                      (c->opts.quirks & FEMX_Q_COH_COUNTS_NODE) ? 1 : 0, dm.bdiag, rbe_cnt, dm.node_count, d_flags + 17);
  {
    size_t tb = 0;
    PCK(cub::DeviceScan::ExclusiveSum(nullptr, tb, rbe_cnt, dm.rbe_ptr, nlnds + 1, st));
    if (tb > cubtmp_bytes) { c->err = "CUB temporary storage exceeds the plan arena"; return FEMX_ERR_ALLOC; }
    PCK(cub::DeviceScan::ExclusiveSum(cubtmp, tb, rbe_cnt, dm.rbe_ptr, nlnds + 1, st));
  }
  // ---- element -> block position (runs while the host fetches bptr / bdiag)
  if ((rc = tracked(c, &dm.bpos, npairs))) return rc;
  if (nconn) bpos_kernel<<<grid_for(nconn), BS, 0, st>>>(nconn, posel, dm.eptr, dm.eb_ptr, dm.conn0, dm.bptr, dm.bcol, dm.bpos);
  if (simplex) simplex_detj_kernel<<<grid_for(nels), BS, 0, st>>>(nels, pdim, dm.eptr, dm.conn0, dm.coords, d_flags + 20);
  int64_t nrbe = 0;
  PCK(cudaMemcpyAsync(h_flags, d_flags, sizeof(h_flags), cudaMemcpyDeviceToHost, st));
  PCK(cudaMemcpyAsync(&nrbe, dm.rbe_ptr + nlnds, sizeof(int64_t), cudaMemcpyDeviceToHost, st));
  P.bptr.resize((size_t)nlnds + 1); P.bdiag.resize(nlnds);
  PCK(cudaMemcpyAsync(P.bptr.data(), dm.bptr, sizeof(int64_t) * ((size_t)nlnds + 1), cudaMemcpyDeviceToHost, st));
  if (nlnds) PCK(cudaMemcpyAsync(P.bdiag.data(), dm.bdiag, sizeof(int32_t) * (size_t)nlnds, cudaMemcpyDeviceToHost, st));
  PCK(cudaStreamSynchronize(st));
  if (h_flags[16] & 1) { c->err = "element node id out of range (expects 1-based local ids)"; return FEMX_ERR_ARG; }
  if (h_flags[17] > 65535) { c->err = "a node couples to more than 65535 nodes"; return FEMX_ERR_UNSUPPORTED; }
  P.max_inc = h_flags[18];
  stage("node statistics, bpos, host bptr");
  // ---- row gather lists of the fused rows kernels
  P.rb_foff = maxnen * d2;
  P.rb_slot = P.rb_foff + pdim;
  if (P.rb_slot % 2 == 0) ++P.rb_slot;   // odd pitch: lane-strided 8-byte stores hit distinct banks
  // the simplex rows kernel scales gradients by sqrt(w detJ mu): admit only positive Jacobians (mu > 0: refresh_materials)
  P.simplex = simplex && !h_flags[20];
  const bool lists = P.max_inc <= 32;
  const bool ri = pdim == 2 && P.max_inc >= 1 && P.max_inc <= 32;
  if ((rc = tracked(c, &dm.rb_off, nblk + nlnds)) || (rc = tracked(c, &dm.rb_ent, nrbe)) || (rc = tracked(c, &dm.sx_ent, P.simplex ? nrbe : 0)) ||
      (rc = tracked(c, &dm.sx_inc, P.simplex ? 4 * nconn : 0)) || (rc = tracked(c, &dm.sx_mat, P.simplex ? nconn : 0)) ||
      (rc = tracked(c, &dm.ri_nodes, ri ? 4 * nconn : 0)) || (rc = tracked(c, &dm.ri_info, ri ? nconn : 0)))
    return rc;
  if (!lists) {
    PCK(cudaMemsetAsync(dm.rb_off, 0, sizeof(uint16_t) * (size_t)std::max<int64_t>(nblk + nlnds, 1), st));
    PCK(cudaMemsetAsync(dm.rb_ent, 0, sizeof(uint16_t) * (size_t)std::max<int64_t>(nrbe, 1), st));
    if (P.simplex) {
      PCK(cudaMemsetAsync(dm.sx_ent, 0, sizeof(uint16_t) * (size_t)std::max<int64_t>(nrbe, 1), st));
      PCK(cudaMemsetAsync(dm.sx_inc, 0, sizeof(int32_t) * 4 * (size_t)std::max<int64_t>(nconn, 1), st));
    }
  } else if (nlnds) {
    row_lists_kernel<<<grid_for(nlnds), BS, 0, st>>>(nlnds, pdim, P.rb_slot, P.simplex ? 1 : 0, dm.bptr, dm.n2e_ptr, dm.n2e_elem, dm.n2e_a, dm.eptr, dm.eb_ptr,
                                                     dm.bpos, dm.rbe_ptr, dm.rb_off, dm.rb_ent, dm.sx_ent);
  }
  if (nconn && ((P.simplex && lists) || ri))
    incidence_records_kernel<<<grid_for(nconn), BS, 0, st>>>(nconn, dm.n2e_elem, dm.n2e_a, dm.eptr, dm.conn0, dm.etype, dm.emat,
                                                             P.simplex && lists ? dm.sx_inc : nullptr, dm.sx_mat, ri ? dm.ri_nodes : nullptr, dm.ri_info);
  stage("row lists");
  // ---- type buckets (ascending element id): stream compaction of the element ids per type present
  c->buckets.clear();
  {
    int32_t *d_all = nullptr;
    int *d_n = nullptr;
    if ((rc = tracked(c, &d_all, nels))) return rc;
    TAKE(d_n, int, 1);
    int64_t at = 0;
    for (int t = FEMX_TRI; t <= FEMX_COH; ++t) {
      const int n = h_flags[8 + t];
      if (!n) continue;
      Bucket b;
      b.type = t; b.n = n; b.d_elem = d_all + at;
      thrust::counting_iterator<int> ids(0);
      size_t tb = 0;
      PCK(cub::DeviceSelect::If(nullptr, tb, ids, b.d_elem, d_n, nels, TypeIs{dm.etype, t}, st));
      if (tb > cubtmp_bytes) { c->err = "CUB temporary storage exceeds the plan arena"; return FEMX_ERR_ALLOC; }
      PCK(cub::DeviceSelect::If(cubtmp, tb, ids, b.d_elem, d_n, nels, TypeIs{dm.etype, t}, st));
      at += n;
      c->buckets.push_back(b);
    }
  }
  stage("type buckets");
  // ---- cohesive overlay (elements with nlMat != 0, src/m_global.F90:130-150, :188-206): O(interface), sorted on the host
  {
    P.coh_elems.assign((size_t)ncoh, 0);
    std::vector<int32_t> rows(4 * (size_t)ncoh);
    std::vector<int64_t> blks(16 * (size_t)ncoh);
    if ((rc = tracked(c, &c->d_coh_elems, ncoh))) return rc;
    if (ncoh) {
      int32_t *d_rows = nullptr;
      int64_t *d_blk = nullptr;
      int *d_n = nullptr;
      TAKE(d_rows, int32_t, rows.size()); TAKE(d_blk, int64_t, blks.size()); TAKE(d_n, int, 1);
      thrust::counting_iterator<int> ids(0);
      size_t tb = 0;
      PCK(cub::DeviceSelect::If(nullptr, tb, ids, c->d_coh_elems, d_n, nels, IsNonlinear{dm.enlmat}, st));
      if (tb > cubtmp_bytes) { c->err = "CUB temporary storage exceeds the plan arena"; return FEMX_ERR_ALLOC; }
      PCK(cub::DeviceSelect::If(cubtmp, tb, ids, c->d_coh_elems, d_n, nels, IsNonlinear{dm.enlmat}, st));
      coh_blocks_kernel<<<grid_for(16 * (int64_t)ncoh), BS, 0, st>>>(ncoh, c->d_coh_elems, dm.eptr, dm.conn0, dm.eb_ptr, dm.bpos, dm.bptr, d_rows, d_blk);
      PCK(cudaMemcpyAsync(P.coh_elems.data(), c->d_coh_elems, sizeof(int32_t) * (size_t)ncoh, cudaMemcpyDeviceToHost, st));
      PCK(cudaMemcpyAsync(rows.data(), d_rows, sizeof(int32_t) * rows.size(), cudaMemcpyDeviceToHost, st));
      PCK(cudaMemcpyAsync(blks.data(), d_blk, sizeof(int64_t) * blks.size(), cudaMemcpyDeviceToHost, st));
      PCK(cudaStreamSynchronize(st));
    }
    struct C { int64_t blk; int32_t row, src; };
    std::vector<C> cs;
    struct N { int32_t node, src; };
    std::vector<N> ns;
    for (int ci = 0; ci < ncoh; ++ci)
      for (int a = 0; a < 4; ++a) {
        const int r = rows[4 * (size_t)ci + a];
        ns.push_back({r, (int32_t)((ci << 4) | a)});
        for (int b = 0; b < 4; ++b) cs.push_back({blks[16 * (size_t)ci + 4 * a + b], r, (int32_t)((ci << 8) | (a << 4) | b)});
      }
    std::stable_sort(cs.begin(), cs.end(), [](const C &x, const C &y) { return x.blk < y.blk; });
    std::stable_sort(ns.begin(), ns.end(), [](const N &x, const N &y) { return x.node < y.node; });
    P.cblk_id.clear(); P.cblk_row.clear(); P.cblk_ptr.assign(1, 0); P.cblk_src.clear();
    for (size_t i = 0; i < cs.size(); ++i) {
      if (i == 0 || cs[i].blk != cs[i - 1].blk) {
        if (i) P.cblk_ptr.push_back((int32_t)i);
        P.cblk_id.push_back(cs[i].blk); P.cblk_row.push_back(cs[i].row);
      }
      P.cblk_src.push_back(cs[i].src);
    }
    if (!cs.empty()) P.cblk_ptr.push_back((int32_t)cs.size());
    P.cnode_id.clear(); P.cnode_ptr.assign(1, 0); P.cnode_src.clear();
    for (size_t i = 0; i < ns.size(); ++i) {
      if (i == 0 || ns[i].node != ns[i - 1].node) {
        if (i) P.cnode_ptr.push_back((int32_t)i);
        P.cnode_id.push_back(ns[i].node);
      }
      P.cnode_src.push_back(ns[i].src);
    }
    if (!ns.empty()) P.cnode_ptr.push_back((int32_t)ns.size());
    UP(c->d_cblk_id, P.cblk_id); UP(c->d_cblk_row, P.cblk_row);
    UP(c->d_cblk_ptr, P.cblk_ptr); UP(c->d_cblk_src, P.cblk_src); UP(c->d_cnode_id, P.cnode_id);
    UP(c->d_cnode_ptr, P.cnode_ptr); UP(c->d_cnode_src, P.cnode_src);
    if ((rc = dev_alloc(c, &c->d_coh_ke, 64 * (int64_t)ncoh))) return rc;
    if ((rc = dev_alloc(c, &c->d_coh_fe, 8 * (int64_t)ncoh))) return rc;
  }
#undef UP
  c->nl2g_identity = c->nlnds == c->nnds;
  for (int i = 0; i < c->nlnds && c->nl2g_identity; ++i) c->nl2g_identity = c->nl2g[i] == i + 1;
  stage("cohesive overlay");
  // ---- hex8 logical grid
  c->hex = HexPlan();
  const int av = c->opts.assembly_variant;
  if (av != 1 && av != 9 && av != 11) {
    if ((rc = hex_plan_device(c, st, A, h_flags[8 + FEMX_HEX] == nels))) return rc;
  }
  PCK(cudaStreamSynchronize(st));
  PCK(cudaGetLastError());
  stage("hex8 logical grid");
  P.build_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  c->host_bcol = false; c->host_bpos = false;
  return FEMX_OK;
}

// host copies of the large plan arrays, fetched when a host-side exporter first needs them
int plan_simplex_localize(femx_ctx *c) {
  femx::Plan &P = c->plan;
  femx::DevMesh &dm = c->dm;
  if (!P.simplex || P.max_inc < 1 || P.max_inc > 32 || !dm.sx_inc || c->nlnds == 0) return FEMX_OK;   // rows_simplex_ok
  const int64_t nconn = c->eptr[c->nels];
  int rc, *d_max = nullptr, h_max = 0;
  if ((rc = dev_alloc(c, &dm.sx_loc, nconn)) || (rc = dev_alloc(c, &d_max, 1))) return rc;
  PCK(cudaMemsetAsync(d_max, 0, sizeof(int), c->stream));
  sx_loc_kernel<<<grid_for(c->nlnds), BS, 0, c->stream>>>(c->nlnds, c->pdim + 1, dm.n2e_ptr, dm.bptr, dm.bcol, dm.sx_inc, dm.sx_loc, d_max);
  c->launches += 1;
  PCK(cudaMemcpyAsync(&h_max, d_max, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  PCK(cudaStreamSynchronize(c->stream));
  PCK(cudaGetLastError());
  P.max_nb = h_max;
  dev_release(c, d_max);
  dev_release(c, dm.sx_inc);
  dm.sx_inc = nullptr;
  return FEMX_OK;
}

int plan_host_blocks(femx_ctx *c) {
  if (c->host_bcol) return FEMX_OK;
  c->plan.bcol.resize((size_t)c->plan.nblk);
  if (c->plan.nblk) PCK(cudaMemcpy(c->plan.bcol.data(), c->dm.bcol, sizeof(int32_t) * (size_t)c->plan.nblk, cudaMemcpyDeviceToHost));
  c->host_bcol = true;
  return FEMX_OK;
}
int plan_host_bpos(femx_ctx *c) {
  if (c->host_bpos) return FEMX_OK;
  const int64_t n = c->plan_npairs;
  c->plan.eb_ptr.resize((size_t)c->nels + 1);
  PCK(cudaMemcpy(c->plan.eb_ptr.data(), c->dm.eb_ptr, sizeof(int64_t) * ((size_t)c->nels + 1), cudaMemcpyDeviceToHost));
  c->plan.bpos.resize((size_t)n);
  if (n) PCK(cudaMemcpy(c->plan.bpos.data(), c->dm.bpos, sizeof(uint16_t) * (size_t)n, cudaMemcpyDeviceToHost));
  c->host_bpos = true;
  return FEMX_OK;
}

}  // namespace femx
