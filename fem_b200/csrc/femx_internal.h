// femx_internal.h -- ctx layout, host plan, launcher prototypes (not installed).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <map>
#include <string>
#include <vector>

#include "../../include/femx.h"

namespace femx {

constexpr int kMaxNen = 8;
__host__ __device__ inline int type_nen(int t) { return t == FEMX_TRI ? 3 : (t == FEMX_HEX ? 8 : 4); }
__host__ __device__ inline int type_nip(int t) {
  return t == FEMX_QUA ? 4 : (t == FEMX_HEX ? 8 : (t == FEMX_COH ? 2 : 1));
}
__host__ __device__ inline int type_dim(int t) { return (t == FEMX_TET || t == FEMX_HEX) ? 3 : 2; }

// PENALTY_PARAM = 1.0e30 is a default-real literal in the reference (src/m_global.F90:15)
constexpr double kPenalty = 1.0000000150474662e30;

// ------------------------------------------------------------------ host plan (plan.cpp)
// One-time integer pass: block (node x node) CSR of the assembled matrix, the
// node -> (element, local node) incidence in ascending element order, and the
// element -> block-position map.  Scalar CSR slots follow in closed form:
//   slot(r,i,k,j) = pd*pd*bptr[r] + i*pd*(bptr[r+1]-bptr[r]) + pd*k + j
struct Plan {
  int pdim = 0, nels = 0, nlnds = 0, nnds = 0;
  std::vector<int32_t> conn0;      // 0-based local nodes, eptr offsets
  std::vector<int64_t> bptr;       // [nlnds+1]
  std::vector<int32_t> bcol;       // [nblk] local node, ascending
  std::vector<int32_t> bdiag;      // [nlnds] position of r in its own block row
  std::vector<int64_t> n2e_ptr;    // [nlnds+1]
  std::vector<int32_t> n2e_elem;   // ascending element per node
  std::vector<uint8_t> n2e_a;      // local node index within that element
  std::vector<int64_t> eb_ptr;     // [nels+1] sum nen^2
  std::vector<uint16_t> bpos;      // block position of (e, a, b) in row conn[e][a]
  std::vector<int64_t> ke_off;     // [nels+1] offsets (doubles) of packed Ke scratch: nen^2*pd^2
  std::vector<int64_t> fe_off;     // [nels+1] offsets of per-element vectors: nen*pd
  int64_t nblk = 0;
  // row gather lists for the fused rows kernel (variant 9): for node r and block k of its row, the contributing
  // (incident element index, local node b) pairs in ascending element order
  std::vector<int64_t> rbe_ptr;    // [nlnds+1] first entry of node r in rb_ent
  std::vector<uint16_t> rb_off;    // node r: nb+1 node-local offsets at [bptr[r] + r ...]
  std::vector<uint16_t> rb_ent;    // shared-memory offset (doubles) of the contributing block: incident index * rb_slot + b * pd^2
  std::vector<uint16_t> sx_ent;    // simplex rows kernel: (incident index << 4) | (a << 2) | b, same order / offsets as rb_ent
  std::vector<int32_t> ri_nodes;   // 2-D meshes, fused rows kernel: per (node, incident element) the element's 0-based nodes padded to 4
  std::vector<int32_t> ri_info;    // ... and type | local node << 4 | 0-based material row << 8 (cohesive: row 0)
  std::vector<int32_t> sx_inc;     // simplex rows kernel: per (node, incident element) the element's 0-based nodes, padded to 4
  std::vector<int32_t> sx_mat;     // ... and its 0-based material row
  bool simplex = false;            // every element is a tri3 / tet4 (one integration point)
  int rb_slot = 0, rb_foff = 0;    // slot pitch per incident element (odd) and offset of its f share inside the slot
  int max_inc = 0;                 // largest number of elements around a node
  int max_nb = 0;                  // widest block row among the rows of the simplex rows kernel (plan_simplex_localize)
  // cohesive overlay: blocks / nodes touched by nonlinear elements, contributions ascending
  std::vector<int32_t> coh_elems;        // local element ids with enlmat != 0
  std::vector<int64_t> cblk_id;          // touched block ids (r-major order)
  std::vector<int32_t> cblk_row;         // row node of each touched block
  std::vector<int32_t> cblk_ptr;         // [ncblk+1]
  std::vector<int32_t> cblk_src;         // (coh index << 8) | (a << 4) | b
  std::vector<int32_t> cnode_id;         // nodes touched by cohesive elements
  std::vector<int32_t> cnode_ptr;        // [ncnode+1]
  std::vector<int32_t> cnode_src;        // (coh index << 4) | a
  double build_seconds = 0.0;
};

// plan_hex.cpp: logical-grid proof + canonical 27-stencil positions for the fused hex8 kernel
struct HexPlan {
  bool ok = false;
  int NI = 0, NJ = 0, NK = 0;       // logical node-grid extents
  std::vector<int32_t> node_at;     // [NK][NJ][NI] local node id or -1
  std::vector<int32_t> elem_at;     // [NK][NJ][NI] local element id or -1 (cell spans i..i+1)
  std::vector<uint8_t> kmap;        // [nlnds][32]: block position of neighbour (dz,dy,dx), 255 = none
  std::string why;                  // reason the mesh does not qualify
};
struct HexRowMeta {                 // 48 bytes, copied to shared memory with six 8-byte cp.async
  long long base;                   // slot of (row i = 0, block position 0) = 9 * bptr[node]
  int nb;                           // blocks in the node's row, 0 = no node at this logical position
  int grow;                         // first global row of the node: 3 * (nl2g - 1)
  unsigned char k[28];              // [dz][dy][dx] -> block position, 255 = absent
};
struct DevHex {
  int NI = 0, NJ = 0, NK = 0;
  double rho = 0.0;
  int32_t *node_at = nullptr, *elem_at = nullptr;
  uint8_t *kmap = nullptr;
  double *coords4 = nullptr;        // [nlnds][4] padded coordinates (128-bit loads)
  double *sqrt_mu = nullptr;        // [nmts]
  double *pos_x = nullptr;          // [NK][NJ][NI][3] coordinates by logical position
  HexRowMeta *pos_meta = nullptr;   // [NK][NJ][NI]
  double *cellmu = nullptr;         // [NK][NJ][NI] sqrt(mu) of the element in that cell, 0 = no element
};
// peer-memory exchange (peer.cu): kernel parameter that tells the assembly kernels where the values of rows another
// rank owns go.  Local value offsets [vb[q], vb[q+1]) (rows ascend with the owner) belong to rank q; vdst[q] is the
// address, in q's receive arena mapped over NVLink, at which offset vb[q] lands (nullptr: keep in the local array).
// fb / fdst: the same for the internal-force entries, numbered pdim * local node + i.
constexpr int kMaxPeers = 8;
struct PeerTab {
  int n = 0, rank = 0;
  long long vb[kMaxPeers + 1] = {};
  long long fb[kMaxPeers + 1] = {};
  double *vdst[kMaxPeers] = {};
  double *fdst[kMaxPeers] = {};
};
// fast path: the value range [lo, hi) / force entry o stays in the local arrays (97 % of the rows)
__device__ __forceinline__ bool peer_local(const PeerTab &pt, long long lo, long long hi) {
  return lo >= pt.vb[pt.rank] && hi <= pt.vb[pt.rank + 1];
}
__device__ __forceinline__ bool peer_local_f(const PeerTab &pt, long long o) { return o >= pt.fb[pt.rank] && o < pt.fb[pt.rank + 1]; }
// address of local value offset `off` (first value of one matrix row); a row never straddles owners
__device__ __forceinline__ double *peer_row(const PeerTab &pt, double *vals, long long off) {
  double *p = vals + off;
#pragma unroll 1
  for (int q = 0; q < pt.n; ++q)
    if (pt.vdst[q] && off >= pt.vb[q] && off < pt.vb[q + 1]) p = pt.vdst[q] + (off - pt.vb[q]);
  return p;
}
__device__ __forceinline__ double *peer_f(const PeerTab &pt, double *local, long long o) {
  double *p = local;
#pragma unroll 1
  for (int q = 0; q < pt.n; ++q)
    if (pt.fdst[q] && o >= pt.fb[q] && o < pt.fb[q + 1]) p = pt.fdst[q] + (o - pt.fb[q]);
  return p;
}

int build_hex_plan(const Plan &P, const int *etype, const int *eptr, HexPlan *H, std::string *why);

int build_plan(int pdim, int nels, const int *etype, const int *eptr, const int *eind,
               const int *enlmat, int nlnds, int nnds, Plan *out, std::string *err);

// ------------------------------------------------------------------ device views
struct Bucket {            // homogeneous element-type bucket, ascending local element id
  int type = 0, n = 0;
  int32_t *d_elem = nullptr;   // [n] local element id
  std::vector<int32_t> h_elem;
};

struct DevMesh {
  int pdim = 0, nels = 0, nlnds = 0, nnds = 0, cp = 0;
  int32_t *etype = nullptr, *eptr = nullptr, *conn0 = nullptr, *emat = nullptr, *enlmat = nullptr;
  int32_t *nl2g = nullptr;     // 1-based global node
  double *coords = nullptr;    // [nlnds][pdim]
  double *lam_mu = nullptr;    // [nmts][2]
  double *cohprops = nullptr;  // [ncoh][6]
  // plan
  int64_t *bptr = nullptr, *n2e_ptr = nullptr, *eb_ptr = nullptr, *ke_off = nullptr, *fe_off = nullptr;
  int32_t *bcol = nullptr, *bdiag = nullptr, *n2e_elem = nullptr;
  uint8_t *n2e_a = nullptr;
  uint16_t *bpos = nullptr;
  int64_t *rbe_ptr = nullptr;
  uint16_t *rb_off = nullptr;
  uint16_t *rb_ent = nullptr;
  uint16_t *sx_ent = nullptr;
  int32_t *sx_inc = nullptr, *sx_mat = nullptr;   // sx_inc: plan intermediate, released once sx_loc exists
  uint32_t *sx_loc = nullptr;      // simplex rows kernel: per (node, incident element) the element's nodes as positions in the node's block row, one byte each
  int32_t *ri_nodes = nullptr, *ri_info = nullptr;
  int32_t *node_count = nullptr;  // stress-average multiplicity per local node
};

// loads of FormRHS (src/m_global.F90:373-436) gathered by destination dof, contributions in the reference's call order
struct DevLoads {
  int ntouched = 0, nfrcs = 0, ntrcs = 0;
  int64_t *dof = nullptr;      // [ntouched] global dof
  int32_t *ptr = nullptr;      // [ntouched+1]
  int32_t *src = nullptr;      // (load index << 3) | (kind << 2) | component; kind 0 = nodal force, 1 = traction
  double *fval = nullptr;      // [nfrcs][pdim+2]
  double *tval = nullptr;      // [ntrcs][pdim+2]
  int32_t *tel = nullptr;      // [ntrcs][2] 0-based local element, 1-based side
};

}  // namespace femx

struct femx_ctx {
  int device = 0;
  femx_opts opts{};
  cudaStream_t stream = nullptr;
  cudaStream_t own_stream = nullptr;   // the stream femx_create made (femx_set_stream may substitute the caller's)
  bool async = false;                 // femx_set_async: calls return without synchronising the stream
  bool ms_pending = false;            // ... the last call's event pair has not been read yet
  cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev_mid = nullptr;   // ev_mid: behind the assembly kernel, before the peer push / barrier / add
  bool mid_valid = false;
  double last_asm_ms = 0.0;
  std::string err;
  int64_t launches = 0;
  double last_ms = 0.0;
  double plan_seconds = 0.0;          // wall time of the last femx_build_plan
  // host copies
  int pdim = 0, nels = 0, nlnds = 0, nnds = 0, nmts = 0, ncoh = 0, nbcs = 0;
  std::vector<int32_t> etype, eptr, eind, emat, enlmat, nl2g;
  std::vector<double> coords, E_nu, cohprops, bcval;
  std::vector<int32_t> bcnode;
  bool has_mesh = false, has_plan = false;
  int64_t plan_npairs = 0;             // sum over elements of nen^2 (device plan; the host plan keeps eb_ptr)
  bool host_bcol = true, host_bpos = true;   // plan.bcol / plan.bpos are present on the host (device plan: fetched on demand, plan_gpu.cu)
  bool mat_ok_simplex = false, mat_ok_hex = false;   // the CURRENT materials admit the simplex rows / hex8 tile kernels (mu > 0, one Poisson ratio)
  std::string mat_why;                                // why not
  int own_nranks = 1, own_rank = 0;            // femx_set_ownership (PETSC_DECIDE row split of this rank)
  int64_t own_row_start = 0, own_row_end = -1;
  bool nl2g_identity = false;   // nl2g[i] == i + 1 for every local node (checked once in femx_build_plan)
  femx::Plan plan;
  femx::HexPlan hex;
  femx::DevHex dhex;
  femx::DevMesh dm;
  std::vector<femx::Bucket> buckets;
  // scratch
  double *d_ke = nullptr;     // packed block-major element matrices
  double *d_fe = nullptr;     // per-element vectors (f_e)
  double *d_sn = nullptr;     // per-element nodal stress [nen][cp]
  double *d_u = nullptr;      // staged u_local
  double *d_vals = nullptr;   // internal values buffer for host-output calls
  double *d_kstage = nullptr; // staged K values (host-input femx_jacobian)
  std::map<double **, int64_t> buf_n;  // capacity of the grow-on-demand buffers above
  double *d_vec[4] = {nullptr, nullptr, nullptr, nullptr};  // global-length staging vectors
  int64_t nnz = 0;
  // BC / cohesive device lists
  int32_t *d_bc_slot_row = nullptr;  // unused placeholder
  int64_t *d_bc_diag = nullptr;      // diag slot per constrained dof
  int64_t *d_bc_dof = nullptr;       // global dof per constrained entry
  double *d_bc_val = nullptr;
  int nbc_dofs = 0;
  int32_t *d_coh_elems = nullptr, *d_cblk_row = nullptr, *d_cblk_ptr = nullptr, *d_cblk_src = nullptr;
  int64_t *d_cblk_id = nullptr;
  int32_t *d_cnode_id = nullptr, *d_cnode_ptr = nullptr, *d_cnode_src = nullptr;
  double *d_coh_ke = nullptr, *d_coh_fe = nullptr;
  femx::DevLoads loads;
  std::vector<void *> allocs;  // everything cudaMalloc'ed by the ctx
  std::vector<void *> plan_allocs;   // ... of which the current plan's arrays (released by the next femx_build_plan)
  // peer-memory exchange (peer.cu)
  struct Peer {
    int nranks = 0, rank = 0;
    char *arena = nullptr;                 // [flags 1 KB | buffer 0: vals, f | buffer 1: vals, f]
    int64_t nv = 0, nf = 0;                // receive capacity (doubles) per buffer
    int64_t nu = 0;                        // ghost-displacement receive area (doubles per buffer, indexed by local dof): pdim * nlnds
    int64_t rnu[femx::kMaxPeers] = {};     // peers' nu
    unsigned long long ustep = 0;          // femx_scatter_u calls done; u-area parity = ustep & 1
    int64_t *d_sptr = nullptr, *d_srow = nullptr;   // femx_peer_set_scatter: per peer the owned global rows whose u it needs
    int64_t udst[femx::kMaxPeers] = {};    // first local dof (in the peer's numbering) of the block this rank sends it
    int64_t n_scatter = 0;
    bool upending = false; double *pend_ulocal = nullptr; const double *pend_uowned = nullptr;
    char *remote[femx::kMaxPeers] = {};    // peers' arenas mapped with cudaIpcOpenMemHandle
    int64_t rnv[femx::kMaxPeers] = {}, rnf[femx::kMaxPeers] = {};   // peers' capacities (their buffer layout)
    int64_t vb[femx::kMaxPeers + 1] = {}, fb[femx::kMaxPeers + 1] = {};
    int64_t vdst[femx::kMaxPeers] = {}, fdst[femx::kMaxPeers] = {};  // offsets (doubles) inside the peer's vals / f receive area
    unsigned long long step = 0;           // assemblies done; buffer parity = step & 1
    int *d_err = nullptr;
    bool connected = false;
    bool local = false;                    // peers are ctxs of this process on the same device (femx_peer_connect_local): no IPC mapping
    bool in_kernel = false;                // this assembly's kernel stores the interface rows itself (PEER instantiation)
    unsigned nbr_mask = 0;                 // ranks this one exchanges a non-empty slice with (either direction): the step barrier's scope
    bool deferred = false;                 // assemblies signal but do not wait / merge; femx_peer_complete does (ranks driven in turn on one GPU)
    bool pending = false;                  // ... and one is outstanding
    double *pend_vals = nullptr, *pend_fint = nullptr;
    // receiver add over the interface rows (femx_peer_set_merge)
    int64_t n_merge = 0, n_fmerge = 0;
    int64_t *d_mown = nullptr, *d_mptr = nullptr, *d_midx = nullptr, *d_frow = nullptr, *d_fptr = nullptr, *d_fidx = nullptr;
    int32_t *d_mown32 = nullptr, *d_mptr32 = nullptr, *d_midx32 = nullptr;   // the same lists in 32 bits when everything fits
    double *d_merged = nullptr;
    // exchange overlapped with the assembly (hex8 tile kernel): the CTAs that own interface rows run first on a high-priority
    // stream, followed there by push, step barrier and receiver add, while the interior CTAs run on the ctx stream
    int *d_cta_list = nullptr;             // linear CTA ids: [0, n_cta_if) own an interface row, [n_cta_if, n_cta) do not
    int n_cta_if = 0, n_cta = 0;
    bool split_valid = false;
    cudaStream_t hi = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  } peer;
};
