// peer.cu -- interface-row exchange through peer memory (replaces MatAssemblyBegin/End's stash
// traffic, src/main.F90:167-168, and the NCCL all_to_all that followed the kernel in the first cut).
//
// Every rank owns a receive arena (cudaMalloc, exported with cudaIpcGetMemHandle, mapped by its peers over
// NVLink / NVSwitch).  Because the rank-local CSR lists rows in ascending global order, the values bound for
// rank q are ONE contiguous range of local value offsets (likewise the internal-force entries in local dof
// order).  Default: peer_push_kernel, launched behind the assembly kernel on the same stream, copies those
// ranges into the owners' arenas with coalesced remote stores and its last CTA runs the step barrier -- this
// rank's step counter is raised in every peer's arena (system-scope release) and the peers' counters are
// awaited (acquire): when femx_assemble_* returns, the rank's own slice and everything its peers sent are
// complete.  Opt-in (opts.reserved[3] = 1): the PEER instantiations of hex_sym2_kernel / fused_rows_kernel /
// rows_simplex_kernel look a row's first offset up in the PeerTab and store it remotely themselves; only the
// barrier kernel follows (measured slower).
// Receive areas are double-buffered on the step parity, so a peer that is one assembly ahead never
// overwrites values the host solver may still be reading; the step barrier keeps it from getting two ahead.
#include <cstdlib>

#include "api_util.h"

namespace {

constexpr int64_t kFlagBytes = 1024;   // kMaxPeers step counters, 64 B apart, + spare

__device__ __forceinline__ void st_release_sys(unsigned long long *p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long *p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

struct FlagPtrs { unsigned long long *remote[kMaxPeers]; };

// lane q: tell NEIGHBOUR rank q (a rank this one exchanges a non-empty slice with, in either direction) that this rank's
// stores of assembly `step` are done, then wait for q's.  Ranks that share no row are neither signalled nor awaited: a
// step costs the slowest neighbour, not the slowest of all ranks.  wait = false: signal only (the caller waits later,
// femx_peer_complete -- used when several ranks are driven one after another on one GPU).
__device__ __forceinline__ void signal_wait(const FlagPtrs &fp, unsigned long long *mine, int n, int rank, unsigned nbr_mask, unsigned long long step,
                                            long long limit_clocks, int *err, bool sig, bool wait) {
  const int q = threadIdx.x;
  if (q >= n || q == rank || !((nbr_mask >> q) & 1u)) return;
  if (sig) {
    __threadfence_system();
    st_release_sys(fp.remote[q] + 8 * rank, step);
  }
  if (!wait) return;
  const long long t0 = clock64();
  while (ld_acquire_sys(mine + 8 * q) < step) {
    if (clock64() - t0 > limit_clocks) { *err = 1 + q; break; }   // a dead peer must not hang the GPU
    __nanosleep(200);
  }
}

__global__ void peer_signal_wait_kernel(FlagPtrs fp, unsigned long long *mine, int n, int rank, unsigned nbr_mask, unsigned long long step,
                                        long long limit_clocks, int *err, bool sig, bool wait) {
  signal_wait(fp, mine, n, rank, nbr_mask, step, limit_clocks, err, sig, wait);
}

// Push form (works behind every assembly variant): the kernel wrote all of its rows into the local arrays; copy the
// ranges other ranks own into their arenas with remote stores (coalesced 8-byte lanes over NVLink), then the last CTA
// to finish runs the step barrier.  fint is indexed by global row, the wire format by local dof (pdim * node + i).
__global__ void __launch_bounds__(256)
peer_push_kernel(PeerTab pt, const double *__restrict__ vals, const double *__restrict__ fint, const int32_t *__restrict__ nl2g, int pdim,
                 FlagPtrs fp, unsigned long long *mine, unsigned nbr_mask, unsigned long long step, long long limit_clocks, int *err, unsigned int *ticket,
                 bool wait) {
  const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x, nth = (long long)gridDim.x * blockDim.x;
  for (int q = 0; q < pt.n; ++q) {
    if (q == pt.rank) continue;
    if (pt.vdst[q]) {
      const double *src = vals + pt.vb[q];
      double *dst = pt.vdst[q];
      const long long len = pt.vb[q + 1] - pt.vb[q];
      long long i = tid;
      for (; i + 3 * nth < len; i += 4 * nth) {   // four independent loads in flight per thread: the copy runs on few CTAs when it overlaps the tile kernel
        const double a0 = src[i], a1 = src[i + nth], a2 = src[i + 2 * nth], a3 = src[i + 3 * nth];
        dst[i] = a0; dst[i + nth] = a1; dst[i + 2 * nth] = a2; dst[i + 3 * nth] = a3;
      }
      for (; i < len; i += nth) dst[i] = src[i];
    }
    if (fint && pt.fdst[q]) {
      double *dst = pt.fdst[q];
      for (long long o = pt.fb[q] + tid; o < pt.fb[q + 1]; o += nth) {
        const long long r = o / pdim;
        dst[o - pt.fb[q]] = fint[(long long)pdim * (nl2g[r] - 1) + (o - r * pdim)];
      }
    }
  }
  __shared__ bool last;
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) last = atomicAdd(ticket, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!last) return;
  if (threadIdx.x == 0) *ticket = 0;
  signal_wait(fp, mine, pt.n, pt.rank, nbr_mask, step, limit_clocks, err, true, wait);
}

// Receiver add (the part of MatAssemblyEnd that sums the stash into the owned rows, src/main.F90:167-168), interface rows
// only: entry s of the merged interface rows = this rank's own value (slot own[s] of the local CSR, -1 = the rank has no
// such entry) + the received values idx[ptr[s]..ptr[s+1]) in list order = (source rank, sender position): deterministic.
template <class I>
__global__ void peer_merge_kernel(const double *__restrict__ vals, const double *__restrict__ recv, const I *__restrict__ own,
                                  const I *__restrict__ ptr, const I *__restrict__ idx, int64_t n, double *__restrict__ out) {
  const int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n) return;
  const I o = own[s], q0 = ptr[s], q1 = ptr[s + 1];
  double v = o >= 0 ? vals[o] : 0.0;
  for (I q = q0; q < q1; ++q) v += recv[idx[q]];
  out[s] = v;
}
// owned internal-force entries: fint[row[s]] += received entries in (source rank, position) order
__global__ void peer_merge_f_kernel(double *__restrict__ fint, const double *__restrict__ recv_f, const int64_t *__restrict__ row,
                                    const int64_t *__restrict__ ptr, const int64_t *__restrict__ idx, int64_t n) {
  const int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n) return;
  double v = fint[row[s]];
  for (int64_t q = ptr[s]; q < ptr[s + 1]; ++q) v += recv_f[idx[q]];
  fint[row[s]] = v;
}

// GetVec_U across ranks (src/m_global.F90:549-564: VecScatter global -> ghosted local): every owner stores the displacement
// entries its peers hold as ghosts straight into their u areas (indexed by the peer's local dof), then signals.
struct ScatterDst { double *u[kMaxPeers]; };
__global__ void __launch_bounds__(256)
peer_scatter_push_kernel(int nranks, int rank, const int64_t *__restrict__ sptr, const int64_t *__restrict__ srow, const double *__restrict__ u_owned,
                         long long row_start, ScatterDst dst, FlagPtrs fp, unsigned long long *mine, unsigned nbr_mask, unsigned long long step,
                         long long limit_clocks, int *err, unsigned int *ticket, bool wait) {
  const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x, nth = (long long)gridDim.x * blockDim.x;
  for (int q = 0; q < nranks; ++q) {
    if (q == rank || !dst.u[q]) continue;
    const long long n = sptr[q + 1] - sptr[q];
    const int64_t *rows = srow + sptr[q];
    for (long long i = tid; i < n; i += nth) dst.u[q][i] = u_owned[rows[i] - row_start];
  }
  __shared__ bool last;
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) last = atomicAdd(ticket, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!last) return;
  if (threadIdx.x == 0) *ticket = 0;
  signal_wait(fp, mine, nranks, rank, nbr_mask, step, limit_clocks, err, true, wait);
}
// the rank's own dofs from its slice of the global vector, the ghosts from the u area the owners filled
__global__ void peer_scatter_local_kernel(double *__restrict__ u_local, const double *__restrict__ u_owned, const double *__restrict__ uarea,
                                          const int32_t *__restrict__ nl2g, int pdim, long long n, long long own_lo, long long own_hi, long long row_start) {
  const long long ld = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (ld >= n) return;
  if (ld >= own_lo && ld < own_hi) {
    const long long node = ld / pdim;
    u_local[ld] = u_owned[(long long)pdim * (nl2g[node] - 1) + (ld - node * pdim) - row_start];
  } else {
    u_local[ld] = uarea[ld];
  }
}

int64_t buffer_doubles(int64_t nv, int64_t nf) { return nv + nf; }

}  // namespace

namespace femx {

// the redirect table of the assembly that is about to be launched (buffer parity of the current step)
PeerTab peer_tab(const femx_ctx *c) {
  PeerTab t;
  const auto &P = c->peer;
  if (!P.connected) return t;
  t.n = P.nranks; t.rank = P.rank;
  const int par = (int)(P.step & 1);
  for (int q = 0; q <= P.nranks; ++q) { t.vb[q] = P.vb[q]; t.fb[q] = P.fb[q]; }
  for (int q = 0; q < P.nranks; ++q) {
    if (q == P.rank || !P.remote[q]) continue;
    double *buf = reinterpret_cast<double *>(P.remote[q] + kFlagBytes) + par * buffer_doubles(P.rnv[q], P.rnf[q]);
    t.vdst[q] = buf + P.vdst[q];
    t.fdst[q] = buf + P.rnv[q] + P.fdst[q];
  }
  return t;
}

static int launch_peer_merge(femx_ctx *c, const double *d_vals, double *d_fint) {
  auto &P = c->peer;
  int n = 0;
  const int par = (int)((P.step - 1) & 1);   // the buffer the assembly that just ran filled
  const double *buf = reinterpret_cast<const double *>(P.arena + kFlagBytes) + par * buffer_doubles(P.nv, P.nf);
  if (P.n_merge > 0) {
    // 32-bit lists when every index fits (femx_peer_set_merge): the add moves 32 instead of 48 bytes per merged value
    if (P.d_mown32) peer_merge_kernel<int32_t><<<(unsigned)((P.n_merge + 255) / 256), 256, 0, c->stream>>>(d_vals, buf, P.d_mown32, P.d_mptr32, P.d_midx32, P.n_merge, P.d_merged);
    else peer_merge_kernel<int64_t><<<(unsigned)((P.n_merge + 255) / 256), 256, 0, c->stream>>>(d_vals, buf, P.d_mown, P.d_mptr, P.d_midx, P.n_merge, P.d_merged);
    ++n;
  }
  if (P.n_fmerge > 0 && d_fint) {
    peer_merge_f_kernel<<<(unsigned)((P.n_fmerge + 255) / 256), 256, 0, c->stream>>>(d_fint, buf + P.nv, P.d_frow, P.d_fptr, P.d_fidx, P.n_fmerge);
    ++n;
  }
  return n;
}

// ---- interface-first order of the hex8 tile kernel --------------------------------------------------------------
// An interface node is one that another rank also holds: its row is sent (the local dof lies in a peer's f_bounds range)
// or receives (it is named by the force merge list).  A CTA of the tile kernel that owns such a row runs in the first,
// high-priority launch; push, step barrier and receiver add follow on that stream while the other CTAs still assemble.
__global__ void mark_recv_nodes_kernel(int64_t nf, int pdim, const int64_t *__restrict__ f_row, unsigned char *__restrict__ markg) {
  const int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s < nf) markg[f_row[s] / pdim] = 1;
}
__global__ void mark_cta_kernel(int NI, int NJ, int NK, int nxt, int nyt, int lz, int gx, int gy, const int32_t *__restrict__ node_at,
                                const int32_t *__restrict__ nl2g, int pdim, long long own_lo, long long own_hi,
                                const unsigned char *__restrict__ markg, int *__restrict__ flag) {
  const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= (int64_t)NI * NJ * NK) return;
  const int r = node_at[q];
  if (r < 0) return;
  const long long dof = (long long)pdim * r;
  if (dof >= own_lo && dof < own_hi && !markg[nl2g[r] - 1]) return;   // owned here and nobody adds to it
  const int i = (int)(q % NI), j = (int)((q / NI) % NJ), k = (int)(q / ((int64_t)NI * NJ));
  flag[((k / lz) * gy + j / nyt) * gx + i / nxt] = 1;
}

int peer_hex_split(femx_ctx *c) {
  auto &P = c->peer;
  if (P.split_valid) return FEMX_OK;
  CK(cudaSetDevice(c->device));
  int gx, gy, gz, nxt, nyt, lz;
  hex_mma_grid(c, &gx, &gy, &gz, &nxt, &nyt, &lz);
  const int n = gx * gy * gz;
  const DevHex &hx = c->dhex;
  unsigned char *d_mark = nullptr;
  int *d_flag = nullptr, rc;
  if ((rc = dev_alloc(c, &d_mark, (int64_t)c->nnds)) || (rc = dev_alloc(c, &d_flag, n))) return rc;
  CK(cudaMemsetAsync(d_mark, 0, (size_t)c->nnds, c->stream));
  CK(cudaMemsetAsync(d_flag, 0, sizeof(int) * (size_t)n, c->stream));
  if (P.n_fmerge > 0)
    mark_recv_nodes_kernel<<<(unsigned)((P.n_fmerge + 255) / 256), 256, 0, c->stream>>>(P.n_fmerge, c->pdim, P.d_frow, d_mark);
  const int64_t npos = (int64_t)hx.NI * hx.NJ * hx.NK;
  mark_cta_kernel<<<(unsigned)((npos + 255) / 256), 256, 0, c->stream>>>(hx.NI, hx.NJ, hx.NK, nxt, nyt, lz, gx, gy, hx.node_at, c->dm.nl2g, c->pdim,
                                                                       (long long)P.fb[P.rank], (long long)P.fb[P.rank + 1], d_mark, d_flag);
  c->launches += 2;
  std::vector<int> flag((size_t)n), list;
  CK(cudaMemcpyAsync(flag.data(), d_flag, sizeof(int) * (size_t)n, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  CK(cudaGetLastError());
  dev_release(c, d_mark); dev_release(c, d_flag);
  list.reserve((size_t)n);
  for (int id = 0; id < n; ++id) if (flag[(size_t)id]) list.push_back(id);
  P.n_cta_if = (int)list.size();
  for (int id = 0; id < n; ++id) if (!flag[(size_t)id]) list.push_back(id);
  P.n_cta = n;
  if (P.d_cta_list) { dev_release(c, P.d_cta_list); P.d_cta_list = nullptr; }
  if ((rc = upload(c, &P.d_cta_list, list))) return rc;
  CK(cudaStreamSynchronize(c->stream));
  if (!P.hi) {
    int least = 0, greatest = 0;
    CK(cudaDeviceGetStreamPriorityRange(&least, &greatest));
    CK(cudaStreamCreateWithPriority(&P.hi, cudaStreamNonBlocking, greatest));
    CK(cudaEventCreateWithFlags(&P.ev_fork, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&P.ev_join, cudaEventDisableTiming));
  }
  P.split_valid = true;
  if (getenv("FEMX_PEER_TIMING"))
    fprintf(stderr, "femx peer (rank %d): %d of the tile kernel's %d CTAs own interface rows and run first\n", P.rank, P.n_cta_if, P.n_cta);
  return FEMX_OK;
}

// fused = the assembly kernel already stored the interface rows at their owners (PEER instantiations)
int launch_peer_finish(femx_ctx *c, bool fused, const double *d_vals, double *d_fint) {
  auto &P = c->peer;
  FlagPtrs fp{};
  for (int q = 0; q < P.nranks; ++q) fp.remote[q] = reinterpret_cast<unsigned long long *>(P.remote[q]);
  unsigned long long *mine = reinterpret_cast<unsigned long long *>(P.arena);
  const long long limit = 20000000000ll;   // ~10 s
  const bool wait = !P.deferred;
  int n = 1;
  // FEMX_PEER_TIMING=1: device time of the push (+ barrier) and of the receiver add, averaged, to stderr every 20 assemblies
  static const bool timing = getenv("FEMX_PEER_TIMING") != nullptr;
  static cudaEvent_t te[3] = {nullptr, nullptr, nullptr};
  static double tacc[2] = {0.0, 0.0};
  static int tn = 0;
  if (timing) {
    if (!te[0]) for (auto &e : te) cudaEventCreate(&e);
    cudaEventRecord(te[0], c->stream);
  }
  if (fused) {
    ++P.step;
    peer_signal_wait_kernel<<<1, 32, 0, c->stream>>>(fp, mine, P.nranks, P.rank, P.nbr_mask, P.step, limit, P.d_err, true, wait);
  } else {
    const PeerTab pt = peer_tab(c);   // destinations of the current step's buffer
    ++P.step;
    int64_t nv = 0;
    for (int q = 0; q < P.nranks; ++q) if (q != P.rank) nv += P.vb[q + 1] - P.vb[q];
    // behind the assembly: two CTAs per SM.  Overlapped with the interior tiles (the stream is the high-priority one): the
    // copy is off the critical path, and every SM it touches is one the tile kernel (one CTA = a whole SM) cannot use meanwhile,
    // so it is confined to a few SMs' worth of CTAs (FEMX_PEER_PUSH_CTAS, default 64 = 8 SMs)
    static const int ov_ctas = getenv("FEMX_PEER_PUSH_CTAS") ? std::max(1, atoi(getenv("FEMX_PEER_PUSH_CTAS"))) : 64;
    const int64_t cap = (P.hi && c->stream == P.hi) ? ov_ctas : 296;
    int grid = (int)std::min<int64_t>(cap, std::max<int64_t>(1, (nv + 255) / 256 / 8));
    peer_push_kernel<<<grid, 256, 0, c->stream>>>(pt, d_vals, d_fint, c->dm.nl2g, c->pdim, fp, mine, P.nbr_mask, P.step, limit, P.d_err,
                                                  reinterpret_cast<unsigned int *>(P.d_err + 1), wait);
  }
  if (timing) cudaEventRecord(te[1], c->stream);
  if (wait) n += launch_peer_merge(c, d_vals, d_fint);
  else { P.pending = true; P.pend_vals = const_cast<double *>(d_vals); P.pend_fint = d_fint; }
  if (timing) {
    cudaEventRecord(te[2], c->stream);
    cudaEventSynchronize(te[2]);
    float a = 0.f, b = 0.f;
    cudaEventElapsedTime(&a, te[0], te[1]); cudaEventElapsedTime(&b, te[1], te[2]);
    tacc[0] += a; tacc[1] += b;
    if (++tn % 20 == 0) {
      fprintf(stderr, "femx peer timing (rank %d, %d assemblies): push + barrier %.4f ms, receiver add %.4f ms\n", P.rank, tn, tacc[0] / 20, tacc[1] / 20);
      tacc[0] = tacc[1] = 0.0;
    }
  }
  return n;
}

}  // namespace femx

extern "C" {

static int launch_scatter_local(femx_ctx *c, const double *d_uown, double *d_uloc);

int femx_peer_arena_create(femx_ctx *c, int64_t n_recv_vals, int64_t n_recv_f, char handle[64]) {
  if (!c) return FEMX_ERR_ARG;
  ARG(n_recv_vals >= 0 && n_recv_f >= 0 && handle, "bad arguments");
  ARG(!c->peer.arena, "the receive arena exists already");
  ARG(c->has_mesh, "femx_set_mesh first");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size");
  CK(cudaSetDevice(c->device));
  auto &P = c->peer;
  P.nv = n_recv_vals; P.nf = n_recv_f;
  P.nu = (int64_t)c->pdim * c->nlnds;      // ghost displacements (femx_scatter_u) land at their local dof
  const size_t bytes = (size_t)kFlagBytes + 2 * sizeof(double) * (size_t)(buffer_doubles(P.nv, P.nf) + P.nu);
  void *p = nullptr;
  CK(cudaMalloc(&p, bytes));
  P.arena = (char *)p;
  CK(cudaMemset(p, 0, bytes));
  int rc = dev_alloc(c, &P.d_err, 2);   // [error flag, CTA ticket of the push kernel]
  if (rc) return rc;
  CK(cudaMemset(P.d_err, 0, 2 * sizeof(int)));
  CK(cudaDeviceSynchronize());
  cudaIpcMemHandle_t h;
  CK(cudaIpcGetMemHandle(&h, p));
  memcpy(handle, &h, 64);
  return FEMX_OK;
}

static int peer_connect_common(femx_ctx *c, int nranks, int rank, const int64_t *recv_vals, const int64_t *recv_f, const int64_t *val_bounds,
                               const int64_t *val_dst, const int64_t *f_bounds, const int64_t *f_dst, const int64_t *recv_from) {
  ARG(c->has_plan, "femx_build_plan first");
  ARG(c->peer.arena, "femx_peer_arena_create first");
  ARG(!c->peer.connected, "already connected (femx_peer_disconnect first)");
  ARG(nranks >= 2 && nranks <= kMaxPeers && rank >= 0 && rank < nranks, "peer exchange handles 2..8 ranks of one node");
  ARG(recv_vals && recv_f && val_bounds && val_dst && f_bounds && f_dst && recv_from, "NULL argument");
  ARG(c->own_row_end < 0 || (c->own_nranks == nranks && c->own_rank == rank), "femx_set_ownership named another rank / rank count");
  ARG(val_bounds[0] == 0 && val_bounds[nranks] == c->nnz, "val_bounds must span the local values");
  ARG(f_bounds[0] == 0 && f_bounds[nranks] == (int64_t)c->pdim * c->nlnds, "f_bounds must span the local dofs");
  auto &P = c->peer;
  P.nranks = nranks; P.rank = rank;
  P.nbr_mask = 0;
  P.split_valid = false;
  for (int q = 0; q <= nranks; ++q) { P.vb[q] = val_bounds[q]; P.fb[q] = f_bounds[q]; }
  for (int q = 0; q < nranks; ++q) {
    ARG(val_bounds[q] <= val_bounds[q + 1] && f_bounds[q] <= f_bounds[q + 1], "bounds must ascend");
    P.rnv[q] = recv_vals[q]; P.rnf[q] = recv_f[q];
    P.vdst[q] = val_dst[q]; P.fdst[q] = f_dst[q];
    if (q == rank) continue;
    ARG(val_dst[q] >= 0 && val_dst[q] + (val_bounds[q + 1] - val_bounds[q]) <= recv_vals[q], "value slice does not fit the peer's arena");
    ARG(f_dst[q] >= 0 && f_dst[q] + (f_bounds[q + 1] - f_bounds[q]) <= recv_f[q], "force slice does not fit the peer's arena");
    // neighbours: a non-empty slice in either direction (the relation is symmetric: q computes the same from its side)
    if (val_bounds[q + 1] > val_bounds[q] || f_bounds[q + 1] > f_bounds[q] || recv_from[q] > 0) P.nbr_mask |= 1u << q;
  }
  return FEMX_OK;
}

int femx_peer_connect(femx_ctx *c, int nranks, int rank, const char *handles, const int64_t *recv_vals, const int64_t *recv_f,
                      const int64_t *val_bounds, const int64_t *val_dst, const int64_t *f_bounds, const int64_t *f_dst,
                      const int64_t *recv_from) {
  if (!c) return FEMX_ERR_ARG;
  ARG(handles, "NULL argument");
  int rc = peer_connect_common(c, nranks, rank, recv_vals, recv_f, val_bounds, val_dst, f_bounds, f_dst, recv_from);
  if (rc) return rc;
  CK(cudaSetDevice(c->device));
  auto &P = c->peer;
  for (int q = 0; q < nranks; ++q) {
    if (q == rank) continue;
    cudaIpcMemHandle_t h;
    memcpy(&h, handles + 64 * (size_t)q, 64);
    void *p = nullptr;
    CK(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    P.remote[q] = (char *)p;
  }
  // the step counter is NOT reset: the arenas' flag words keep the steps of earlier connections, and every rank has run
  // the same number of (collective) assemblies
  P.local = false;
  P.connected = true;
  return FEMX_OK;
}

// Ranks that are ctxs of ONE process on ONE device (tests on a single GPU; a host that drives several partitions per
// GPU): the peers' arenas are plain device pointers.  Such ranks run one after another on the device, so a kernel must
// never wait for a peer that has not been launched yet: assemblies only push and signal (deferred = 1), and
// femx_peer_complete -- called on every rank after all of them have assembled -- checks the flags and runs the receiver add.
int femx_peer_connect_local(femx_ctx *c, int nranks, int rank, femx_ctx *const *peers, const int64_t *recv_vals, const int64_t *recv_f,
                            const int64_t *val_bounds, const int64_t *val_dst, const int64_t *f_bounds, const int64_t *f_dst,
                            const int64_t *recv_from) {
  if (!c) return FEMX_ERR_ARG;
  ARG(peers, "NULL argument");
  int rc = peer_connect_common(c, nranks, rank, recv_vals, recv_f, val_bounds, val_dst, f_bounds, f_dst, recv_from);
  if (rc) return rc;
  auto &P = c->peer;
  for (int q = 0; q < nranks; ++q) {
    if (q == rank) continue;
    ARG(peers[q] && peers[q]->peer.arena && peers[q]->device == c->device, "local peers must be ctxs with an arena on the same device");
    ARG(peers[q]->peer.nv == recv_vals[q] && peers[q]->peer.nf == recv_f[q], "peer arena capacity mismatch");
    P.remote[q] = peers[q]->peer.arena;
  }
  P.local = true;
  P.deferred = true;
  P.connected = true;
  return FEMX_OK;
}

int femx_peer_set_merge(femx_ctx *c, int64_t n, const int64_t *own_slot, const int64_t *ptr, const int64_t *idx, int64_t nf,
                        const int64_t *f_row, const int64_t *f_ptr, const int64_t *f_idx) {
  if (!c) return FEMX_ERR_ARG;
  ARG(c->peer.arena, "femx_peer_arena_create first");
  ARG(n >= 0 && nf >= 0 && (n == 0 || (own_slot && ptr && idx)) && (nf == 0 || (f_row && f_ptr && f_idx)), "bad merge lists");
  CK(cudaSetDevice(c->device));
  auto &P = c->peer;
  for (int64_t s = 0; s < n; ++s) ARG(own_slot[s] >= -1 && own_slot[s] < c->nnz && ptr[s] <= ptr[s + 1], "merge list: bad own slot / ptr");
  for (int64_t q = 0; q < (n ? ptr[n] : 0); ++q) ARG(idx[q] >= 0 && idx[q] < P.nv, "merge list: received index outside the arena");
  for (int64_t s = 0; s < nf; ++s) ARG(f_row[s] >= 0 && f_row[s] < (int64_t)c->pdim * c->nnds && f_ptr[s] <= f_ptr[s + 1], "force merge list: bad row / ptr");
  for (int64_t q = 0; q < (nf ? f_ptr[nf] : 0); ++q) ARG(f_idx[q] >= 0 && f_idx[q] < P.nf, "force merge list: received index outside the arena");
  for (void *q : {(void *)P.d_mown, (void *)P.d_mptr, (void *)P.d_midx, (void *)P.d_frow, (void *)P.d_fptr, (void *)P.d_fidx, (void *)P.d_merged,
                  (void *)P.d_mown32, (void *)P.d_mptr32, (void *)P.d_midx32}) dev_release(c, q);
  P.d_mown = P.d_mptr = P.d_midx = P.d_frow = P.d_fptr = P.d_fidx = nullptr; P.d_merged = nullptr;
  P.d_mown32 = P.d_mptr32 = P.d_midx32 = nullptr;
  P.n_merge = n; P.n_fmerge = nf;
  P.split_valid = false;
  int rc;
  if (n > 0) {
    if (c->nnz < 2147483647ll && P.nv < 2147483647ll && ptr[n] < 2147483647ll) {   // every slot, arena position and list offset fits 32 bits
      if ((rc = upload(c, &P.d_mown32, std::vector<int32_t>(own_slot, own_slot + n))) || (rc = upload(c, &P.d_mptr32, std::vector<int32_t>(ptr, ptr + n + 1))) ||
          (rc = upload(c, &P.d_midx32, std::vector<int32_t>(idx, idx + ptr[n]))))
        return rc;
    } else if ((rc = upload(c, &P.d_mown, std::vector<int64_t>(own_slot, own_slot + n))) || (rc = upload(c, &P.d_mptr, std::vector<int64_t>(ptr, ptr + n + 1))) ||
               (rc = upload(c, &P.d_midx, std::vector<int64_t>(idx, idx + ptr[n]))))
      return rc;
    if ((rc = dev_alloc(c, &P.d_merged, n))) return rc;
  }
  if (nf > 0) {
    if ((rc = upload(c, &P.d_frow, std::vector<int64_t>(f_row, f_row + nf))) || (rc = upload(c, &P.d_fptr, std::vector<int64_t>(f_ptr, f_ptr + nf + 1))) ||
        (rc = upload(c, &P.d_fidx, std::vector<int64_t>(f_idx, f_idx + f_ptr[nf]))))
      return rc;
  }
  CK(cudaStreamSynchronize(c->stream));
  return FEMX_OK;
}

int femx_peer_merged(femx_ctx *c, const double **merged, int64_t *n) {
  if (!c) return FEMX_ERR_ARG;
  if (merged) *merged = c->peer.d_merged;
  if (n) *n = c->peer.n_merge;
  return FEMX_OK;
}

static int peer_check_error(femx_ctx *c) {
  int e = 0;
  CK(cudaMemcpy(&e, c->peer.d_err, sizeof(int), cudaMemcpyDeviceToHost));
  if (e) {
    CK(cudaMemset(c->peer.d_err, 0, sizeof(int)));   // reported once, not sticky
    c->err = "peer exchange: rank " + std::to_string(e - 1) + " did not reach the step barrier";
    return FEMX_ERR_CUDA;
  }
  return FEMX_OK;
}

// Several PROCESSES that share one GPU (a host that runs more MPI ranks than it has GPUs) are in the position of local
// ranks: their kernels are time-sliced, so a step barrier spun on the device waits for a context switch.  Deferred mode
// on an IPC connection: the assembly pushes and signals, the host synchronises its stream and meets the other ranks
// (MPI_Barrier), femx_peer_complete checks the flags and runs the receiver add.
int femx_peer_set_deferred(femx_ctx *c, int on) {
  if (!c) return FEMX_ERR_ARG;
  auto &P = c->peer;
  ARG(P.connected && !P.local, "femx_peer_set_deferred belongs to femx_peer_connect connections (local peers are always deferred)");
  ARG(!P.pending && !P.upending, "an assembly / scatter is outstanding (femx_peer_complete)");
  P.deferred = on != 0;
  return FEMX_OK;
}

int femx_peer_complete(femx_ctx *c) {
  if (!c) return FEMX_ERR_ARG;
  auto &P = c->peer;
  ARG(P.connected && P.deferred, "femx_peer_complete belongs to deferred peer connections");
  ARG(P.pending || P.upending, "no assembly / scatter is outstanding");
  CK(cudaSetDevice(c->device));
  FlagPtrs fp{};
  if (P.upending) {
    peer_signal_wait_kernel<<<1, 32, 0, c->stream>>>(fp, reinterpret_cast<unsigned long long *>(P.arena + 512), P.nranks, P.rank, P.nbr_mask, P.ustep, 2000000ll,
                                                     P.d_err, false, true);
    c->launches += 1 + launch_scatter_local(c, P.pend_uowned, P.pend_ulocal);
    P.upending = false;
    CK(cudaStreamSynchronize(c->stream));
    CK(cudaGetLastError());
    int rc = peer_check_error(c);
    if (rc || !P.pending) return rc;
  }
  for (int q = 0; q < P.nranks; ++q) fp.remote[q] = reinterpret_cast<unsigned long long *>(P.remote[q]);
  // every peer has already signalled (its assembly ran before this call): a short limit turns a protocol error into a message
  peer_signal_wait_kernel<<<1, 32, 0, c->stream>>>(fp, reinterpret_cast<unsigned long long *>(P.arena), P.nranks, P.rank, P.nbr_mask, P.step, 2000000ll,
                                                   P.d_err, false, true);
  c->launches += 1 + launch_peer_merge(c, P.pend_vals, P.pend_fint);
  P.pending = false;
  CK(cudaStreamSynchronize(c->stream));
  CK(cudaGetLastError());
  return peer_check_error(c);
}

int femx_peer_last_error(femx_ctx *c) {   // used by the assemble path
  if (!c) return FEMX_ERR_ARG;
  return peer_check_error(c);
}

int femx_peer_set_scatter(femx_ctx *c, const int64_t *send_ptr, const int64_t *send_rows, const int64_t *u_dst, const int64_t *recv_u) {
  if (!c) return FEMX_ERR_ARG;
  auto &P = c->peer;
  ARG(P.connected, "femx_peer_connect first");
  ARG(send_ptr && u_dst && recv_u && send_ptr[0] == 0, "bad scatter lists");
  ARG(c->own_row_end >= c->own_row_start && c->own_row_end >= 0, "femx_set_ownership first");
  CK(cudaSetDevice(c->device));
  const int64_t n = send_ptr[P.nranks];
  ARG(n == 0 || send_rows, "bad scatter lists");
  for (int q = 0; q < P.nranks; ++q) {
    ARG(send_ptr[q] <= send_ptr[q + 1], "scatter ptr must ascend");
    ARG(q == P.rank || u_dst[q] + (send_ptr[q + 1] - send_ptr[q]) <= recv_u[q], "ghost block does not fit the peer's u area");
    P.udst[q] = u_dst[q]; P.rnu[q] = recv_u[q];
    if (q != P.rank && send_ptr[q + 1] > send_ptr[q]) P.nbr_mask |= 1u << q;
  }
  for (int64_t i = 0; i < n; ++i) ARG(send_rows[i] >= c->own_row_start && send_rows[i] < c->own_row_end, "scatter row outside the owned range");
  // ghosts of mine: their owners are neighbours too (they will signal me)
  for (int q = 0; q < P.nranks; ++q)
    if (q != P.rank && P.fb[q + 1] > P.fb[q]) P.nbr_mask |= 1u << q;
  dev_release(c, P.d_sptr); dev_release(c, P.d_srow);
  P.d_sptr = P.d_srow = nullptr;
  int rc;
  if ((rc = upload(c, &P.d_sptr, std::vector<int64_t>(send_ptr, send_ptr + P.nranks + 1)))) return rc;
  if ((rc = upload(c, &P.d_srow, std::vector<int64_t>(send_rows, send_rows + n)))) return rc;
  P.n_scatter = n;
  CK(cudaStreamSynchronize(c->stream));
  return FEMX_OK;
}

static const double *my_uarea(const femx_ctx *c, int par) {
  const auto &P = c->peer;
  return reinterpret_cast<const double *>(P.arena + kFlagBytes) + 2 * buffer_doubles(P.nv, P.nf) + par * P.nu;
}

static int launch_scatter_local(femx_ctx *c, const double *d_uown, double *d_uloc) {
  auto &P = c->peer;
  const long long n = (long long)c->pdim * c->nlnds;
  if (!n) return 0;
  const int par = (int)((P.ustep - 1) & 1);
  peer_scatter_local_kernel<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>(d_uloc, d_uown, my_uarea(c, par), c->dm.nl2g, c->pdim, n, P.fb[P.rank],
                                                                                P.fb[P.rank + 1], c->own_row_start);
  return 1;
}

// GetVec_U (src/m_global.F90:549-564; scatter built at src/main.F90:183-190): u_owned = this rank's slice [row_start, row_end)
// of the global solution vector, u_local [pdim * nlnds] = the ghosted local vector every kernel reads.
int femx_scatter_u(femx_ctx *c, const double *u_owned, double *u_local, int where) {
  if (!c) return FEMX_ERR_ARG;
  auto &P = c->peer;
  ARG(P.connected && P.d_sptr, "femx_peer_connect / femx_peer_set_scatter first");
  ARG(u_owned && u_local, "NULL argument");
  ARG(!P.upending, "a deferred femx_scatter_u is outstanding (femx_peer_complete)");
  CK(cudaSetDevice(c->device));
  const int64_t nown = c->own_row_end - c->own_row_start, nloc = (int64_t)c->pdim * c->nlnds;
  View uo(c, u_owned, sizeof(double) * (size_t)nown, where, true, false);
  View ul(c, u_local, sizeof(double) * (size_t)nloc, where, false, true);
  if (uo.rc || ul.rc) return uo.rc | ul.rc;
  int rc;
  if ((rc = begin_timing(c))) return rc;
  const int par = (int)(P.ustep & 1);
  ++P.ustep;
  ScatterDst dst{};
  FlagPtrs fp{};
  for (int q = 0; q < P.nranks; ++q) {
    if (q == P.rank || !P.remote[q]) continue;
    // the flag words of the u exchange sit behind those of the assembly exchange (512 bytes further)
    fp.remote[q] = reinterpret_cast<unsigned long long *>(P.remote[q] + 512);
    dst.u[q] = reinterpret_cast<double *>(P.remote[q] + kFlagBytes) + 2 * buffer_doubles(P.rnv[q], P.rnf[q]) + par * P.rnu[q] + P.udst[q];
  }
  const int grid = (int)std::min<int64_t>(296, std::max<int64_t>(1, (P.n_scatter + 255) / 256 / 4));
  const bool wait = !P.deferred;
  peer_scatter_push_kernel<<<grid, 256, 0, c->stream>>>(P.nranks, P.rank, P.d_sptr, P.d_srow, uo.as<double>(), c->own_row_start, dst, fp,
                                                        reinterpret_cast<unsigned long long *>(P.arena + 512), P.nbr_mask, P.ustep, 20000000000ll, P.d_err,
                                                        reinterpret_cast<unsigned int *>(P.d_err + 1), wait);
  c->launches += 1;
  if (wait) {
    c->launches += launch_scatter_local(c, uo.as<double>(), ul.as<double>());
    if ((rc = end_timing(c))) return rc;
    if ((rc = ul.finish())) return rc;
    if ((rc = finish_call(c, true))) return rc;
    return peer_check_error(c);
  }
  ARG(where == FEMX_DEVICE, "deferred (local) peers: femx_scatter_u works on device buffers");
  P.upending = true; P.pend_uowned = u_owned; P.pend_ulocal = u_local;
  if ((rc = end_timing(c))) return rc;
  return finish_call(c, true);
}

int femx_peer_recv(femx_ctx *c, const double **vals_recv, const double **f_recv) {
  if (!c) return FEMX_ERR_ARG;
  ARG(c->peer.connected && c->peer.step > 0, "no peer assembly has run");
  const auto &P = c->peer;
  const int par = (int)((P.step - 1) & 1);   // the buffer the last assembly filled
  const double *buf = reinterpret_cast<const double *>(P.arena + kFlagBytes) + par * buffer_doubles(P.nv, P.nf);
  if (vals_recv) *vals_recv = buf;
  if (f_recv) *f_recv = buf + P.nv;
  return FEMX_OK;
}

int femx_peer_disconnect(femx_ctx *c) {
  if (!c) return FEMX_ERR_ARG;
  auto &P = c->peer;
  cudaSetDevice(c->device);
  cudaStreamSynchronize(c->stream);
  for (int q = 0; q < kMaxPeers; ++q)
    if (P.remote[q]) { if (!P.local) cudaIpcCloseMemHandle(P.remote[q]); P.remote[q] = nullptr; }
  P.connected = false;   // the arena stays until femx_destroy: peers may still have it mapped
  P.deferred = false; P.pending = false;
  P.split_valid = false;
  return FEMX_OK;
}

}  // extern "C"
