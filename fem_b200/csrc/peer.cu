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

// lane q: tell rank q that this rank's stores of assembly `step` are done, then wait for q's
__device__ __forceinline__ void signal_wait(const FlagPtrs &fp, unsigned long long *mine, int n, int rank, unsigned long long step,
                                            long long limit_clocks, int *err) {
  const int q = threadIdx.x;
  if (q >= n || q == rank) return;
  __threadfence_system();
  st_release_sys(fp.remote[q] + 8 * rank, step);
  const long long t0 = clock64();
  while (ld_acquire_sys(mine + 8 * q) < step) {
    if (clock64() - t0 > limit_clocks) { *err = 1 + q; break; }   // a dead peer must not hang the GPU
    __nanosleep(200);
  }
}

__global__ void peer_signal_wait_kernel(FlagPtrs fp, unsigned long long *mine, int n, int rank, unsigned long long step,
                                        long long limit_clocks, int *err) {
  signal_wait(fp, mine, n, rank, step, limit_clocks, err);
}

// Push form (works behind every assembly variant): the kernel wrote all of its rows into the local arrays; copy the
// ranges other ranks own into their arenas with remote stores (coalesced 8-byte lanes over NVLink), then the last CTA
// to finish runs the step barrier.  fint is indexed by global row, the wire format by local dof (pdim * node + i).
__global__ void __launch_bounds__(256)
peer_push_kernel(PeerTab pt, const double *__restrict__ vals, const double *__restrict__ fint, const int32_t *__restrict__ nl2g, int pdim,
                 FlagPtrs fp, unsigned long long *mine, unsigned long long step, long long limit_clocks, int *err, unsigned int *ticket) {
  const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x, nth = (long long)gridDim.x * blockDim.x;
  for (int q = 0; q < pt.n; ++q) {
    if (q == pt.rank) continue;
    if (pt.vdst[q]) {
      const double *src = vals + pt.vb[q];
      double *dst = pt.vdst[q];
      const long long len = pt.vb[q + 1] - pt.vb[q];
      for (long long i = tid; i < len; i += nth) dst[i] = src[i];
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
  signal_wait(fp, mine, pt.n, pt.rank, step, limit_clocks, err);
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

// fused = the assembly kernel already stored the interface rows at their owners (PEER instantiations)
int launch_peer_finish(femx_ctx *c, bool fused, const double *d_vals, const double *d_fint) {
  auto &P = c->peer;
  FlagPtrs fp{};
  for (int q = 0; q < P.nranks; ++q) fp.remote[q] = reinterpret_cast<unsigned long long *>(P.remote[q]);
  unsigned long long *mine = reinterpret_cast<unsigned long long *>(P.arena);
  const long long limit = 20000000000ll;   // ~10 s
  if (fused) {
    ++P.step;
    peer_signal_wait_kernel<<<1, 32, 0, c->stream>>>(fp, mine, P.nranks, P.rank, P.step, limit, P.d_err);
    return 1;
  }
  const PeerTab pt = peer_tab(c);   // destinations of the current step's buffer
  ++P.step;
  int64_t n = 0;
  for (int q = 0; q < P.nranks; ++q) if (q != P.rank) n += P.vb[q + 1] - P.vb[q];
  int grid = (int)std::min<int64_t>(296, std::max<int64_t>(1, (n + 255) / 256 / 8));
  peer_push_kernel<<<grid, 256, 0, c->stream>>>(pt, d_vals, d_fint, c->dm.nl2g, c->pdim, fp, mine, P.step, limit, P.d_err,
                                                reinterpret_cast<unsigned int *>(P.d_err + 1));
  return 1;
}

}  // namespace femx

extern "C" {

int femx_peer_arena_create(femx_ctx *c, int64_t n_recv_vals, int64_t n_recv_f, char handle[64]) {
  if (!c) return FEMX_ERR_ARG;
  ARG(n_recv_vals >= 0 && n_recv_f >= 0 && handle, "bad arguments");
  ARG(!c->peer.arena, "the receive arena exists already");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size");
  CK(cudaSetDevice(c->device));
  auto &P = c->peer;
  P.nv = n_recv_vals; P.nf = n_recv_f;
  const size_t bytes = (size_t)kFlagBytes + 2 * sizeof(double) * (size_t)buffer_doubles(P.nv, P.nf);
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

int femx_peer_connect(femx_ctx *c, int nranks, int rank, const char *handles, const int64_t *recv_vals, const int64_t *recv_f,
                      const int64_t *val_bounds, const int64_t *val_dst, const int64_t *f_bounds, const int64_t *f_dst) {
  if (!c) return FEMX_ERR_ARG;
  ARG(c->has_plan, "femx_build_plan first");
  ARG(c->peer.arena, "femx_peer_arena_create first");
  ARG(nranks >= 2 && nranks <= kMaxPeers && rank >= 0 && rank < nranks, "peer exchange handles 2..8 ranks of one node");
  ARG(handles && recv_vals && recv_f && val_bounds && val_dst && f_bounds && f_dst, "NULL argument");
  ARG(c->own_row_end < 0 || (c->own_nranks == nranks && c->own_rank == rank), "femx_set_ownership named another rank / rank count");
  ARG(val_bounds[0] == 0 && val_bounds[nranks] == c->nnz, "val_bounds must span the local values");
  ARG(f_bounds[0] == 0 && f_bounds[nranks] == (int64_t)c->pdim * c->nlnds, "f_bounds must span the local dofs");
  CK(cudaSetDevice(c->device));
  auto &P = c->peer;
  P.nranks = nranks; P.rank = rank;
  for (int q = 0; q <= nranks; ++q) { P.vb[q] = val_bounds[q]; P.fb[q] = f_bounds[q]; }
  for (int q = 0; q < nranks; ++q) {
    ARG(val_bounds[q] <= val_bounds[q + 1] && f_bounds[q] <= f_bounds[q + 1], "bounds must ascend");
    P.rnv[q] = recv_vals[q]; P.rnf[q] = recv_f[q];
    P.vdst[q] = val_dst[q]; P.fdst[q] = f_dst[q];
    if (q == rank) continue;
    ARG(val_dst[q] >= 0 && val_dst[q] + (val_bounds[q + 1] - val_bounds[q]) <= recv_vals[q], "value slice does not fit the peer's arena");
    ARG(f_dst[q] >= 0 && f_dst[q] + (f_bounds[q + 1] - f_bounds[q]) <= recv_f[q], "force slice does not fit the peer's arena");
    cudaIpcMemHandle_t h;
    memcpy(&h, handles + 64 * (size_t)q, 64);
    void *p = nullptr;
    CK(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    P.remote[q] = (char *)p;
  }
  P.step = 0;
  P.connected = true;
  return FEMX_OK;
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
    if (P.remote[q]) { cudaIpcCloseMemHandle(P.remote[q]); P.remote[q] = nullptr; }
  P.connected = false;   // the arena stays until femx_destroy: peers may still have it mapped
  return FEMX_OK;
}

}  // extern "C"
