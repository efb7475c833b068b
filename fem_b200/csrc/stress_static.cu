// stress_static.cu -- femx_stress_noalloc: nodal-averaged stress (RecoverStress, src/m_global.F90:243-301, as femx_stress
// computes it) for callers that may not allocate: code under CUDA graph capture, tight serving loops, harnesses that forbid
// cudaMalloc.  No counterpart in the reference.
//
// femx_stress keeps the per-element nodal stresses (pass 1) in a context buffer grown on demand with cudaMalloc.  Here that
// scratch is a file-scope device array of a fixed, modest size (FEMX_STRESS_NOALLOC_MAX_DOUBLES doubles = 16 MiB; every GPU
// that loads libfemx.so holds it until the process exits).  The two kernels are the ones femx_stress launches
// (mesh_stress_kernel, node_stress_gather_kernel in kernels.cu), so the results are bit-identical.
//
//   * Host code must not hand the array's NAME to a kernel: on the host the symbol is a stand-in, not a device address.  The
//     address is resolved with cudaGetSymbolAddress, once per device, with that device current, and kept in s_scratch[].
//   * Every call on a device shares the one array and finds the previous call's contents in it.  Nothing has to start at
//     zero: pass 1 writes every (element, node) slot pass 2 reads, in the same call, in stream order.  Calls that overlap on
//     the GPU (two contexts / two streams of one device running at the same time) race on it: serialise them.
//   * No cudaMalloc / cudaFree / allocating library call on this path: inputs and outputs are the caller's device buffers,
//     timing events and the stream exist since femx_create.
#include <mutex>

#include "api_util.h"

namespace {

constexpr int64_t kMaxDoubles = FEMX_STRESS_NOALLOC_MAX_DOUBLES;
constexpr int kMaxDevices = 64;

static __device__ __align__(256) double g_stress_scratch[kMaxDoubles];

std::mutex s_mu;
double *s_scratch[kMaxDevices] = {};   // device address of g_stress_scratch on device i (nullptr: not resolved yet)

// the current device must be `device`
int scratch_address(femx_ctx *c, int device, double **out) {
  if (device < 0 || device >= kMaxDevices) { c->err = "femx_stress_noalloc: device index beyond 63"; return FEMX_ERR_ARG; }
  std::lock_guard<std::mutex> lk(s_mu);
  if (!s_scratch[device]) {
    double *p = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void **)&p, g_stress_scratch);
    if (e != cudaSuccess || !p) {
      c->err = std::string("femx_stress_noalloc: cudaGetSymbolAddress(g_stress_scratch) -> ") + cudaGetErrorString(e);
      return FEMX_ERR_CUDA;
    }
    s_scratch[device] = p;
  }
  *out = s_scratch[device];
  return FEMX_OK;
}

}  // namespace

extern "C" {

int64_t femx_stress_noalloc_max_doubles(void) { return kMaxDoubles; }

int femx_stress_noalloc(femx_ctx *c, const double *d_uu_local, double *d_stress_node, int layout) {
  if (!c) return FEMX_ERR_ARG;
  ARG(c->has_plan && d_uu_local && d_stress_node, "femx_stress_noalloc: femx_build_plan first / NULL argument");
  ARG(layout == FEMX_ROW_MAJOR || layout == FEMX_COL_MAJOR, "femx_stress_noalloc: layout");
  const int cp = c->pdim == 2 ? 3 : 6;
  const int64_t need = (int64_t)c->eptr[c->nels] * cp;   // one stress vector per (element, node)
  if (need > kMaxDoubles) {
    char b[256];
    snprintf(b, sizeof b, "femx_stress_noalloc: mesh needs %lld doubles of scratch, the static array holds %lld (use femx_stress)",
             (long long)need, (long long)kMaxDoubles);
    c->err = b;
    return FEMX_ERR_UNSUPPORTED;
  }
  cudaPointerAttributes pa;
  if (cudaPointerGetAttributes(&pa, d_uu_local) != cudaSuccess || pa.type == cudaMemoryTypeUnregistered ||
      cudaPointerGetAttributes(&pa, d_stress_node) != cudaSuccess || pa.type == cudaMemoryTypeUnregistered) {
    cudaGetLastError();
    c->err = "femx_stress_noalloc: inputs and outputs must be device (or pinned / managed) buffers -- pageable host memory would need a staging allocation";
    return FEMX_ERR_ARG;
  }
  CK(cudaSetDevice(c->device));
  double *d_sn = nullptr;
  int rc;
  if ((rc = scratch_address(c, c->device, &d_sn))) return rc;
  if ((rc = begin_timing(c))) return rc;
  c->launches += launch_mesh_stress(c, d_uu_local, d_sn, d_stress_node, layout == FEMX_COL_MAJOR, 1);
  if ((rc = end_timing(c))) return rc;
  return finish_call(c, true);
}

}  // extern "C"
