// api_util.h -- helpers shared by the C-ABI translation units (error latching, staging views).
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>

#include "elem_math.cuh"
#include "launch.h"

using namespace femx;

#define CK(call)                                                                         \
  do {                                                                                   \
    cudaError_t _e = (call);                                                             \
    if (_e != cudaSuccess) {                                                             \
      char _b[512];                                                                      \
      snprintf(_b, sizeof _b, "%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(_e)); \
      c->err = _b;                                                                       \
      return FEMX_ERR_CUDA;                                                              \
    }                                                                                    \
  } while (0)

#define ARG(cond, msg)                 \
  do {                                 \
    if (!(cond)) {                     \
      c->err = msg;                    \
      return FEMX_ERR_ARG;             \
    }                                  \
  } while (0)

namespace {

template <class T>
int dev_alloc(femx_ctx *c, T **p, int64_t n) {
  *p = nullptr;
  if (n <= 0) n = 1;
  void *q = nullptr;
  cudaError_t e = cudaMalloc(&q, sizeof(T) * (size_t)n);
  if (e != cudaSuccess) {
    c->err = std::string("cudaMalloc failed: ") + cudaGetErrorString(e);
    return FEMX_ERR_ALLOC;
  }
  c->allocs.push_back(q);
  *p = (T *)q;
  return FEMX_OK;
}

void dev_release(femx_ctx *c, void *p) {
  if (!p) return;
  auto it = std::find(c->allocs.begin(), c->allocs.end(), p);
  if (it != c->allocs.end()) c->allocs.erase(it);
  cudaFree(p);
}

template <class T>
int upload(femx_ctx *c, T **d, const std::vector<T> &h) {
  int rc = dev_alloc(c, d, (int64_t)h.size());
  if (rc) return rc;
  if (!h.empty()) CK(cudaMemcpyAsync(*d, h.data(), sizeof(T) * h.size(), cudaMemcpyHostToDevice, c->stream));
  return FEMX_OK;
}

// scoped device view of a caller buffer
struct View {
  femx_ctx *c;
  void *dev = nullptr;
  void *host = nullptr;
  size_t bytes = 0;
  bool temp = false, out = false;
  int rc = FEMX_OK;
  View(femx_ctx *c_, const void *p, size_t bytes_, int where, bool in, bool out_) : c(c_), bytes(bytes_), out(out_) {
    if (!p) return;
    if (where == FEMX_DEVICE) { dev = const_cast<void *>(p); return; }
    host = const_cast<void *>(p);
    temp = true;
    if (cudaMalloc(&dev, bytes ? bytes : 1) != cudaSuccess) { rc = FEMX_ERR_ALLOC; c->err = "cudaMalloc (staging) failed"; dev = nullptr; return; }
    if (in && bytes && cudaMemcpyAsync(dev, host, bytes, cudaMemcpyHostToDevice, c->stream) != cudaSuccess) {
      rc = FEMX_ERR_CUDA; c->err = "H2D staging copy failed";
    }
  }
  int finish() {  // D2H for outputs; call before destruction when the result matters
    if (temp && out && dev && bytes)
      if (cudaMemcpyAsync(host, dev, bytes, cudaMemcpyDeviceToHost, c->stream) != cudaSuccess) { c->err = "D2H copy failed"; return FEMX_ERR_CUDA; }
    return FEMX_OK;
  }
  ~View() {
    if (temp && dev) { cudaStreamSynchronize(c->stream); cudaFree(dev); }
  }
  template <class T> T *as() { return (T *)dev; }
};

void invert(int n, const double *A, double *inv) {  // Gauss-Jordan with partial pivoting (setup only)
  std::vector<double> M(n * 2 * n, 0.0);
  for (int i = 0; i < n; ++i) {
    for (int j = 0; j < n; ++j) M[i * 2 * n + j] = A[i * n + j];
    M[i * 2 * n + n + i] = 1.0;
  }
  for (int col = 0; col < n; ++col) {
    int p = col;
    for (int i = col + 1; i < n; ++i)
      if (std::fabs(M[i * 2 * n + col]) > std::fabs(M[p * 2 * n + col])) p = i;
    if (p != col)
      for (int j = 0; j < 2 * n; ++j) std::swap(M[col * 2 * n + j], M[p * 2 * n + j]);
    const double d = 1.0 / M[col * 2 * n + col];
    for (int j = 0; j < 2 * n; ++j) M[col * 2 * n + j] *= d;
    for (int i = 0; i < n; ++i)
      if (i != col) {
        const double f = M[i * 2 * n + col];
        if (f != 0.0)
          for (int j = 0; j < 2 * n; ++j) M[i * 2 * n + j] -= f * M[col * 2 * n + j];
      }
  }
  for (int i = 0; i < n; ++i)
    for (int j = 0; j < n; ++j) inv[i * n + j] = M[i * 2 * n + n + j];
}

// N2(j,:) = N(xi_j) of ShapeFuncPrecompQua/Hex (src/m_elems.F90:332-354, :538-566), inverted once
void nodal_stress_tables(double *qua16, double *hex64) {
  const double g = FEMX_G;
  double N[64];
  for (int ip = 0; ip < 4; ++ip) {
    const double e = (ip & 2) ? g : -g, n = (ip & 1) ? g : -g;
    N[ip * 4 + 0] = 0.25 * (1 - e) * (1 - n); N[ip * 4 + 1] = 0.25 * (1 + e) * (1 - n);
    N[ip * 4 + 2] = 0.25 * (1 + e) * (1 + n); N[ip * 4 + 3] = 0.25 * (1 - e) * (1 + n);
  }
  invert(4, N, qua16);
  const int sx[8] = {-1, 1, 1, -1, -1, 1, 1, -1}, sy[8] = {-1, -1, 1, 1, -1, -1, 1, 1}, sz[8] = {-1, -1, -1, -1, 1, 1, 1, 1};
  for (int ip = 0; ip < 8; ++ip)
    for (int a = 0; a < 8; ++a)
      N[ip * 8 + a] = 0.125 * (1 + sx[a] * sx[ip] * g) * (1 + sy[a] * sy[ip] * g) * (1 + sz[a] * sz[ip] * g);
  invert(8, N, hex64);
}

int begin_timing(femx_ctx *c) { CK(cudaEventRecord(c->ev0, c->stream)); return FEMX_OK; }
int end_timing(femx_ctx *c) { CK(cudaEventRecord(c->ev1, c->stream)); return FEMX_OK; }
int finish_call(femx_ctx *c, bool timed) {
  if (c->async) {                       // femx_set_async: the caller orders later work on the stream / calls femx_synchronize
    c->ms_pending = timed;
    CK(cudaPeekAtLastError());
    return FEMX_OK;
  }
  CK(cudaStreamSynchronize(c->stream));
  CK(cudaGetLastError());
  if (timed) {
    float ms = 0.f;
    CK(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
    c->last_ms = ms;
  }
  return FEMX_OK;
}

}  // namespace
