// Micro-benchmarks that decide the fused hex8 kernel's data path (DESIGN.md "measured inputs"):
// LDS.128/LDS.64 wavefront cost under broadcast patterns, SHFL rate, DMMA m8n8k4 rate.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench tools/ubench.cu
#include <cstdio>
#include <cuda_runtime.h>

#define ITERS 4096

// pattern: 0 = 32 distinct 16B (conflict-free), 1 = one address per quarter-warp, 2 = one per warp,
// 3 = lanes l and l+8.. share (address = lane%8): 8 distinct spread over all quarters
template <int PAT>
__global__ void lds128_kernel(double *out) {
  extern __shared__ double2 sm[];
  for (int i = threadIdx.x; i < 4096; i += blockDim.x) sm[i] = make_double2(i, 1.0);
  __syncthreads();
  const int lane = threadIdx.x & 31;
  int idx = PAT == 0 ? lane : (PAT == 1 ? lane / 8 : (PAT == 2 ? 0 : lane % 8));
  idx += (threadIdx.x >> 5) * 33;
  double2 acc = make_double2(0, 0);
#pragma unroll 8
  for (int it = 0; it < ITERS; ++it) {
    double2 v = sm[(idx + it * 37) & 4095];
    acc.x += v.x;
    acc.y += v.y;
  }
  if (acc.x == 1.2345) out[0] = acc.y;
}

template <int PAT>
__global__ void lds64_kernel(double *out) {
  extern __shared__ double sm1[];
  for (int i = threadIdx.x; i < 8192; i += blockDim.x) sm1[i] = i;
  __syncthreads();
  const int lane = threadIdx.x & 31;
  int idx = PAT == 0 ? lane : (PAT == 1 ? lane / 8 : (PAT == 2 ? 0 : lane % 8));
  idx += (threadIdx.x >> 5) * 33;
  double acc = 0;
#pragma unroll 8
  for (int it = 0; it < ITERS; ++it) acc += sm1[(idx + it * 37) & 8191];
  if (acc == 1.2345) out[0] = acc;
}

__global__ void shfl_kernel(double *out) {
  int v[8];
  for (int i = 0; i < 8; ++i) v[i] = threadIdx.x + i;
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = __shfl_xor_sync(0xffffffffu, v[i], 1 + (i & 3));
  }
  int s = 0;
  for (int i = 0; i < 8; ++i) s += v[i];
  if (s == 123456789) out[0] = s;
}

__global__ void dmma_kernel(double *out) {
  double c[8][2];
  for (int i = 0; i < 8; ++i) c[i][0] = c[i][1] = 0.0;
  double a = threadIdx.x * 1e-3, b = 1.0 + threadIdx.x * 1e-4;
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i)
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                   : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(a), "d"(b));
  }
  double s = 0;
  for (int i = 0; i < 8; ++i) s += c[i][0] + c[i][1];
  if (s == 1.2345) out[0] = s;
}

// DFMA chain interleaved with LDS.128: do the two pipes overlap?
__global__ void dfma_lds_kernel(double *out, int with_lds) {
  extern __shared__ double2 sm[];
  for (int i = threadIdx.x; i < 4096; i += blockDim.x) sm[i] = make_double2(1e-9, 1e-9);
  __syncthreads();
  double x[12];
  for (int i = 0; i < 12; ++i) x[i] = threadIdx.x * 1e-3 + i;
  int idx = threadIdx.x;
  double2 v = make_double2(1e-9, 1e-9);
  for (int it = 0; it < ITERS; ++it) {
    if (with_lds) v = sm[(idx + it * 37) & 4095];
#pragma unroll
    for (int i = 0; i < 12; ++i) x[i] = fma(x[i], 1.0000001, i & 1 ? v.x : v.y);
  }
  double s = 0;
  for (int i = 0; i < 12; ++i) s += x[i];
  if (s == 1.2345) out[0] = s;
}

template <class F>
float timeit(F f) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  f();
  cudaEventRecord(e0);
  f();
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  return ms;
}

int main() {
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, 0);
  const int sms = p.multiProcessorCount;
  const double clk = p.clockRate * 1e3;
  double *d;
  cudaMalloc(&d, 1024);
  const int threads = 512, blocks = sms * 2, warps_per_sm = 2 * threads / 32;
  auto report = [&](const char *name, float ms, double instr_per_warp) {
    const double cyc = ms * 1e-3 * clk;
    printf("%-34s %8.3f ms  %.3f clk per warp-instr per SM\n", name, ms, cyc / (instr_per_warp * warps_per_sm));
  };
  cudaFuncSetAttribute(lds128_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  cudaFuncSetAttribute(lds128_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  cudaFuncSetAttribute(lds128_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  cudaFuncSetAttribute(lds128_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  report("LDS.128 32 distinct", timeit([&] { lds128_kernel<0><<<blocks, threads, 65536>>>(d); }), ITERS);
  report("LDS.128 1 addr per quarter-warp", timeit([&] { lds128_kernel<1><<<blocks, threads, 65536>>>(d); }), ITERS);
  report("LDS.128 1 addr per warp", timeit([&] { lds128_kernel<2><<<blocks, threads, 65536>>>(d); }), ITERS);
  report("LDS.128 8 addrs (lane%8)", timeit([&] { lds128_kernel<3><<<blocks, threads, 65536>>>(d); }), ITERS);
  report("LDS.64 32 distinct", timeit([&] { lds64_kernel<0><<<blocks, threads, 65536>>>(d); }), ITERS);
  report("LDS.64 1 addr per quarter-warp", timeit([&] { lds64_kernel<1><<<blocks, threads, 65536>>>(d); }), ITERS);
  report("LDS.64 1 addr per warp", timeit([&] { lds64_kernel<2><<<blocks, threads, 65536>>>(d); }), ITERS);
  report("LDS.64 8 addrs (lane%8)", timeit([&] { lds64_kernel<3><<<blocks, threads, 65536>>>(d); }), ITERS);
  report("SHFL.32 (xor)", timeit([&] { shfl_kernel<<<blocks, threads>>>(d); }), ITERS * 8.0);
  {
    float ms = timeit([&] { dmma_kernel<<<blocks, threads>>>(d); });
    report("DMMA m8n8k4", ms, ITERS * 8.0);
    printf("   -> DMMA %.2f TFLOP/s\n", 2.0 * 256 * ITERS * 8.0 * (blocks * threads / 32) / ms * 1e-9);
  }
  cudaFuncSetAttribute(dfma_lds_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  report("12 DFMA (no LDS), per iteration", timeit([&] { dfma_lds_kernel<<<blocks, threads, 65536>>>(d, 0); }), ITERS);
  report("12 DFMA + 1 LDS.128, per iteration", timeit([&] { dfma_lds_kernel<<<blocks, threads, 65536>>>(d, 1); }), ITERS);
  printf("err=%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
