// Measures the FP64 FMA-pipe peak (register-resident DFMA chains) and HBM write/copy bandwidth.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/fp64_peak tools/fp64_peak.cu
#include <cstdio>
#include <cuda_runtime.h>

__global__ void dfma_kernel(double *out, int iters, double a, double b) {
  double x[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) x[i] = threadIdx.x * 1e-3 + i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) x[i] = fma(x[i], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += x[i];
  if (s == 123.456) out[0] = s;
}

__global__ void write_kernel(double2 *out, size_t n) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) out[i] = make_double2(1.0, 2.0);
}

int main() {
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, 0);
  printf("device %s, %d SMs, clock %d kHz\n", p.name, p.multiProcessorCount, p.clockRate);
  double *d;
  size_t n = (size_t)1 << 28;  // 4 GiB of double2
  cudaMalloc(&d, n * 16);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int threads : {256, 512, 1024}) {
    const int blocks = p.multiProcessorCount * (2048 / threads), iters = 20000;
    dfma_kernel<<<blocks, threads>>>(d, 100, 1.0000001, 1e-9);
    cudaEventRecord(e0);
    dfma_kernel<<<blocks, threads>>>(d, iters, 1.0000001, 1e-9);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    double fl = 2.0 * 16 * iters * (double)blocks * threads;
    printf("DFMA threads/block=%d: %.2f TFLOP/s (%.3f ms)\n", threads, fl / ms * 1e-9, ms);
  }
  for (int rep = 0; rep < 3; ++rep) {
    cudaEventRecord(e0);
    write_kernel<<<p.multiProcessorCount * 8, 512>>>((double2 *)d, n);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    printf("HBM write: %.1f GB/s\n", n * 16.0 / ms * 1e-6);
  }
  for (int rep = 0; rep < 3; ++rep) {
    cudaEventRecord(e0);
    cudaMemcpyAsync(d, (char *)d + n * 8, n * 8, cudaMemcpyDeviceToDevice);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    printf("HBM copy (r+w): %.1f GB/s\n", 2.0 * n * 8.0 / ms * 1e-6);
  }
  return 0;
}
