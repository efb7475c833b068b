// Round-2 micro-benchmarks that decide the hex8 kernel's data path:
//   1. does DMMA.8x8x4 share the FP64 FMA pipe?  (DMMA alone, DFMA alone, both interleaved in one warp, both on
//      different warps of one SM)
//   2. DMMA dependent-chain latency
//   3. shared-memory load rates (LDS.64 / LDS.128, conflict-free) with the results kept live
//   4. global store paths: warp-contiguous STG.64 runs vs cp.async.bulk (shared -> global) issued by one thread,
//      alone and under concurrent LDS load from the other warps
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench2 tools/ubench2.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define ITERS 2048

__device__ __forceinline__ void dmma(double &c0, double &c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// mode 0: 8 independent DMMA per iteration; 1: 16 independent DFMA (x2 = 32 per iteration: same 64-lane-clk budget? see report);
// 2: both interleaved in every warp; 3: even warps DMMA, odd warps DFMA
__global__ void pipe_kernel(double *out, int mode) {
  double c[8][2], x[16];
  for (int i = 0; i < 8; ++i) c[i][0] = c[i][1] = 0.0;
  for (int i = 0; i < 16; ++i) x[i] = threadIdx.x * 1e-3 + i;
  const double a = threadIdx.x * 1e-3, b = 1.0 + threadIdx.x * 1e-4;
  const int w = threadIdx.x >> 5;
  const bool do_mma = mode == 0 || mode == 2 || (mode == 3 && (w & 1) == 0);
  const bool do_fma = mode == 1 || mode == 2 || (mode == 3 && (w & 1) == 1);
  for (int it = 0; it < ITERS; ++it) {
    if (do_mma) {
#pragma unroll
      for (int i = 0; i < 8; ++i) dmma(c[i][0], c[i][1], a, b);
    }
    if (do_fma) {
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int i = 0; i < 16; ++i) x[i] = fma(x[i], 1.0000001, b);
    }
  }
  double s = 0;
  for (int i = 0; i < 8; ++i) s += c[i][0] + c[i][1];
  for (int i = 0; i < 16; ++i) s += x[i];
  if (s == 1.2345) out[0] = s;
}

__global__ void dmma_chain_kernel(double *out, long long *cyc) {
  double c0 = 0, c1 = 0;
  const double a = threadIdx.x * 1e-3, b = 1.0 + threadIdx.x * 1e-4;
  const long long t0 = clock64();
#pragma unroll 16
  for (int it = 0; it < 1024; ++it) dmma(c0, c1, a, b);
  const long long t1 = clock64();
  if (c0 + c1 == 1.2345) out[0] = c0;
  if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}

template <int W>   // W = 8 or 16 bytes per lane
__global__ void lds_kernel(double *out) {
  extern __shared__ __align__(16) double sm[];
  for (int i = threadIdx.x; i < 8192; i += blockDim.x) sm[i] = i;
  __syncthreads();
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  double acc0 = 0, acc1 = 0, acc2 = 0, acc3 = 0;
  int idx = (lane + w * 32) * (W / 8);
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const int o = (idx + (it * 4 + r) * 64) & 8191 & ~(W / 8 - 1);
      if (W == 8) {
        const double v = sm[o];
        if (r & 1) acc1 += v; else acc0 += v;
      } else {
        const double2 v = *reinterpret_cast<const double2 *>(&sm[o]);
        if (r & 1) { acc2 += v.x; acc3 += v.y; } else { acc0 += v.x; acc1 += v.y; }
      }
    }
  }
  if (acc0 + acc1 + acc2 + acc3 == 1.2345) out[0] = acc0;
}

// ---- store paths.  Every CTA owns a contiguous slab of `per_cta` bytes and writes it `reps` times.
// mode 0: warp-contiguous STG.64 (216-byte runs of 27 lanes, like the gather of variant 12)
// mode 1: full-warp STG.64 (256 B per instruction)
// mode 2: full-warp STG.128 (512 B per instruction)
__global__ void stg_kernel(double *dst, long long per_cta_doubles, int mode) {
  double *base = dst + blockIdx.x * per_cta_doubles;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
  if (mode == 0) {
    for (long long o = (long long)w * 27; o + 27 <= per_cta_doubles; o += (long long)nw * 27)
      if (lane < 27) base[o + lane] = (double)lane;
  } else if (mode == 1) {
    for (long long o = (long long)w * 32; o + 32 <= per_cta_doubles; o += (long long)nw * 32) base[o + lane] = (double)lane;
  } else {
    for (long long o = (long long)w * 64; o + 64 <= per_cta_doubles; o += (long long)nw * 64)
      *reinterpret_cast<double2 *>(&base[o + 2 * lane]) = make_double2(lane, lane);
  }
}

// cp.async.bulk shared -> global: warp 0 lane 0 issues chunks of `chunk` bytes from a shared staging buffer; the other
// warps optionally hammer shared memory with LDS.128 (lds_load = 1) to see whether the two interfere
__global__ void bulk_kernel(double *dst, long long per_cta_bytes, int chunk, int lds_load, double *out) {
  extern __shared__ __align__(128) unsigned char smraw[];
  double *sm = reinterpret_cast<double *>(smraw);
  const int nd = 2 * chunk / 8;   // two staging buffers
  for (int i = threadIdx.x; i < nd; i += blockDim.x) sm[i] = i;
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  if (threadIdx.x == 0) {
    char *g = reinterpret_cast<char *>(dst) + blockIdx.x * per_cta_bytes;
    const unsigned s0 = (unsigned)__cvta_generic_to_shared(sm);
    int k = 0;
    for (long long o = 0; o + chunk <= per_cta_bytes; o += chunk, ++k) {
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(g + o), "r"(s0 + (k & 1) * chunk), "r"(chunk) : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  } else if (lds_load && threadIdx.x >= 32) {
    double acc0 = 0, acc1 = 0;
    const int n2 = nd / 2;
    for (int it = 0; it < lds_load; ++it) {
      const double2 v = *reinterpret_cast<const double2 *>(&sm[2 * ((threadIdx.x + it * 67) % n2)]);
      acc0 += v.x; acc1 += v.y;
    }
    if (acc0 + acc1 == 1.2345) out[0] = acc0;
  }
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
  printf("device %s, %d SMs, %.0f MHz\n", p.name, sms, clk * 1e-6);
  double *d;
  cudaMalloc(&d, 1024);
  long long *dc;
  cudaMalloc(&dc, 64);
  for (int threads : {128, 256, 512}) {
    const int blocks = sms;
    const double warps = threads / 32.0;
    const char *nm[4] = {"DMMA only (8/iter)", "DFMA only (64/iter)", "DMMA 8 + DFMA 64 same warp", "DMMA even warps, DFMA odd warps"};
    for (int mode = 0; mode < 4; ++mode) {
      const float ms = timeit([&] { pipe_kernel<<<blocks, threads>>>(d, mode); });
      const double cyc = ms * 1e-3 * clk;
      double mma_w = (mode == 0 || mode == 2) ? warps : (mode == 3 ? warps / 2 : 0), fma_w = (mode == 1 || mode == 2) ? warps : (mode == 3 ? warps / 2 : 0);
      const double mma_flop = mma_w * ITERS * 8 * 512.0, fma_flop = fma_w * ITERS * 64 * 64.0;
      printf("threads %4d  %-34s %8.3f ms  DMMA %.1f + DFMA %.1f = %.1f flop/clk/SM\n", threads, nm[mode], ms, mma_flop / cyc, fma_flop / cyc, (mma_flop + fma_flop) / cyc);
    }
  }
  {
    dmma_chain_kernel<<<1, 32>>>(d, dc);
    long long h;
    cudaMemcpy(&h, dc, 8, cudaMemcpyDeviceToHost);
    printf("DMMA dependent chain: %.1f clk per DMMA\n", h / 1024.0);
  }
  {
    cudaFuncSetAttribute(lds_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
    cudaFuncSetAttribute(lds_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
    for (int threads : {256, 512, 1024}) {
      float ms = timeit([&] { lds_kernel<8><<<sms, threads, 65536>>>(d); });
      printf("LDS.64  threads %4d: %.2f B/clk/SM\n", threads, threads * 8.0 * ITERS * 4 / (ms * 1e-3 * clk));
      ms = timeit([&] { lds_kernel<16><<<sms, threads, 65536>>>(d); });
      printf("LDS.128 threads %4d: %.2f B/clk/SM\n", threads, threads * 16.0 * ITERS * 4 / (ms * 1e-3 * clk));
    }
  }
  {
    const long long total = 2ll << 30;   // 2 GiB
    double *dst;
    cudaMalloc(&dst, total);
    for (int ctas_per_sm : {1, 2}) {
      const int blocks = sms * ctas_per_sm;
      const long long per_cta = (total / blocks) & ~4095ll;
      for (int mode = 0; mode < 3; ++mode) {
        const float ms = timeit([&] { stg_kernel<<<blocks, 512>>>(dst, per_cta / 8, mode); });
        printf("STG mode %d (%s), %d CTA/SM x 512 thr: %.1f GB/s\n", mode, mode == 0 ? "27-lane 216 B runs" : (mode == 1 ? "STG.64 256 B" : "STG.128 512 B"), ctas_per_sm,
               (double)per_cta * blocks / ms * 1e-6);
      }
      for (int chunk : {2048, 4096, 16384, 32768}) {
        cudaFuncSetAttribute(bulk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * 32768);
        for (int lds : {0, 1}) {
          const int lds_iters = lds ? (int)(per_cta / chunk) * 40 : 0;
          const float ms = timeit([&] { bulk_kernel<<<blocks, 512, 2 * chunk>>>(dst, per_cta, chunk, lds_iters, d); });
          printf("bulk store chunk %5d B, %d CTA/SM, %s: %.1f GB/s\n", chunk, ctas_per_sm, lds ? "15 warps LDS.128" : "idle warps      ", (double)(per_cta / chunk * chunk) * blocks / ms * 1e-6);
        }
      }
    }
  }
  printf("err=%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
