// tcgen05.ld throughput per shape (all variants move 4 KB per warp instruction, 32 registers per thread).
#include <cstdio>
#include <cuda_runtime.h>
#include "ptx.cuh"
using namespace inksam;

#define REGS32(v) "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), \
  "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), \
  "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), \
  "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
#define LIST32 "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}"

template <int SHAPE>
__device__ __forceinline__ void ld(uint32_t a, uint32_t (&v)[32]) {
  if constexpr (SHAPE == 0) asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 " LIST32 ", [%32];" : REGS32(v) : "r"(a) : "memory");
  if constexpr (SHAPE == 1) asm volatile("tcgen05.ld.sync.aligned.16x256b.x8.b32 " LIST32 ", [%32];" : REGS32(v) : "r"(a) : "memory");
  if constexpr (SHAPE == 2) asm volatile("tcgen05.ld.sync.aligned.16x128b.x16.b32 " LIST32 ", [%32];" : REGS32(v) : "r"(a) : "memory");
  if constexpr (SHAPE == 3) asm volatile("tcgen05.ld.sync.aligned.16x64b.x32.b32 " LIST32 ", [%32];" : REGS32(v) : "r"(a) : "memory");
}

template <int SHAPE>
__global__ void __launch_bounds__(512, 1) k(int iters, int nwarps, long long* out, float* sink) {
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) { tmem_alloc<1>(smem_u32(&slot), 512); tmem_relinquish<1>(); }
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t base = slot + ((uint32_t)((warp & 3) * 32) << 16) + ((warp >> 2) * 128) % 512;
  uint32_t v[32];
  float acc = 0.f;
  __syncthreads();
  const long long t0 = clock64();
  if (warp < nwarps) {
    for (int i = 0; i < iters; ++i) {
      ld<SHAPE>(base + (i & 1) * 32, v);
      tmem_ld_wait();
      acc += __uint_as_float(v[0] ^ v[5] ^ v[13] ^ v[31]);
    }
  }
  __syncthreads();
  const long long t1 = clock64();
  if (threadIdx.x == 0) out[0] = t1 - t0;
  if (acc == 123.456f) sink[0] = acc;
  tc_fence_before(); __syncthreads();
  if (warp == 0) tmem_dealloc<1>(slot, 512);
}

// store throughput (32x32b.x32)
__global__ void __launch_bounds__(512, 1) kst(int iters, int nwarps, long long* out) {
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) { tmem_alloc<1>(smem_u32(&slot), 512); tmem_relinquish<1>(); }
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t base = slot + ((uint32_t)((warp & 3) * 32) << 16) + ((warp >> 2) * 128) % 512;
  uint32_t v[32];
  for (int i = 0; i < 32; ++i) v[i] = threadIdx.x + i;
  __syncthreads();
  const long long t0 = clock64();
  if (warp < nwarps) {
    for (int i = 0; i < iters; ++i) { tmem_st_32x32(base + (i & 1) * 32, v); tmem_st_wait(); }
  }
  __syncthreads();
  const long long t1 = clock64();
  if (threadIdx.x == 0) out[0] = t1 - t0;
  tc_fence_before(); __syncthreads();
  if (warp == 0) tmem_dealloc<1>(slot, 512);
}

int main() {
  long long* d; float* sink; cudaMalloc(&d, 8); cudaMalloc(&sink, 4);
  long long h;
  const int iters = 4000;
  const char* names[4] = {"32x32b.x32", "16x256b.x8", "16x128b.x16", "16x64b.x32"};
  for (int nw : {1, 4, 8, 16}) {
    double r[5];
    k<0><<<1, 512>>>(iters, nw, d, sink); cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost); r[0] = (double)nw * iters * 4096 / h;
    k<1><<<1, 512>>>(iters, nw, d, sink); cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost); r[1] = (double)nw * iters * 4096 / h;
    k<2><<<1, 512>>>(iters, nw, d, sink); cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost); r[2] = (double)nw * iters * 4096 / h;
    k<3><<<1, 512>>>(iters, nw, d, sink); cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost); r[3] = (double)nw * iters * 4096 / h;
    kst<<<1, 512>>>(iters, nw, d); cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost); r[4] = (double)nw * iters * 4096 / h;
    printf("warps %2d B/clk/SM:", nw);
    for (int i = 0; i < 4; ++i) printf("  ld %s %6.1f", names[i], r[i]);
    printf("  st 32x32b.x32 %6.1f\n", r[4]);
  }
  // all SMs busy at once (does the figure hold chip-wide?)
  k<0><<<148, 512>>>(iters, 8, d, sink); cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
  printf("148 CTAs, 8 warps each: ld 32x32b.x32 %6.1f B/clk/SM\n", 8.0 * iters * 4096 / h);
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
