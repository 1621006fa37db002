// Microbenchmark: tcgen05.ld throughput per SM (how many bytes per clock TMEM -> registers) and MUFU.EX2 throughput, on one
// CTA per SM.  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I inklayer_b200/csrc tools/micro/tmem_bw.cu -o /tmp/tmem_bw
#include <cstdio>
#include <cuda_runtime.h>
#include "ptx.cuh"
using namespace inksam;

__global__ void __launch_bounds__(512, 1) tmem_ld_kernel(int iters, int nwarps, long long* out, float* sink) {
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) { tmem_alloc<1>(smem_u32(&slot), 512); tmem_relinquish<1>(); }
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t base = slot + ((uint32_t)((warp & 3) * 32) << 16);
  uint32_t v[32];
  float acc = 0.f;
  __syncthreads();
  const long long t0 = clock64();
  if (warp < nwarps) {
    for (int i = 0; i < iters; ++i) {
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        tmem_ld_32x32(base + ((warp >> 2) * 128 + c * 32) % 512, v);
        tmem_ld_wait();
        acc += __uint_as_float(v[lane & 31]);
      }
    }
  }
  __syncthreads();
  const long long t1 = clock64();
  if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = t1 - t0;
  if (acc == 123.456f) sink[0] = acc;
  tc_fence_before(); __syncthreads();
  if (warp == 0) tmem_dealloc<1>(slot, 512);
}

// same, 4 loads in flight before one wait
__global__ void __launch_bounds__(512, 1) tmem_ld4_kernel(int iters, int nwarps, long long* out, float* sink) {
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) { tmem_alloc<1>(smem_u32(&slot), 512); tmem_relinquish<1>(); }
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t base = slot + ((uint32_t)((warp & 3) * 32) << 16);
  uint32_t v0[32], v1[32], v2[32], v3[32];
  float acc = 0.f;
  __syncthreads();
  const long long t0 = clock64();
  if (warp < nwarps) {
    for (int i = 0; i < iters; ++i) {
      const uint32_t b = base + ((warp >> 2) * 128) % 512;
      tmem_ld_32x32(b, v0); tmem_ld_32x32(b + 32, v1); tmem_ld_32x32(b + 64, v2); tmem_ld_32x32(b + 96, v3);
      tmem_ld_wait();
      acc += __uint_as_float(v0[lane]) + __uint_as_float(v1[lane]) + __uint_as_float(v2[lane]) + __uint_as_float(v3[lane]);
    }
  }
  __syncthreads();
  const long long t1 = clock64();
  if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = t1 - t0;
  if (acc == 123.456f) sink[0] = acc;
  tc_fence_before(); __syncthreads();
  if (warp == 0) tmem_dealloc<1>(slot, 512);
}

__global__ void __launch_bounds__(512, 1) mufu_kernel(int iters, int nwarps, long long* out, float* sink) {
  const int warp = threadIdx.x >> 5;
  float x[8];
  for (int i = 0; i < 8; ++i) x[i] = -0.001f * (threadIdx.x + i);
  __syncthreads();
  const long long t0 = clock64();
  if (warp < nwarps) {
    for (int i = 0; i < iters; ++i) {
#pragma unroll
      for (int j = 0; j < 8; ++j) x[j] = ex2_approx(x[j]) - 1.0f;
    }
  }
  __syncthreads();
  const long long t1 = clock64();
  if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = t1 - t0;
  float s = 0; for (int i = 0; i < 8; ++i) s += x[i];
  if (s == 123.456f) sink[0] = s;
}

int main() {
  long long* d; float* sink; cudaMalloc(&d, 8); cudaMalloc(&sink, 4);
  long long h;
  const int iters = 2000;
  for (int nw : {1, 2, 4, 8, 16}) {
    tmem_ld_kernel<<<1, 512>>>(iters, nw, d, sink); cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
    const double b1 = (double)nw * iters * 4 * 4096 / h;
    tmem_ld4_kernel<<<1, 512>>>(iters, nw, d, sink); cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
    const double b4 = (double)nw * iters * 4 * 4096 / h;
    mufu_kernel<<<1, 512>>>(iters, nw, d, sink); cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
    const double m = (double)nw * iters * 8 * 32 / h;
    printf("warps %2d: tcgen05.ld x32 (wait each) %7.1f B/clk/SM   (4 in flight) %7.1f B/clk/SM   ex2 %6.2f lanes/clk/SM\n", nw, b1, b4, m);
  }
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
