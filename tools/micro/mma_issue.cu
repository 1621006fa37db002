// How fast can ONE thread issue small tcgen05.mma?  N back-to-back MMAs (M = 128, K = 16, bf16) of width n into TMEM from
// (uninitialised) shared memory, one commit at the end; prints cycles per MMA for n = 16 .. 256 with the operand A in smem
// and in TMEM.  The tensor-pipe floor is 128 * n / 256 cycles (cta_group::1).
#include <cstdio>
#include <cuda_runtime.h>
#include "ptx.cuh"
using namespace inksam;

template <int TS>
__global__ void __launch_bounds__(128, 1) k(int iters, int n, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  __shared__ uint32_t slot;
  __shared__ uint64_t bar;
  const int warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) { mbar_init(smem_u32(&bar), 1); fence_mbar_init(); }
  if (warp == 0) { tmem_alloc<1>(smem_u32(&slot), 512); tmem_relinquish<1>(); }
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tm = slot;
  if (threadIdx.x == 32) {
    const uint32_t idesc = make_idesc(1u, 128, (uint32_t)n);
    const uint64_t ad = make_kmajor_sw128_desc(base), bd = make_kmajor_sw128_desc(base + 16384);
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
      if (TS) umma_ts(tm, tm + 256, bd, idesc, 1u);
      else umma<1, 0>(tm, ad, bd, idesc, 1u);
    }
    umma_commit<1>(smem_u32(&bar));
    const long long t1 = clock64();
    mbar_wait(smem_u32(&bar), 0);
    const long long t2 = clock64();
    out[0] = t1 - t0; out[1] = t2 - t0;
  }
  tc_fence_before(); __syncthreads();
  if (warp == 0) tmem_dealloc<1>(tm, 512);
}

int main() {
  long long* d; cudaMalloc(&d, 16);
  cudaFuncSetAttribute(k<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  cudaFuncSetAttribute(k<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  long long h[2];
  const int iters = 2000;
  for (int n : {16, 32, 64, 80, 128, 208, 256}) {
    k<0><<<1, 128, 100 * 1024>>>(iters, n, d); cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
    const double i0 = (double)h[0] / iters, e0 = (double)h[1] / iters;
    k<1><<<1, 128, 100 * 1024>>>(iters, n, d); cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
    printf("N %3d: A in smem: issue %6.1f clk/MMA, issue+drain %6.1f | A in TMEM: issue %6.1f, issue+drain %6.1f | floor %5.1f\n", n, i0,
           e0, (double)h[0] / iters, (double)h[1] / iters, 128.0 * n / 256);
  }
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
