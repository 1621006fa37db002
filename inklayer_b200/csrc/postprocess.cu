// Fused mask post-process (HBM-bound): Sam.postprocess_masks + threshold in one pass
// (SA/modeling/sam.py:133-162, SA/predictor.py:238-241):
//   low-res logits [K][256][256] --bilinear--> 1024x1024 --crop [:h,:w]--> --bilinear--> (H,W) --> (> thr) as uint8.
// Both interpolations are align_corners=False without antialiasing; source indices, clamping and the lambda
// arithmetic follow ATen's upsample_bilinear2d (area_pixel_compute_source_index): the 1024x1024 intermediate is never
// materialised, each output pixel evaluates its 4 taps on the virtual 1024 grid, each tap 4 taps on the 256 grid.
#include "common.h"

namespace inksam {

struct Tap { int i0, i1; float l0, l1; };

__device__ __forceinline__ Tap src_tap(int dst, float scale, int in_size) {
  float r = scale * ((float)dst + 0.5f) - 0.5f;
  if (r < 0.f) r = 0.f;
  Tap t;
  t.i0 = (int)r;
  if (t.i0 > in_size - 1) t.i0 = in_size - 1;
  t.i1 = t.i0 + (t.i0 < in_size - 1 ? 1 : 0);
  t.l1 = fminf(fmaxf(r - (float)t.i0, 0.f), 1.f);
  t.l0 = 1.f - t.l1;
  return t;
}

// value of the virtual 1024x1024 map at (Y, X)
__device__ __forceinline__ float up1024(const float* __restrict__ low, int Y, int X) {
  const Tap ty = src_tap(Y, 0.25f, 256), tx = src_tap(X, 0.25f, 256);
  const float* r0 = low + ty.i0 * 256;
  const float* r1 = low + ty.i1 * 256;
  return ty.l0 * (tx.l0 * r0[tx.i0] + tx.l1 * r0[tx.i1]) + ty.l1 * (tx.l0 * r1[tx.i0] + tx.l1 * r1[tx.i1]);
}

__global__ void __launch_bounds__(256) postprocess_kernel(const float* __restrict__ low, unsigned char* __restrict__ masks,
                                                          float* __restrict__ logits, int in_h, int in_w, int H, int W,
                                                          float scale_y, float scale_x, float thr, int identity) {
  const int box = blockIdx.z, oy = blockIdx.y;
  const int x0 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (x0 >= W) return;
  const float* lo = low + (size_t)box * 65536;
  float v[4];
  Tap ty;
  if (!identity) ty = src_tap(oy, scale_y, in_h);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int ox = x0 + i;
    if (ox >= W) { v[i] = 0.f; continue; }
    if (identity) {
      v[i] = up1024(lo, oy, ox);
    } else {
      const Tap tx = src_tap(ox, scale_x, in_w);
      v[i] = ty.l0 * (tx.l0 * up1024(lo, ty.i0, tx.i0) + tx.l1 * up1024(lo, ty.i0, tx.i1)) +
             ty.l1 * (tx.l0 * up1024(lo, ty.i1, tx.i0) + tx.l1 * up1024(lo, ty.i1, tx.i1));
    }
  }
  const size_t o = ((size_t)box * H + oy) * W + x0;
  if (masks != nullptr) {
    if ((W & 3) == 0) {
      uchar4 m;
      m.x = v[0] > thr; m.y = v[1] > thr; m.z = v[2] > thr; m.w = v[3] > thr;
      *reinterpret_cast<uchar4*>(masks + o) = m;
    } else {
#pragma unroll
      for (int i = 0; i < 4; ++i)
        if (x0 + i < W) masks[o + i] = v[i] > thr;
    }
  }
  if (logits != nullptr) {
#pragma unroll
    for (int i = 0; i < 4; ++i)
      if (x0 + i < W) logits[o + i] = v[i];
  }
}

// low: [K][256][256] fp32; masks: [K][H][W] uint8 (0/1) or null; logits: [K][H][W] fp32 or null.
int launch_postprocess(const float* low, unsigned char* masks, float* logits, int K, int in_h, int in_w, int H, int W,
                       float thr, cudaStream_t st) {
  if (K <= 0) return 0;
  if (in_h < 1 || in_w < 1 || in_h > 1024 || in_w > 1024 || H < 1 || W < 1)
    return set_error("postprocess: bad sizes input %dx%d original %dx%d", in_h, in_w, H, W);
  const float sy = (float)in_h / (float)H, sx = (float)in_w / (float)W;
  const int identity = (in_h == H && in_w == W) ? 1 : 0;
  dim3 grid((unsigned)((W + 1023) / 1024), (unsigned)H, (unsigned)K);
  postprocess_kernel<<<grid, 256, 0, st>>>(low, masks, logits, in_h, in_w, H, W, sy, sx, thr, identity);
  CUDA_TRY(cudaGetLastError());
  return 0;
}

}  // namespace inksam
