// HBM-bound helper kernels of the image encoder: pixel normalisation + patch gather, LayerNorm -> bf16,
// fp32 -> bf16 cast, and the bias rows of padded window tokens.
#include <cuda_fp16.h>
#include "common.h"
#include "ptx.cuh"

namespace inksam {

// ------------------------------------------------------------------------------------------------ patchify
// Sam.preprocess + the im2col of PatchEmbed (SA/modeling/sam.py:164-174, image_encoder.py:391-395):
//   A[(b,py,px)][c*256 + ky*16 + kx] = bf16((pixel(b, c, 16py+ky, 16px+kx) - mean[c]) / std[c]),
// zero where the pixel lies in the right / bottom padding (the reference pads AFTER normalising).
struct PatchifyArgs {
  const void* images;
  int dtype;               // 0: uint8 HWC, 1: float32 CHW
  long long img_stride;    // bytes between images
  long long row_pitch;     // bytes between pixel rows
  long long plane_stride;  // bytes between channel planes (dtype 1)
  int h[32], w[32];        // valid (resized) extent of each image
  float mean[3], inv_std[3];
  __nv_bfloat16* out;      // [B*4096][768]
  int B;
};

__global__ void __launch_bounds__(256) patchify_kernel(const PatchifyArgs a) {
  const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long total = (long long)a.B * 4096 * 96;
  if (gid >= total) return;
  const int chunk = (int)(gid % 96);
  const long long m = gid / 96;
  const int b = (int)(m >> 12), py = (int)((m >> 6) & 63), px = (int)(m & 63);
  const int c = chunk >> 5, ky = (chunk >> 1) & 15, kx0 = (chunk & 1) * 8;
  const int y = py * 16 + ky, x0 = px * 16 + kx0;
  const int h = a.h[b], w = a.w[b];
  float v[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = 0.f;
  if (y < h) {
    const unsigned char* base = static_cast<const unsigned char*>(a.images) + (long long)b * a.img_stride;
    if (a.dtype == 0) {
      const unsigned char* row = base + (long long)y * a.row_pitch;
#pragma unroll
      for (int i = 0; i < 8; ++i)
        if (x0 + i < w) v[i] = ((float)row[(x0 + i) * 3 + c] - a.mean[c]) * a.inv_std[c];
    } else {
      const float* row = reinterpret_cast<const float*>(base + (long long)c * a.plane_stride + (long long)y * a.row_pitch);
#pragma unroll
      for (int i = 0; i < 8; ++i)
        if (x0 + i < w) v[i] = (row[x0 + i] - a.mean[c]) * a.inv_std[c];
    }
  }
  uint4 pk;
  __nv_bfloat162 t;
  t = __floats2bfloat162_rn(v[0], v[1]); pk.x = *reinterpret_cast<uint32_t*>(&t);
  t = __floats2bfloat162_rn(v[2], v[3]); pk.y = *reinterpret_cast<uint32_t*>(&t);
  t = __floats2bfloat162_rn(v[4], v[5]); pk.z = *reinterpret_cast<uint32_t*>(&t);
  t = __floats2bfloat162_rn(v[6], v[7]); pk.w = *reinterpret_cast<uint32_t*>(&t);
  *reinterpret_cast<uint4*>(a.out + m * 768 + chunk * 8) = pk;
}

int launch_patchify(const void* images, int dtype, int B, const int* hw, long long img_stride, long long row_pitch,
                    long long plane_stride, const float* mean, const float* stdv, void* out, cudaStream_t st) {
  if (B > 32) return set_error("patchify: at most 32 images per launch (got %d)", B);
  PatchifyArgs a{};
  a.images = images; a.dtype = dtype; a.img_stride = img_stride; a.row_pitch = row_pitch; a.plane_stride = plane_stride;
  for (int i = 0; i < B; ++i) {
    a.h[i] = hw[2 * i]; a.w[i] = hw[2 * i + 1];
    if (a.h[i] < 1 || a.w[i] < 1 || a.h[i] > 1024 || a.w[i] > 1024)
      return set_error("patchify: image %d has size %dx%d, expected 1..1024", i, a.h[i], a.w[i]);
  }
  for (int c = 0; c < 3; ++c) { a.mean[c] = mean[c]; a.inv_std[c] = 1.0f / stdv[c]; }
  a.out = static_cast<__nv_bfloat16*>(out); a.B = B;
  const long long total = (long long)B * 4096 * 96;
  patchify_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(a);
  CUDA_TRY(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------------ LayerNorm
// nn.LayerNorm over the last dim, fp32 statistics, biased variance (SA/modeling/image_encoder.py:168,180,
// eps 1e-6 from build_sam.py:72) -> bf16 activations for the next GEMM.  One warp per token, row held in registers.
// plain != 0: no normalisation, just the bf16 cast; with `stats` also the row's (sum, sum of squares) -- the form the
// first block's qkv GEMM reads when the LayerNorm is folded into it (gemm_args.h; later blocks get both from the
// epilogue of the GEMM that wrote the residual stream).
template <int MAXC>
__global__ void __launch_bounds__(256) layernorm_bf16_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                             const float* __restrict__ bsh, __nv_bfloat16* __restrict__ out,
                                                             int M, int D, float eps, int plain, float2* __restrict__ stats) {
  pdl_launch_dependents();  // the consuming GEMM (launched with PDL) may run its prologue under this kernel's last wave
  const int row = blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= M) return;
  const int nch = D >> 2;
  const float4* xr = reinterpret_cast<const float4*>(x + (size_t)row * D);
  float4 v[MAXC];
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < MAXC; ++i) {
    const int c = lane + 32 * i;
    if (c < nch) {
      v[i] = xr[c];
      sum += (v[i].x + v[i].y) + (v[i].z + v[i].w);
    }
  }
  float mean = 0.f, rstd = 1.f;
  if (plain && stats != nullptr) {
    float sq = 0.f;
#pragma unroll
    for (int i = 0; i < MAXC; ++i) {
      const int c = lane + 32 * i;
      if (c < nch) sq += (v[i].x * v[i].x + v[i].y * v[i].y) + (v[i].z * v[i].z + v[i].w * v[i].w);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      sum += __shfl_xor_sync(0xffffffffu, sum, o);
      sq += __shfl_xor_sync(0xffffffffu, sq, o);
    }
    if (lane == 0) stats[row] = make_float2(sum, sq);
  }
  if (!plain) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    mean = sum / (float)D;
    float sq = 0.f;
#pragma unroll
    for (int i = 0; i < MAXC; ++i) {
      const int c = lane + 32 * i;
      if (c < nch) {
        const float a = v[i].x - mean, b = v[i].y - mean, cc = v[i].z - mean, d = v[i].w - mean;
        sq += (a * a + b * b) + (cc * cc + d * d);
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);
    rstd = rsqrtf(sq / (float)D + eps);
  }
  uint2* orow = reinterpret_cast<uint2*>(out + (size_t)row * D);
#pragma unroll
  for (int i = 0; i < MAXC; ++i) {
    const int c = lane + 32 * i;
    if (c < nch) {
      float4 y = v[i];
      if (!plain) {
        const float4 g = reinterpret_cast<const float4*>(w)[c];
        const float4 be = reinterpret_cast<const float4*>(bsh)[c];
        y.x = (y.x - mean) * rstd * g.x + be.x;
        y.y = (y.y - mean) * rstd * g.y + be.y;
        y.z = (y.z - mean) * rstd * g.z + be.z;
        y.w = (y.w - mean) * rstd * g.w + be.w;
      }
      __nv_bfloat162 lo = __floats2bfloat162_rn(y.x, y.y), hi = __floats2bfloat162_rn(y.z, y.w);
      uint2 pk;
      pk.x = *reinterpret_cast<uint32_t*>(&lo);
      pk.y = *reinterpret_cast<uint32_t*>(&hi);
      orow[c] = pk;
    }
  }
}

int launch_layernorm_bf16(const float* x, const float* w, const float* b, void* out, int M, int D, float eps, int plain,
                          float* stats, cudaStream_t st) {
  if (stats != nullptr && !plain) return set_error("layernorm: row statistics are only written in the plain (cast) mode");
  if (D % 4 != 0 || D > 1280) return set_error("layernorm: D=%d must be a multiple of 4 and <= 1280", D);
  const unsigned grid = (unsigned)((M + 7) / 8);
  auto* o = static_cast<__nv_bfloat16*>(out);
  if (D <= 256) layernorm_bf16_kernel<2><<<grid, 256, 0, st>>>(x, w, b, o, M, D, eps, plain, reinterpret_cast<float2*>(stats));
  else if (D <= 768) layernorm_bf16_kernel<6><<<grid, 256, 0, st>>>(x, w, b, o, M, D, eps, plain, reinterpret_cast<float2*>(stats));
  else layernorm_bf16_kernel<10><<<grid, 256, 0, st>>>(x, w, b, o, M, D, eps, plain, reinterpret_cast<float2*>(stats));
  CUDA_TRY(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------------ LayerNorm fold
// Weight side of folding `Linear(LayerNorm(x))` into one GEMM on the raw bf16 x (gemm_args.h):
//   w_out[n][k] = bf16(gamma[k] * W[n][k]);  colsum[n] = sum_k w_out[n][k]  (of the ROUNDED values, so that the mean term
//   cancels exactly what the tensor cores accumulate);  bias_out[n] = bias[n] + sum_k W[n][k] * beta[k].
// One warp per output row; runs once per weight load.
__global__ void __launch_bounds__(256) fold_ln_weights_kernel(const float* __restrict__ W, const float* __restrict__ gamma,
                                                              const float* __restrict__ beta, const float* __restrict__ bias,
                                                              __nv_bfloat16* __restrict__ w_out, float* __restrict__ colsum,
                                                              float* __restrict__ bias_out, int N, int K) {
  const int n = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (n >= N) return;
  const float* wr = W + (size_t)n * K;
  __nv_bfloat16* orow = w_out + (size_t)n * K;
  float cs = 0.f, bb = 0.f;
  for (int k = lane; k < K; k += 32) {
    const float w = wr[k];
    const __nv_bfloat16 r = __float2bfloat16_rn(w * gamma[k]);
    orow[k] = r;
    cs += __bfloat162float(r);
    bb = fmaf(w, beta[k], bb);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    cs += __shfl_xor_sync(0xffffffffu, cs, o);
    bb += __shfl_xor_sync(0xffffffffu, bb, o);
  }
  if (lane == 0) {
    colsum[n] = cs;
    bias_out[n] = bias[n] + bb;
  }
}

int launch_fold_ln_weights(const float* W, const float* gamma, const float* beta, const float* bias, void* w_out, float* colsum,
                           float* bias_out, int N, int K, cudaStream_t st) {
  fold_ln_weights_kernel<<<(unsigned)((N + 7) / 8), 256, 0, st>>>(W, gamma, beta, bias, static_cast<__nv_bfloat16*>(w_out), colsum,
                                                                 bias_out, N, K);
  CUDA_TRY(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------------ window pad rows
// In a windowed block the grid is zero-padded 64 -> 70 after norm1 (SA/modeling/image_encoder.py:258-259), so the
// qkv of a padded token is exactly the qkv bias.  The qkv GEMM stores real tokens in window order; this fills the
// rows of the padded tokens of that matrix ([B*25][196][3D] bf16) with bf16(bias).
// One block = 4 of the 804 padded tokens of an image (y >= 64 or x >= 64 on the 70x70 grid); a thread converts 8 bias
// values once and stores them (16 bytes) into the 4 rows.
__global__ void __launch_bounds__(128) fill_pad_rows_kernel(__nv_bfloat16* __restrict__ qkv, const float* __restrict__ bias,
                                                            int cols) {
  const int b = blockIdx.y, p0 = blockIdx.x * 4;
  size_t row[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int p = p0 + i;  // 0..803: the 6 bottom rows of the padded grid (420 tokens), then the 6 right columns of the rest
    const int y = p < 420 ? 64 + p / 70 : (p - 420) / 6, x = p < 420 ? p % 70 : 64 + (p - 420) % 6;
    const int win = (y / 14) * 5 + x / 14, t = (y % 14) * 14 + x % 14;
    row[i] = ((size_t)(b * 25 + win) * 196 + t) * (size_t)cols;
  }
  for (int c = threadIdx.x * 8; c < cols; c += blockDim.x * 8) {
    const float4 f0 = *reinterpret_cast<const float4*>(bias + c), f1 = *reinterpret_cast<const float4*>(bias + c + 4);
    __nv_bfloat162 h0 = __floats2bfloat162_rn(f0.x, f0.y), h1 = __floats2bfloat162_rn(f0.z, f0.w);
    __nv_bfloat162 h2 = __floats2bfloat162_rn(f1.x, f1.y), h3 = __floats2bfloat162_rn(f1.z, f1.w);
    uint4 pk;
    pk.x = *reinterpret_cast<uint32_t*>(&h0); pk.y = *reinterpret_cast<uint32_t*>(&h1);
    pk.z = *reinterpret_cast<uint32_t*>(&h2); pk.w = *reinterpret_cast<uint32_t*>(&h3);
#pragma unroll
    for (int i = 0; i < 4; ++i) *reinterpret_cast<uint4*>(qkv + row[i] + c) = pk;
  }
}

int launch_fill_pad_rows(void* qkv, const float* bias, int B, int cols, cudaStream_t st) {
  if (cols % 8 != 0) return set_error("fill_pad_rows: cols=%d must be a multiple of 8", cols);
  fill_pad_rows_kernel<<<dim3(804 / 4, B), 128, 0, st>>>(static_cast<__nv_bfloat16*>(qkv), bias, cols);
  CUDA_TRY(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------------ weight repack
// out[i0][i1][i2][i3] = in[perm], at most 4 dims; out dtype bf16 or fp32.  `perm[d]` is the input dim that output dim d
// walks (torch.permute convention).  Used once per tensor at load time.
struct RepackArgs {
  const float* in;
  void* out;
  int out_bf16;
  long long n;
  int oshape[4];
  long long istride_of_odim[4];
};

__global__ void __launch_bounds__(256) repack_kernel(const RepackArgs a) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < a.n; i += (long long)gridDim.x * blockDim.x) {
    long long rem = i, src = 0;
#pragma unroll
    for (int d = 3; d >= 0; --d) {
      const long long c = rem % a.oshape[d];
      rem /= a.oshape[d];
      src += c * a.istride_of_odim[d];
    }
    const float v = a.in[src];
    if (a.out_bf16 == 1) static_cast<__nv_bfloat16*>(a.out)[i] = __float2bfloat16_rn(v);
    else if (a.out_bf16 == 2) static_cast<__half*>(a.out)[i] = __float2half_rn(v);  // (decoder weights: |w| far below 65504)
    else static_cast<float*>(a.out)[i] = v;
  }
}

int launch_repack(const float* in, void* out, int out_bf16, const long long* ishape, int ndim, const int* perm,
                  cudaStream_t st) {
  if (ndim < 1 || ndim > 4) return set_error("repack: ndim %d not in 1..4", ndim);
  long long istride[4] = {1, 1, 1, 1};
  for (int d = ndim - 2; d >= 0; --d) istride[d] = istride[d + 1] * ishape[d + 1];
  RepackArgs a{};
  a.in = in; a.out = out; a.out_bf16 = out_bf16; a.n = 1;
  const int pad = 4 - ndim;
  for (int d = 0; d < 4; ++d) { a.oshape[d] = 1; a.istride_of_odim[d] = 0; }
  for (int d = 0; d < ndim; ++d) {
    const int src = perm ? perm[d] : d;
    a.oshape[pad + d] = (int)ishape[src];
    a.istride_of_odim[pad + d] = istride[src];
    a.n *= ishape[src];
  }
  long long blocks = (a.n + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  if (blocks < 1) blocks = 1;
  repack_kernel<<<(unsigned)blocks, 256, 0, st>>>(a);
  CUDA_TRY(cudaGetLastError());
  return 0;
}

}  // namespace inksam
