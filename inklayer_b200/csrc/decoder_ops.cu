// Mask-decoder kernels that are not GEMMs: the 7-token side of the two-way transformer (fp32 on CUDA cores, one
// CTA per box), token->image and image->token attention over the 4096 image tokens, prompt / dense positional
// encodings.  The image-side linears (k/v/q projections of the 4096 keys, out_proj + norm4, the two transposed
// convolutions) run on the tcgen05 GEMM in tf32 (gemm_tc.cuh); see engine.cu for the order.
//
// Reference: SA/modeling/transformer.py:62-106,151-182,218-240; mask_decoder.py:112-149,154-176;
// prompt_encoder.py:62-71,93-100,185-214.
#include <cuda_fp16.h>
#include "common.h"
#include "decoder_ops.h"
#include "ptx.cuh"

namespace inksam {

constexpr int E = 256;       // transformer dim
constexpr int NT = 7;        // iou token + 4 mask tokens + 2 box corners
constexpr int TOK_THREADS = 512;

// out[t][n] = act(sum_k in[t][k] * W[n][k] + b[n]) for t < T; in/out in shared memory, W row-major in global.
// One warp per group of R = 4 output rows, lanes split k (coalesced 512-byte weight reads).  The weights come from L2
// (every CTA = box re-streams them): with one 16-byte load per row in flight the loop was bound by the L2 round trip
// (~600 cycles, 220 us for the two MLP phases of 128 boxes), so each lane now has R x U = 8 weight loads in flight.
template <int T>
__device__ __forceinline__ void linear_tok(float* __restrict__ out, int out_ld, const float* __restrict__ in, int in_ld,
                                           const float* __restrict__ W, const float* __restrict__ bias, int N, int Kd,
                                           bool relu, bool accumulate) {
  constexpr int R = 4, U = 2;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  const int k4n = Kd / 4;
  for (int n0 = warp * R; n0 < N; n0 += nwarps * R) {
    // two partial sums per (row, token) -- even and odd pairs of k -- so that the inner product runs on fma.rn.f32x2: half
    // the FMA-pipe instructions of this FMA-bound loop
    uint64_t acc2[R][T];
#pragma unroll
    for (int r = 0; r < R; ++r)
#pragma unroll
      for (int t = 0; t < T; ++t) acc2[r][t] = f2_pack(0.f, 0.f);
    const float4* w[R];
#pragma unroll
    for (int r = 0; r < R; ++r) w[r] = reinterpret_cast<const float4*>(W + (size_t)(n0 + r < N ? n0 + r : n0) * Kd);
    for (int k4 = lane; k4 < k4n; k4 += 32 * U) {
      float4 a[U][R];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int kk = k4 + 32 * u;
#pragma unroll
        for (int r = 0; r < R; ++r) a[u][r] = kk < k4n ? __ldg(w[r] + kk) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int kk = k4 + 32 * u;
        if (kk < k4n) {
#pragma unroll
          for (int t = 0; t < T; ++t) {
            const float4 x = *reinterpret_cast<const float4*>(in + t * in_ld + kk * 4);
            const uint64_t x01 = f2_pack(x.x, x.y), x23 = f2_pack(x.z, x.w);
#pragma unroll
            for (int r = 0; r < R; ++r)
              acc2[r][t] = f2_fma(f2_pack(a[u][r].z, a[u][r].w), x23, f2_fma(f2_pack(a[u][r].x, a[u][r].y), x01, acc2[r][t]));
          }
        }
      }
    }
    float acc[R][T];
#pragma unroll
    for (int r = 0; r < R; ++r)
#pragma unroll
      for (int t = 0; t < T; ++t) {
        float lo, hi;
        f2_unpack(acc2[r][t], lo, hi);
        acc[r][t] = lo + hi;
      }
    if constexpr (R * T > 16 && R * T <= 32) {
      // Lane-sum of R * T (= 28) values per lane as a reduce-scatter: at offset 16 every lane hands half of its values to
      // its partner and adds the other half's, then 8 values at offset 8, ... -- 31 shuffles and adds instead of 140, and lane
      // i ends with the total of value i (= acc[i / T][i % T]) and writes it itself.  (With K = 256 the 140 shuffle-add pairs
      // per 4 output rows outnumbered the 224 FMAs.)
      float v[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] = i < R * T ? acc[i / T][i % T] : 0.f;
#pragma unroll
      for (int off = 16; off >= 1; off >>= 1) {
        const bool up = (lane & off) != 0;
#pragma unroll
        for (int i = 0; i < off; ++i) {
          const float send = up ? v[i] : v[i + off];
          const float keep = up ? v[i + off] : v[i];
          v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
        }
      }
      const int r = lane / T, t = lane % T;
      if (lane < R * T && n0 + r < N) {
        float o = v[0] + (bias ? bias[n0 + r] : 0.f);
        if (relu) o = fmaxf(o, 0.f);
        if (accumulate) out[t * out_ld + n0 + r] += o; else out[t * out_ld + n0 + r] = o;
      }
    } else {
#pragma unroll
      for (int r = 0; r < R; ++r)
#pragma unroll
        for (int t = 0; t < T; ++t)
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) acc[r][t] += __shfl_xor_sync(0xffffffffu, acc[r][t], o);
      if (lane == 0) {
#pragma unroll
        for (int r = 0; r < R; ++r) {
          if (n0 + r < N) {
#pragma unroll
            for (int t = 0; t < T; ++t) {
              float v = acc[r][t] + (bias ? bias[n0 + r] : 0.f);
              if (relu) v = fmaxf(v, 0.f);
              if (accumulate) out[t * out_ld + n0 + r] += v; else out[t * out_ld + n0 + r] = v;
            }
          }
        }
      }
    }
  }
}

// nn.LayerNorm(256, eps 1e-5) on each of the T rows of x (shared memory), in place. One warp per row.
__device__ __forceinline__ void layernorm_tok(float* x, int T, const float* __restrict__ w, const float* __restrict__ b) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp < T) {
    float v[8], s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) { v[i] = x[warp * E + lane + 32 * i]; s += v[i]; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    const float mean = s * (1.f / E);
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) { const float d = v[i] - mean; q += d * d; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
    const float rstd = rsqrtf(q * (1.f / E) + 1e-5f);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int c = lane + 32 * i;
      x[warp * E + c] = (v[i] - mean) * rstd * w[c] + b[c];
    }
  }
}

// Self-attention over the 7 tokens, 8 heads of 32 (SA/modeling/transformer.py:218-240). q,k,v: [7][256] in smem;
// result overwrites q.  sc: scratch [8][7][8].
__device__ __forceinline__ void self_attn_tok(float* q, const float* k, const float* v, float* sc) {
  const int tid = threadIdx.x;
  if (tid < 8 * NT * NT) {
    const int h = tid / (NT * NT), i = (tid / NT) % NT, j = tid % NT;
    float s = 0.f;
#pragma unroll
    for (int d = 0; d < 32; ++d) s += q[i * E + h * 32 + d] * k[j * E + h * 32 + d];
    sc[(h * NT + i) * 8 + j] = s * 0.17677669529663687f;  // 1/sqrt(32)
  }
  __syncthreads();
  if (tid < 8 * NT) {
    float* r = sc + tid * 8;
    float m = r[0];
#pragma unroll
    for (int j = 1; j < NT; ++j) m = fmaxf(m, r[j]);
    float e[NT], s = 0.f;
#pragma unroll
    for (int j = 0; j < NT; ++j) { e[j] = expf(r[j] - m); s += e[j]; }
    const float inv = 1.f / s;
#pragma unroll
    for (int j = 0; j < NT; ++j) r[j] = e[j] * inv;
  }
  __syncthreads();
  for (int idx = tid; idx < NT * E; idx += blockDim.x) {
    const int i = idx / E, c = idx % E, h = c / 32;
    float o = 0.f;
#pragma unroll
    for (int j = 0; j < NT; ++j) o += sc[(h * NT + i) * 8 + j] * v[j * E + c];
    q[idx] = o;  // each (i, c) reads only scores, never q: safe in place
  }
  __syncthreads();
}

// random-Fourier positional encoding of a point in [0,1]^2 (SA/modeling/prompt_encoder.py:185-192), channel c of 256
__device__ __forceinline__ float pe_channel(const float* __restrict__ gauss, float x01, float y01, int c) {
  const int k = c & 127;
  float v = (2.f * x01 - 1.f) * gauss[k] + (2.f * y01 - 1.f) * gauss[128 + k];
  v *= 6.283185307179586f;
  return c < 128 ? sinf(v) : cosf(v);
}

// ------------------------------------------------------------------------------------------------ token kernel
// One CTA per box.  phase:
//   0  build tokens from the box, layer-0 self-attention (replaces queries, no PE), norm1, q of token->image attn
//   1  layer-1 self-attention (with PE, residual), norm1, q of token->image attn
//   2  after token->image attn of layer `layer`: out_proj + residual, norm2, MLP (ReLU) + residual, norm3,
//      then k and v of image->token attn
//   3  q of the final token->image attn
//   4  after the final attn: out_proj + residual, norm_final, hypernetwork MLP of mask token 0, IoU head
__global__ void __launch_bounds__(TOK_THREADS, 1) decoder_token_kernel(const DecWeights w, const DecState s, int phase, int layer) {
  extern __shared__ float sm[];
  float* Q = sm;                 // [7][256] queries
  float* PE = Q + NT * E;        // [7][256] query_pe (= the initial tokens)
  float* A = PE + NT * E;        // scratch [7][256]
  float* Bq = A + NT * E;
  float* Cq = Bq + NT * E;
  float* H = Cq + NT * E;        // [7][2048] MLP hidden / misc scratch
  const int box = blockIdx.x, tid = threadIdx.x;
  float* gQ = s.queries + (size_t)box * NT * E;
  float* gPE = s.query_pe + (size_t)box * NT * E;

  if (phase == 0) {
    const float4 bx = *reinterpret_cast<const float4*>(s.boxes + box * 4);
    for (int idx = tid; idx < NT * E; idx += blockDim.x) {
      const int t = idx / E, c = idx % E;
      float v;
      if (t == 0) v = w.iou_token[c];
      else if (t < 5) v = w.mask_tokens[(t - 1) * E + c];
      else {  // box corner t-5: (x0,y0) / (x1,y1), +0.5 then / img_size (prompt_encoder.py:93-100,207-214)
        const float px = ((t == 5 ? bx.x : bx.z) + 0.5f) * (1.f / 1024.f);
        const float py = ((t == 5 ? bx.y : bx.w) + 0.5f) * (1.f / 1024.f);
        v = pe_channel(w.gauss, px, py, c) + (t == 5 ? w.corner0[c] : w.corner1[c]);
      }
      Q[idx] = v; PE[idx] = v; gPE[idx] = v;
    }
    __syncthreads();
  } else {
    for (int idx = tid; idx < NT * E; idx += blockDim.x) { Q[idx] = gQ[idx]; PE[idx] = gPE[idx]; }
    __syncthreads();
  }

  if (phase == 0 || phase == 1) {
    const DecLayerW& L = w.layer[phase];
    float* QQ = H;  // queries + pe for layer 1
    if (phase == 1) {
      for (int idx = tid; idx < NT * E; idx += blockDim.x) QQ[idx] = Q[idx] + PE[idx];
      __syncthreads();
    }
    const float* qin = phase == 1 ? QQ : Q;
    linear_tok<NT>(A, E, qin, E, L.sa_q_w, L.sa_q_b, E, E, false, false);
    linear_tok<NT>(Bq, E, qin, E, L.sa_k_w, L.sa_k_b, E, E, false, false);
    linear_tok<NT>(Cq, E, Q, E, L.sa_v_w, L.sa_v_b, E, E, false, false);
    __syncthreads();
    self_attn_tok(A, Bq, Cq, H + NT * E);
    // layer 0: queries = out_proj(attn) (skip_first_layer_pe, transformer.py:155-156); layer 1: residual
    linear_tok<NT>(Q, E, A, E, L.sa_o_w, L.sa_o_b, E, E, false, phase == 1);
    __syncthreads();
    layernorm_tok(Q, NT, L.n1_w, L.n1_b);
    __syncthreads();
    for (int idx = tid; idx < NT * E; idx += blockDim.x) { gQ[idx] = Q[idx]; A[idx] = Q[idx] + PE[idx]; }
    __syncthreads();
    linear_tok<NT>(Bq, 128, A, E, L.t2i_q_w, L.t2i_q_b, 128, E, false, false);
    __syncthreads();
    for (int idx = tid; idx < NT * 128; idx += blockDim.x) s.q_t2i[(size_t)box * NT * 128 + idx] = Bq[idx];
  } else if (phase == 2) {
    const DecLayerW& L = w.layer[layer];
    for (int idx = tid; idx < NT * 128; idx += blockDim.x) A[idx] = s.attn_t2i[(size_t)box * NT * 128 + idx];
    __syncthreads();
    linear_tok<NT>(Q, E, A, 128, L.t2i_o_w, L.t2i_o_b, E, 128, false, true);
    __syncthreads();
    layernorm_tok(Q, NT, L.n2_w, L.n2_b);
    __syncthreads();
    linear_tok<NT>(H, 2048, Q, E, L.mlp1_w, L.mlp1_b, 2048, E, true, false);
    __syncthreads();
    linear_tok<NT>(Q, E, H, 2048, L.mlp2_w, L.mlp2_b, E, 2048, false, true);
    __syncthreads();
    layernorm_tok(Q, NT, L.n3_w, L.n3_b);
    __syncthreads();
    for (int idx = tid; idx < NT * E; idx += blockDim.x) { gQ[idx] = Q[idx]; A[idx] = Q[idx] + PE[idx]; }
    __syncthreads();
    linear_tok<NT>(Bq, 128, A, E, L.i2t_k_w, L.i2t_k_b, 128, E, false, false);
    linear_tok<NT>(Cq, 128, Q, E, L.i2t_v_w, L.i2t_v_b, 128, E, false, false);
    __syncthreads();
    float* kv = s.kv_i2t + (size_t)box * 2 * NT * 128;
    for (int idx = tid; idx < NT * 128; idx += blockDim.x) { kv[idx] = Bq[idx]; kv[NT * 128 + idx] = Cq[idx]; }
  } else if (phase == 3) {
    for (int idx = tid; idx < NT * E; idx += blockDim.x) A[idx] = Q[idx] + PE[idx];
    __syncthreads();
    linear_tok<NT>(Bq, 128, A, E, w.fin_q_w, w.fin_q_b, 128, E, false, false);
    __syncthreads();
    for (int idx = tid; idx < NT * 128; idx += blockDim.x) s.q_t2i[(size_t)box * NT * 128 + idx] = Bq[idx];
  } else {
    for (int idx = tid; idx < NT * 128; idx += blockDim.x) A[idx] = s.attn_t2i[(size_t)box * NT * 128 + idx];
    __syncthreads();
    linear_tok<NT>(Q, E, A, 128, w.fin_o_w, w.fin_o_b, E, 128, false, true);
    __syncthreads();
    layernorm_tok(Q, NT, w.nf_w, w.nf_b);
    __syncthreads();
    // multimask_output=False keeps mask 0 / iou 0 only (mask_decoder.py:105-107): token 1 -> hypernetwork 0, token 0 -> IoU
    linear_tok<1>(A, E, Q + 1 * E, E, w.hyp_w[0], w.hyp_b[0], E, E, true, false);
    linear_tok<1>(Bq, E, Q, E, w.iou_w[0], w.iou_b[0], E, E, true, false);
    __syncthreads();
    linear_tok<1>(Cq, E, A, E, w.hyp_w[1], w.hyp_b[1], E, E, true, false);
    linear_tok<1>(Cq + E, E, Bq, E, w.iou_w[1], w.iou_b[1], E, E, true, false);
    __syncthreads();
    linear_tok<1>(A, 32, Cq, E, w.hyp_w[2], w.hyp_b[2], 32, E, false, false);
    linear_tok<1>(Bq, 4, Cq + E, E, w.iou_w[2], w.iou_b[2], 4, E, false, false);
    __syncthreads();
    if (tid < 32) s.hyper[box * 32 + tid] = A[tid];
    if (tid == 0) s.iou[box] = Bq[0];
  }
}

int launch_decoder_tokens(const DecWeights& w, const DecState& s, int boxes, int phase, int layer, cudaStream_t st) {
  constexpr int SMEM = (5 * NT * E + NT * 2048) * 4;
  static PerDeviceFlag attr_done;
  if (!attr_done()) {
    CUDA_TRY(cudaFuncSetAttribute(decoder_token_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
    attr_done() = true;
  }
  decoder_token_kernel<<<boxes, TOK_THREADS, SMEM, st>>>(w, s, phase, layer);
  CUDA_TRY(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------------ token -> image
// 7 queries x 4096 keys per (box, head), internal dim 128 = 8 heads x 16 (downsample_rate 2, transformer.py:204-206).
// kv rows: [4096][ld] fp32 with k at column 0 and v at column 128 (+ head*16); box_stride 0 when all boxes share keys.
// box_img != null: the keys of box b are those of image box_img[b] (kv + box_img[b] * box_stride), shared by its boxes.
//
// HBM-bound (the per-box keys are read once: 4 MB per box and launch).  A CTA takes 512 consecutive keys of one box for
// ALL heads; thread = (key row of an 8-row pass, head, QUARTER of the head's 16 dims): a warp reads one key's 512
// contiguous bytes of k and of v with one 16-byte load per lane, the thread's 4 dims of the 7 queries and of the 7
// accumulators live in registers, and the four partial dot products of a head meet through two shuffles.  (The first
// version had thread = (key, head): q re-read from shared memory for every key and 16-byte loads at a 64-byte stride kept
// the L1 data pipe 75 % busy -- 243 k wavefronts per SM, half of them shared-memory reads -- at a third of the HBM peak, and
// the launch over L2-resident shared keys took as long as the ones over per-box keys.)  The 8 key ranges of a box are
// merged by the last CTA to finish (ticket counter), in fixed order, so the result does not depend on scheduling.
__device__ __forceinline__ float fast_exp2(float x) {  // ex2.approx: one MUFU instruction, 2 ulp; exact 0 for x = -inf
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// 4-term dot product and 4-wide axpy on fma.rn.f32x2 (the attention kernels below are issue-bound)
__device__ __forceinline__ float dot4_f2(const float4& a, const float4& b) {
  float lo, hi;
  f2_unpack(f2_fma(f2_pack(a.z, a.w), f2_pack(b.z, b.w), f2_mul(f2_pack(a.x, a.y), f2_pack(b.x, b.y))), lo, hi);
  return lo + hi;
}
__device__ __forceinline__ void axpy4_f2(float4& acc, float p, const float4& v) {
  const uint64_t p2 = f2_pack(p, p);
  f2_unpack(f2_fma(p2, f2_pack(v.x, v.y), f2_pack(acc.x, acc.y)), acc.x, acc.y);
  f2_unpack(f2_fma(p2, f2_pack(v.z, v.w), f2_pack(acc.z, acc.w)), acc.z, acc.w);
}

constexpr int T2I_SPLITS = 8;                       // key ranges per box
constexpr int T2I_PART = 8 * NT * 18;               // floats of one partial: [head][token][16 acc, l, m]

// 4 consecutive elements of a k|v|q row for this lane: fp32 (16 bytes) or fp16 (8 bytes; GemmArgs::out_f16).  The prefetch
// queues of the kernels below hold the raw loaded words (and are twice as deep for fp16, so that the bytes in flight per SM
// stay the same: with half the bytes per load the fp16 variants were latency-bound at 1.3 TB/s).
template <bool HALF> struct Row4 { using Raw = float4; static constexpr int DEPTH = 2; };
template <> struct Row4<true> { using Raw = uint2; static constexpr int DEPTH = 4; };
template <bool HALF>
__device__ __forceinline__ typename Row4<HALF>::Raw load_row4(const void* row, int idx4) {
  return *(reinterpret_cast<const typename Row4<HALF>::Raw*>(row) + idx4);
}
__device__ __forceinline__ float4 cvt_row4(const float4& r) { return r; }
__device__ __forceinline__ float4 cvt_row4(const uint2& u) {
  const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&u.x));
  const float2 b = __half22float2(*reinterpret_cast<const __half2*>(&u.y));
  return make_float4(a.x, a.y, b.x, b.y);
}

template <bool HALF>
__global__ void __launch_bounds__(256) decoder_t2i_kernel(const float* __restrict__ q, const void* __restrict__ kv_,
                                                          long long box_stride, int ld, const int* __restrict__ box_img,
                                                          float* __restrict__ out, float* __restrict__ partial,
                                                          int* __restrict__ ticket) {
  constexpr int ES = HALF ? 2 : 4;  // bytes per element of the kv rows (box_stride and ld are in elements)
  const int box = blockIdx.y, split = blockIdx.x, tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31, h = lane >> 2, qt = lane & 3;  // lane = float4 index inside a 128-float row
  __shared__ float red_m[8][8 * NT];
  __shared__ float red[8][8 * NT * 17];
  __shared__ int last;
  float4 q4[NT], acc[NT];
  float m[NT], l[NT];
#pragma unroll
  for (int t = 0; t < NT; ++t) {  // 1/sqrt(16) and log2(e) folded into q
    q4[t] = __ldg(reinterpret_cast<const float4*>(q + (size_t)box * NT * 128 + t * 128) + lane);
    q4[t].x *= 0.25f * 1.4426950408889634f; q4[t].y *= 0.25f * 1.4426950408889634f;
    q4[t].z *= 0.25f * 1.4426950408889634f; q4[t].w *= 0.25f * 1.4426950408889634f;
    acc[t] = make_float4(0.f, 0.f, 0.f, 0.f);
    m[t] = 0.f; l[t] = 0.f;
  }
  const char* base = static_cast<const char*>(kv_) + ((size_t)(box_img ? box_img[box] : box) * box_stride + (size_t)split * 512 * ld) * ES;
  // 64 passes of 8 keys (one per warp) with the next two passes in flight.  The softmax shift of a token is lazy: it is set
  // by the first key and moves (with a rescale of l and acc) only when a later score exceeds it by more than 2^24 -- tested
  // once per key for all 7 tokens -- so the common path is one ex2.approx and 9 FMAs per (token, key) and thread.
  constexpr int PD = Row4<HALF>::DEPTH;
  typename Row4<HALF>::Raw nk[PD], nv[PD];
#pragma unroll
  for (int i = 0; i < PD; ++i) {
    const char* r = base + (size_t)(warp + 8 * i) * ld * ES;
    nk[i] = load_row4<HALF>(r, lane); nv[i] = load_row4<HALF>(r, 32 + lane);
  }
#pragma unroll
  for (int t = 0; t < NT; ++t) {  // the shift of a token starts at its score with this warp's first key
    const float4 k0 = cvt_row4(nk[0]);
    float x = dot4_f2(q4[t], k0);
    x += __shfl_xor_sync(0xffffffffu, x, 1);
    x += __shfl_xor_sync(0xffffffffu, x, 2);
    m[t] = x;
  }
#pragma unroll PD
  for (int i = 0; i < 64; ++i) {
    const float4 kk = cvt_row4(nk[i % PD]), vv = cvt_row4(nv[i % PD]);
    if (i + PD < 64) {
      const char* r = base + (size_t)(warp + 8 * (i + PD)) * ld * ES;
      nk[i % PD] = load_row4<HALF>(r, lane); nv[i % PD] = load_row4<HALF>(r, 32 + lane);
    }
    float sc[NT], over = -INFINITY;
#pragma unroll
    for (int t = 0; t < NT; ++t) {
      float x = dot4_f2(q4[t], kk);
      x += __shfl_xor_sync(0xffffffffu, x, 1);
      x += __shfl_xor_sync(0xffffffffu, x, 2);  // the four lanes of a head hold the same score
      sc[t] = x - m[t];
      over = fmaxf(over, sc[t]);
    }
    if (over > 24.f) {  // one test per key (cold): a score far above some token's shift -> move that shift there
#pragma unroll
      for (int t = 0; t < NT; ++t) {
        if (sc[t] > 24.f) {
          const float corr = fast_exp2(-sc[t]);
          m[t] += sc[t];
          l[t] *= corr;
          acc[t].x *= corr; acc[t].y *= corr; acc[t].z *= corr; acc[t].w *= corr;
          sc[t] = 0.f;
        }
      }
    }
#pragma unroll
    for (int t = 0; t < NT; ++t) {
      const float p = fast_exp2(sc[t]);
      l[t] += p;
      axpy4_f2(acc[t], p, vv);
    }
  }
  // merge the 8 warps (each holds the softmax state of its 64 keys for every head and token)
  if (qt == 0) {
#pragma unroll
    for (int t = 0; t < NT; ++t) red_m[warp][h * NT + t] = m[t];
  }
  __syncthreads();
#pragma unroll
  for (int t = 0; t < NT; ++t) {
    float gm = red_m[0][h * NT + t];
#pragma unroll
    for (int w2 = 1; w2 < 8; ++w2) gm = fmaxf(gm, red_m[w2][h * NT + t]);
    const float sc = exp2f(m[t] - gm);
    float* r = &red[warp][(h * NT + t) * 17];
    r[qt * 4] = acc[t].x * sc; r[qt * 4 + 1] = acc[t].y * sc; r[qt * 4 + 2] = acc[t].z * sc; r[qt * 4 + 3] = acc[t].w * sc;
    if (qt == 0) r[16] = l[t] * sc;
  }
  __syncthreads();
  // this key range's partial: [head][token][16 acc, l, m]
  float* part = partial + ((size_t)box * T2I_SPLITS + split) * T2I_PART;
  for (int i = tid; i < 8 * NT * 17; i += 256) {
    float a = 0.f;
#pragma unroll
    for (int w2 = 0; w2 < 8; ++w2) a += red[w2][i];
    part[(i / 17) * 18 + i % 17] = a;
  }
  if (tid < 8 * NT) {
    float v = red_m[0][tid];
#pragma unroll
    for (int w2 = 1; w2 < 8; ++w2) v = fmaxf(v, red_m[w2][tid]);
    part[tid * 18 + 17] = v;
  }
  __threadfence();
  __syncthreads();
  if (tid == 0) {
    const int tk = atomicAdd(ticket + box, 1);
    last = (tk == T2I_SPLITS - 1);
    if (last) ticket[box] = 0;  // ready for the next launch
  }
  __syncthreads();
  if (!last) return;
  __threadfence();
  const float* pb = partial + (size_t)box * T2I_SPLITS * T2I_PART;
  for (int i = tid; i < 8 * NT * 16; i += 256) {
    const int ht = i >> 4, d = i & 15;  // ht = head * NT + token
    float M = -INFINITY;
#pragma unroll
    for (int sp = 0; sp < T2I_SPLITS; ++sp) M = fmaxf(M, __ldcg(pb + sp * T2I_PART + ht * 18 + 17));
    float a = 0.f, ls = 0.f;
#pragma unroll
    for (int sp = 0; sp < T2I_SPLITS; ++sp) {
      const float* pp = pb + sp * T2I_PART + ht * 18;
      const float w = exp2f(__ldcg(pp + 17) - M);
      a += __ldcg(pp + d) * w;
      ls += __ldcg(pp + 16) * w;
    }
    const int hh = ht / NT, t = ht % NT;
    out[(size_t)box * NT * 128 + t * 128 + hh * 16 + d] = a / ls;
  }
}

// partial: boxes * T2I_SPLITS * T2I_PART floats of scratch; ticket: boxes ints, zero before the first launch (the
// kernel leaves them zero).
int launch_decoder_t2i(const float* q, const void* kv, int kv_half, long long box_stride, int ld, const int* box_img, float* out,
                       float* partial, int* ticket, int boxes, cudaStream_t st) {
  if (kv_half) decoder_t2i_kernel<true><<<dim3(T2I_SPLITS, boxes), 256, 0, st>>>(q, kv, box_stride, ld, box_img, out, partial, ticket);
  else decoder_t2i_kernel<false><<<dim3(T2I_SPLITS, boxes), 256, 0, st>>>(q, kv, box_stride, ld, box_img, out, partial, ticket);
  CUDA_TRY(cudaGetLastError());
  return 0;
}

size_t decoder_t2i_scratch_floats(int boxes) { return (size_t)boxes * T2I_SPLITS * T2I_PART; }

// ------------------------------------------------------------------------------------------------ image -> token
// Every image token attends to the 7 tokens (8 heads x 16). q: column `qcol` of the kvq rows (stride ld, box_stride 0
// when shared); kv: [box][2][7][128]. out: [box*4096][128] fp32, the A operand of the out_proj GEMM.
//
// HBM-bound by design (512 B read + 512 B written per pixel), but the first version -- thread = (pixel, head), k / v of the
// 7 tokens re-read from shared memory for every pixel (56 LDS.128 per thread and pixel) and 16-byte global accesses at a
// 64-byte stride -- kept the L1 data pipe 80 % busy (281 k wavefronts per SM, ncu) and reached a third of the HBM peak.
// Now thread = (pixel, head, QUARTER of the head's 16 dims): the thread's 4 dims of k and v of the 7 tokens live in registers
// for the whole block (no shared memory in the loop), a warp reads / writes one pixel's 512 contiguous bytes with one
// 16-byte access per lane, and the four partial dot products of a head meet through two shuffles.
constexpr int I2T_PIX = 512;  // pixels per block

template <bool HALF, bool OUT_HALF>
__global__ void __launch_bounds__(256) decoder_i2t_kernel(const void* __restrict__ qsrc, long long box_stride, int ld, int qcol,
                                                          const int* __restrict__ box_img, const float* __restrict__ kv,
                                                          void* __restrict__ out) {
  constexpr int ES = HALF ? 2 : 4;
  const int box = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  // lane = head * 4 + quarter = float4 index inside the 128-float row
  float4 k4[NT], v4[NT];
  {
    const float4* kp = reinterpret_cast<const float4*>(kv + (size_t)box * 2 * NT * 128);
#pragma unroll
    for (int t = 0; t < NT; ++t) {
      k4[t] = __ldg(kp + t * 32 + lane);
      v4[t] = __ldg(kp + NT * 32 + t * 32 + lane);
      k4[t].x *= 0.25f * 1.4426950408889634f; k4[t].y *= 0.25f * 1.4426950408889634f;  // 1/sqrt(16) and log2(e) folded into k
      k4[t].z *= 0.25f * 1.4426950408889634f; k4[t].w *= 0.25f * 1.4426950408889634f;
    }
  }
  const char* qbase = static_cast<const char*>(qsrc) + ((size_t)(box_img ? box_img[box] : box) * box_stride + qcol) * ES;
  const int pix0 = blockIdx.x * I2T_PIX + warp;
  // out: [box*4096][128] fp32, or fp16 when the out_proj GEMM that consumes it takes fp16 operands (GemmArgs::ab_f16)
  char* obase = static_cast<char*>(out) + (size_t)box * 4096 * 128 * (OUT_HALF ? 2 : 4);
  constexpr int NP = I2T_PIX / 8;  // pixels per warp
  constexpr int PD = Row4<HALF>::DEPTH;
  typename Row4<HALF>::Raw qn[PD];
#pragma unroll
  for (int i = 0; i < PD; ++i) qn[i] = load_row4<HALF>(qbase + (size_t)(pix0 + 8 * i) * ld * ES, lane);
#pragma unroll PD
  for (int i = 0; i < NP; ++i) {
    const int pix = pix0 + i * 8;
    const float4 q = cvt_row4(qn[i % PD]);
    if (i + PD < NP) qn[i % PD] = load_row4<HALF>(qbase + (size_t)(pix + 8 * PD) * ld * ES, lane);
    float sc[NT], mx = -INFINITY;
#pragma unroll
    for (int t = 0; t < NT; ++t) {
      float x = dot4_f2(q, k4[t]);
      x += __shfl_xor_sync(0xffffffffu, x, 1);
      x += __shfl_xor_sync(0xffffffffu, x, 2);  // all four lanes of the head hold the same sum
      sc[t] = x;
      mx = fmaxf(mx, x);
    }
    float sum = 0.f;
#pragma unroll
    for (int t = 0; t < NT; ++t) { sc[t] = fast_exp2(sc[t] - mx); sum += sc[t]; }
    const float inv = __fdividef(1.f, sum);
    float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int t = 0; t < NT; ++t) {
      axpy4_f2(o, sc[t] * inv, v4[t]);
    }
    if constexpr (OUT_HALF) {
      uint2 h2;
      h2.x = pack_f16x2_sat(o.x, o.y); h2.y = pack_f16x2_sat(o.z, o.w);
      reinterpret_cast<uint2*>(obase)[(size_t)pix * 32 + lane] = h2;
    } else {
      reinterpret_cast<float4*>(obase)[(size_t)pix * 32 + lane] = o;
    }
  }
}

int launch_decoder_i2t(const void* qsrc, int q_half, long long box_stride, int ld, int qcol, const int* box_img, const float* kv,
                       void* out, int out_half, int boxes, cudaStream_t st) {
  const dim3 grid(4096 / I2T_PIX, boxes);
  if (q_half && out_half) decoder_i2t_kernel<true, true><<<grid, 256, 0, st>>>(qsrc, box_stride, ld, qcol, box_img, kv, out);
  else if (q_half) decoder_i2t_kernel<true, false><<<grid, 256, 0, st>>>(qsrc, box_stride, ld, qcol, box_img, kv, out);
  else if (out_half) decoder_i2t_kernel<false, true><<<grid, 256, 0, st>>>(qsrc, box_stride, ld, qcol, box_img, kv, out);
  else decoder_i2t_kernel<false, false><<<grid, 256, 0, st>>>(qsrc, box_stride, ld, qcol, box_img, kv, out);
  CUDA_TRY(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------------ keys / PE
// keys0[tok][c] = embedding[c][tok] + no_mask_embed[c]   (mask_decoder.py:126-127 with the dense prompt of
// prompt_encoder.py:164-166); transposes NCHW -> token-major through a padded shared tile.
__global__ void __launch_bounds__(256) decoder_keys0_kernel(const float* __restrict__ emb, const float* __restrict__ nomask,
                                                            float* __restrict__ keys) {
  __shared__ float tile[32][33];
  const int t0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  emb += (size_t)blockIdx.z * E * 4096;
  keys += (size_t)blockIdx.z * 4096 * E;
  for (int i = ty; i < 32; i += 8) tile[i][tx] = emb[(size_t)(c0 + i) * 4096 + t0 + tx];
  __syncthreads();
  for (int i = ty; i < 32; i += 8) keys[(size_t)(t0 + i) * E + c0 + tx] = tile[tx][i] + nomask[c0 + tx];
}

// emb: [B][256][4096] (NCHW embeddings); keys: [B][4096][256]
int launch_decoder_keys0(const float* emb, const float* nomask, float* keys, int B, cudaStream_t st) {
  decoder_keys0_kernel<<<dim3(128, 8, B), 256, 0, st>>>(emb, nomask, keys);
  CUDA_TRY(cudaGetLastError());
  return 0;
}

// box_img[k] = image that box k belongs to; boxes are grouped by image, image i owns boxes [start[i], start[i+1]).
__global__ void box_image_index_kernel(const BoxSegments seg, int* __restrict__ box_img) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= seg.start[seg.n]) return;
  int i = 0;
  while (i + 1 < seg.n && k >= seg.start[i + 1]) ++i;
  box_img[k] = i;
}

int launch_box_image_index(const BoxSegments& seg, int* box_img, cudaStream_t st) {
  const int total = seg.start[seg.n];
  if (total <= 0) return 0;
  box_image_index_kernel<<<(total + 255) / 256, 256, 0, st>>>(seg, box_img);
  CUDA_TRY(cudaGetLastError());
  return 0;
}

// dense positional encoding of the 64x64 grid at pixel centres, token-major [4096][256] (prompt_encoder.py:62-71,194-205)
__global__ void __launch_bounds__(256) dense_pe_kernel(const float* __restrict__ gauss, float* __restrict__ pe) {
  const int tok = blockIdx.x, c = threadIdx.x;
  const float x = ((tok & 63) + 0.5f) * (1.f / 64.f), y = ((tok >> 6) + 0.5f) * (1.f / 64.f);
  pe[(size_t)tok * E + c] = pe_channel(gauss, x, y, c);
}

int launch_dense_pe(const float* gauss, float* pe, cudaStream_t st) {
  dense_pe_kernel<<<4096, 256, 0, st>>>(gauss, pe);
  CUDA_TRY(cudaGetLastError());
  return 0;
}

}  // namespace inksam
