// Mask-decoder kernels that are not GEMMs: the 7-token side of the two-way transformer (fp32 on CUDA cores, one
// CTA per box), token->image and image->token attention over the 4096 image tokens, prompt / dense positional
// encodings.  The image-side linears (k/v/q projections of the 4096 keys, out_proj + norm4, the two transposed
// convolutions) run on the tcgen05 GEMM in tf32 (gemm_tc.cuh); see engine.cu for the order.
//
// Reference: SA/modeling/transformer.py:62-106,151-182,218-240; mask_decoder.py:112-149,154-176;
// prompt_encoder.py:62-71,93-100,185-214.
#include "common.h"
#include "decoder_ops.h"

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
    float acc[R][T];
#pragma unroll
    for (int r = 0; r < R; ++r)
#pragma unroll
      for (int t = 0; t < T; ++t) acc[r][t] = 0.f;
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
#pragma unroll
            for (int r = 0; r < R; ++r) acc[r][t] += a[u][r].x * x.x + a[u][r].y * x.y + a[u][r].z * x.z + a[u][r].w * x.w;
          }
        }
      }
    }
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
// ALL heads: thread = (key row of a 32-row pass, head), so the 8 threads of a key read its 512 contiguous bytes of k
// and of v and the CTA streams one contiguous 512-row range (a CTA per (box, head) read 64-byte slices at a 1-1.5 KB
// stride and sat at a third of the HBM peak on DRAM page misses).  The 8 key ranges of a box are merged by the last
// CTA to finish (ticket counter), in fixed order, so the result does not depend on scheduling.
constexpr int T2I_SPLITS = 8;                       // key ranges per box
constexpr int T2I_PART = 8 * NT * 18;               // floats of one partial: [head][token][16 acc, l, m]
constexpr int T2I_QP = 20;                          // padded head row of q in shared memory (conflict-free LDS.128)

__global__ void __launch_bounds__(256) decoder_t2i_kernel(const float* __restrict__ q, const float* __restrict__ kv,
                                                          long long box_stride, int ld, const int* __restrict__ box_img,
                                                          float* __restrict__ out, float* __restrict__ partial,
                                                          int* __restrict__ ticket) {
  const int box = blockIdx.y, split = blockIdx.x, tid = threadIdx.x;
  const int h = tid & 7, kr = tid >> 3, warp = tid >> 5, lane = tid & 31;
  __shared__ __align__(16) float qs[NT * 8 * T2I_QP];
  __shared__ float red_m[8][8 * NT];
  __shared__ float red[8][8 * NT * 17];
  __shared__ int last;
  for (int i = tid; i < NT * 128; i += 256) {  // 1/sqrt(16) and log2(e) folded into q
    const int t = i >> 7, c = i & 127;
    qs[(t * 8 + (c >> 4)) * T2I_QP + (c & 15)] = q[(size_t)box * NT * 128 + i] * (0.25f * 1.4426950408889634f);
  }
  __syncthreads();
  const float* base = kv + (size_t)(box_img ? box_img[box] : box) * box_stride + (size_t)split * 512 * ld + h * 16;
  const uint32_t qs_addr = (uint32_t)__cvta_generic_to_shared(&qs[0]) + (uint32_t)(h * T2I_QP * 4);
  float m[NT], l[NT], acc[NT][16];
#pragma unroll
  for (int t = 0; t < NT; ++t) {
    m[t] = -INFINITY; l[t] = 0.f;
#pragma unroll
    for (int d = 0; d < 16; ++d) acc[t][d] = 0.f;
  }
  // 16 passes of 32 keys with the next pass already in flight.  The softmax shift of a token is lazy: it is set by the
  // first key and moves (with an exact rescale of l and acc) only when a later score exceeds it by more than 2^24, so
  // the common path is one exp2 and 17 FMAs per (token, key).
  float4 nk[4], nv[4];
  {
    const float4* r = reinterpret_cast<const float4*>(base + (size_t)kr * ld);
#pragma unroll
    for (int i = 0; i < 4; ++i) { nk[i] = r[i]; nv[i] = r[32 + i]; }
  }
#pragma unroll 1
  for (int j = kr; j < 512; j += 32) {
    float kk[16], vv[16];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      kk[4 * i] = nk[i].x; kk[4 * i + 1] = nk[i].y; kk[4 * i + 2] = nk[i].z; kk[4 * i + 3] = nk[i].w;
      vv[4 * i] = nv[i].x; vv[4 * i + 1] = nv[i].y; vv[4 * i + 2] = nv[i].z; vv[4 * i + 3] = nv[i].w;
    }
    if (j + 32 < 512) {
      const float4* r = reinterpret_cast<const float4*>(base + (size_t)(j + 32) * ld);
#pragma unroll
      for (int i = 0; i < 4; ++i) { nk[i] = r[i]; nv[i] = r[32 + i]; }
    }
#pragma unroll
    for (int t = 0; t < NT; ++t) {
      float sc = 0.f;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        // q stays in shared memory (LDS.128, re-read per key): in registers it would be 112 more live values on top
        // of the 112 accumulators
        float4 q4;
        asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                     : "=f"(q4.x), "=f"(q4.y), "=f"(q4.z), "=f"(q4.w)
                     : "r"(qs_addr + (uint32_t)((t * 8 * T2I_QP + 4 * i) * 4)));
        sc += q4.x * kk[4 * i] + q4.y * kk[4 * i + 1] + q4.z * kk[4 * i + 2] + q4.w * kk[4 * i + 3];
      }
      if (sc > m[t] + 24.f) {  // first key (m = -inf) or a much larger score: move the shift
        const float corr = exp2f(m[t] - sc);  // 0 for the first key
        m[t] = sc;
        l[t] *= corr;
#pragma unroll
        for (int d = 0; d < 16; ++d) acc[t][d] *= corr;
      }
      const float p = exp2f(sc - m[t]);
      l[t] += p;
#pragma unroll
      for (int d = 0; d < 16; ++d) acc[t][d] += p * vv[d];
    }
  }
  // merge the 32 partial softmaxes of every head: the lanes of a warp with equal (lane & 7) share a head
  float gm[NT];
#pragma unroll
  for (int t = 0; t < NT; ++t) {
    float v = m[t];
    v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 8));
    v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, 16));
    if (lane < 8) red_m[warp][h * NT + t] = v;
  }
  __syncthreads();
#pragma unroll
  for (int t = 0; t < NT; ++t) {
    float v = red_m[0][h * NT + t];
#pragma unroll
    for (int w2 = 1; w2 < 8; ++w2) v = fmaxf(v, red_m[w2][h * NT + t]);
    gm[t] = v;
  }
#pragma unroll
  for (int t = 0; t < NT; ++t) {
    const float sc = exp2f(m[t] - gm[t]);
    float v = l[t] * sc;
    v += __shfl_xor_sync(0xffffffffu, v, 8);
    v += __shfl_xor_sync(0xffffffffu, v, 16);
    if (lane < 8) red[warp][(h * NT + t) * 17 + 16] = v;
#pragma unroll
    for (int d = 0; d < 16; ++d) {
      float a = acc[t][d] * sc;
      a += __shfl_xor_sync(0xffffffffu, a, 8);
      a += __shfl_xor_sync(0xffffffffu, a, 16);
      if (lane < 8) red[warp][(h * NT + t) * 17 + d] = a;
    }
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
int launch_decoder_t2i(const float* q, const float* kv, long long box_stride, int ld, const int* box_img, float* out,
                       float* partial, int* ticket, int boxes, cudaStream_t st) {
  decoder_t2i_kernel<<<dim3(T2I_SPLITS, boxes), 256, 0, st>>>(q, kv, box_stride, ld, box_img, out, partial, ticket);
  CUDA_TRY(cudaGetLastError());
  return 0;
}

size_t decoder_t2i_scratch_floats(int boxes) { return (size_t)boxes * T2I_SPLITS * T2I_PART; }

// ------------------------------------------------------------------------------------------------ image -> token
// Every image token attends to the 7 tokens (8 heads x 16). q: column `qcol` of the kvq rows (stride ld, box_stride 0
// when shared); kv: [box][2][7][128]. out: [box*4096][128] fp32, the A operand of the out_proj GEMM.
__global__ void __launch_bounds__(256) decoder_i2t_kernel(const float* __restrict__ qsrc, long long box_stride, int ld, int qcol,
                                                          const int* __restrict__ box_img, const float* __restrict__ kv,
                                                          float* __restrict__ out) {
  const int box = blockIdx.y, tid = threadIdx.x;
  // k / v of the 7 tokens, head rows padded 16 -> 20 floats: the 8 lanes of a quarter warp (8 heads of one pixel) then
  // read 8 disjoint 4-bank groups with one LDS.128 each (conflict-free; an unpadded row is a 4-way conflict)
  constexpr int HP = 20;
  __shared__ __align__(16) float ks[NT * 8 * HP], vs[NT * 8 * HP];
  for (int i = tid; i < NT * 128; i += 256) {
    const int t = i >> 7, c = i & 127;
    ks[(t * 8 + (c >> 4)) * HP + (c & 15)] = kv[(size_t)box * 2 * NT * 128 + i];
    vs[(t * 8 + (c >> 4)) * HP + (c & 15)] = kv[(size_t)box * 2 * NT * 128 + NT * 128 + i];
  }
  __syncthreads();
  const int h = tid & 7;
#pragma unroll 1
  for (int rep = 0; rep < 4; ++rep) {  // 4 x 32 pixels per block
    const int pix = blockIdx.x * 128 + rep * 32 + (tid >> 3);
    const float4* qr = reinterpret_cast<const float4*>(qsrc + (size_t)(box_img ? box_img[box] : box) * box_stride +
                                                       (size_t)pix * ld + qcol + h * 16);
    float qv[16];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float4 a = qr[i];
      qv[4 * i] = a.x; qv[4 * i + 1] = a.y; qv[4 * i + 2] = a.z; qv[4 * i + 3] = a.w;
    }
    float sc[NT], mx = -INFINITY;
#pragma unroll
    for (int t = 0; t < NT; ++t) {
      const float4* kr = reinterpret_cast<const float4*>(ks + (t * 8 + h) * HP);
      float s = 0.f;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float4 k4 = kr[i];
        s += qv[4 * i] * k4.x + qv[4 * i + 1] * k4.y + qv[4 * i + 2] * k4.z + qv[4 * i + 3] * k4.w;
      }
      sc[t] = s * 0.25f;
      mx = fmaxf(mx, sc[t]);
    }
    float sum = 0.f;
#pragma unroll
    for (int t = 0; t < NT; ++t) { sc[t] = expf(sc[t] - mx); sum += sc[t]; }
    const float inv = 1.f / sum;
    float o[16];
#pragma unroll
    for (int d = 0; d < 16; ++d) o[d] = 0.f;
#pragma unroll
    for (int t = 0; t < NT; ++t) {
      const float p = sc[t] * inv;
      const float4* vr = reinterpret_cast<const float4*>(vs + (t * 8 + h) * HP);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float4 v4 = vr[i];
        o[4 * i] += p * v4.x; o[4 * i + 1] += p * v4.y; o[4 * i + 2] += p * v4.z; o[4 * i + 3] += p * v4.w;
      }
    }
    float4* dst = reinterpret_cast<float4*>(out + ((size_t)box * 4096 + pix) * 128 + h * 16);
#pragma unroll
    for (int i = 0; i < 4; ++i) dst[i] = make_float4(o[4 * i], o[4 * i + 1], o[4 * i + 2], o[4 * i + 3]);
  }
}

int launch_decoder_i2t(const float* qsrc, long long box_stride, int ld, int qcol, const int* box_img, const float* kv,
                       float* out, int boxes, cudaStream_t st) {
  decoder_i2t_kernel<<<dim3(32, boxes), 256, 0, st>>>(qsrc, box_stride, ld, qcol, box_img, kv, out);
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
