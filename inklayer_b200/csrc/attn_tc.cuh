// Flash-style encoder attention for sm_100a: tcgen05 MMAs with S / P / O in tensor memory, TMA-fed operands,
// decomposed relative-position bias computed on the tensor cores and folded into the softmax.
//
// Replaces Attention.forward + add_decomposed_rel_pos + get_rel_pos of the reference image encoder
// (SA/modeling/image_encoder.py:224-240, 292-361) and, for windowed blocks, window_partition /
// window_unpartition (:243-289), which become TMA box coordinates and store addressing.
//
//   attn_global_kernel   one CTA = 128 queries of one (image, head); loops over 32 key tiles of 128 keys
//   attn_window_kernel   one CTA = 128 query rows of one (image, window, head); a single 208-key tile
//
// Both run 128 threads (thread == query row == TMEM lane) with 2 CTAs resident per SM (256 TMEM columns and
// < 113 KB smem each), so one CTA's softmax (MUFU-bound) overlaps the other's MMAs.  Thread 0 issues TMA and
// tcgen05.mma.  Operand tiles are the head's 64 (+16 for head_dim 80) columns of the bf16 qkv matrix:
// a 128B-swizzled [rows][64] tile and, for head_dim 80, a 32B-swizzled [rows][16] tile.
//
//   S = Q K^T          A, B K-major from smem, D -> TMEM fp32
//   T = Q REL^T        same MMA against the rel_pos table; T[q][r] = q . rel_pos[r] (the UNSCALED q, :234)
//   s = S*scale + T_h[q][qh-kh+S-1] + T_w[q][qw-kw+S-1]   (:354-359), all in log2 units
//   P = exp2(s - m)    bf16, stored back into TMEM over S (tcgen05.st), consumed as the A operand (TS MMA)
//   O += P V           V is the MN-major B operand straight from the [keys][head_dim] tile
//
// The softmax shift m only has to keep exp2 in range (softmax is shift-invariant and fp32/bf16 precision is
// relative), so m is the row maximum of the first 32 keys; the accumulator is rescaled only if a later key
// exceeds it by more than 2^32 (kept exact, never taken in practice).
#pragma once
#include "ptx.cuh"

namespace inksam {

struct AttnArgs {
  int B, heads, D;      // images, heads, embed dim
  float c1;             // head_dim^-0.5 * log2(e)
  __nv_bfloat16* out;   // [B*4096][D] bf16, natural token order, heads concatenated
  long long* trace;     // optional (debug): clock64() of pipeline events of CTA 0, [item][16] (inksam_op_attention_trace)
};

constexpr float kLog2e = 1.4426950408889634f;

template <int HD>
struct AttnTile {
  static constexpr bool H80 = (HD == 80);
  static_assert(HD == 64 || HD == 80, "head_dim must be 64 or 80");
  __host__ __device__ static constexpr int bytes(int rows) { return rows * 128 + (H80 ? ((rows * 32 + 1023) / 1024) * 1024 : 0); }
};

// D[tmem, 128 x N] = A[128 x HD] * B[N x HD]^T, both K-major tiles in smem (64-col SW128 part + 16-col SW32 part).
template <int HD>
__device__ __forceinline__ void issue_qk(uint32_t d_tmem, uint32_t a64, uint32_t a16, uint32_t b64, uint32_t b16,
                                         uint32_t idesc) {
  const uint64_t ad = make_kmajor_sw128_desc(a64), bd = make_kmajor_sw128_desc(b64);
#pragma unroll
  for (int k = 0; k < 4; ++k) umma<1, 0>(d_tmem, ad + 2u * k, bd + 2u * k, idesc, (uint32_t)(k != 0));
  if constexpr (HD == 80) umma<1, 0>(d_tmem, make_sw32_desc(a16), make_sw32_desc(b16), idesc, 1u);
}

// O[tmem, 128 x HD] (+)= P[tmem, 128 x 16*KSTEPS] * V[16*KSTEPS x HD]; V tiles are [keys][64] SW128 (+ [keys][16] SW32).
template <int HD, int KSTEPS>
__device__ __forceinline__ void issue_pv(uint32_t o_tmem, uint32_t p_tmem, uint32_t v64, uint32_t v16, bool accumulate) {
  constexpr uint32_t ID64 = make_idesc_bmn(1u, 128, 64);
  constexpr uint32_t ID16 = make_idesc_bmn(1u, 128, 16);
  const uint64_t d64 = make_kmajor_sw128_desc(v64);
  const uint64_t d16 = make_sw32_desc(v16);
#pragma unroll
  for (int ks = 0; ks < KSTEPS; ++ks) {
    const uint32_t acc = (uint32_t)(accumulate || ks != 0);
    umma_ts(o_tmem, p_tmem + 8u * ks, d64 + (uint64_t)(ks * (16 * 128 / 16)), ID64, acc);
    if constexpr (HD == 80) umma_ts(o_tmem + 64, p_tmem + 8u * ks, d16 + (uint64_t)(ks * (16 * 32 / 16)), ID16, acc);
  }
}

// O row (HD fp32 columns of this thread's TMEM lane) * inv -> bf16 -> 32-byte vector stores.
template <int HD>
__device__ __forceinline__ void store_o_row(uint32_t o_taddr, float inv, __nv_bfloat16* dst, bool valid) {
  uint32_t v[32];
  float* f = reinterpret_cast<float*>(v);
#pragma unroll
  for (int c = 0; c < 2; ++c) {
    tmem_ld_32x32(o_taddr + c * 32, v);
    tmem_ld_wait();
    if (valid) {
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        uint32_t pk[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) pk[j] = pack_bf16x2(f[h * 16 + 2 * j] * inv, f[h * 16 + 2 * j + 1] * inv);
        st_global_v8(dst + c * 32 + h * 16, pk);
      }
    }
  }
  if constexpr (HD == 80) {
    uint32_t w[16];
    float* g = reinterpret_cast<float*>(w);
    tmem_ld_32x16(o_taddr + 64, w);
    tmem_ld_wait();
    if (valid) {
      uint32_t pk[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) pk[j] = pack_bf16x2(g[2 * j] * inv, g[2 * j + 1] * inv);
      st_global_v8(dst + 64, pk);
    }
  }
}

// Same, with every tcgen05.ld issued before a single wait (the row's HD fp32 values sit in registers at once).
template <int HD>
__device__ __forceinline__ void store_o_row_once(uint32_t o_taddr, float inv, __nv_bfloat16* dst, bool valid) {
  uint32_t v0[32], v1[32], w[16];
  tmem_ld_32x32(o_taddr, v0);
  tmem_ld_32x32(o_taddr + 32, v1);
  if constexpr (HD == 80) tmem_ld_32x16(o_taddr + 64, w);
  tmem_ld_wait();
  if (!valid) return;
#pragma unroll
  for (int c = 0; c < 2; ++c) {
    const uint32_t (&v)[32] = c ? v1 : v0;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      uint32_t pk[8];
#pragma unroll
      for (int j = 0; j < 8; ++j)
        pk[j] = pack_bf16x2(__uint_as_float(v[h * 16 + 2 * j]) * inv, __uint_as_float(v[h * 16 + 2 * j + 1]) * inv);
      st_global_v8(dst + c * 32 + h * 16, pk);
    }
  }
  if constexpr (HD == 80) {
    uint32_t pk[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) pk[j] = pack_bf16x2(__uint_as_float(w[2 * j]) * inv, __uint_as_float(w[2 * j + 1]) * inv);
    st_global_v8(dst + 64, pk);
  }
}

// ------------------------------------------------------------------------------------------------ global
// Tensor maps over the qkv matrix [B*4096][3D] (natural token order): tm64 box {64 cols, 128 rows} SW128,
// tm16 box {16, 128} SW32; over the rel table [256][HD] (rows 0..126 rel_pos_h, 128..254 rel_pos_w):
// tr64 box {64, 128}, tr16 box {16, 128}.
template <int HD>
__global__ void __launch_bounds__(128, 2)
attn_global_kernel(const __grid_constant__ CUtensorMap tm64, const __grid_constant__ CUtensorMap tm16,
                   const __grid_constant__ CUtensorMap tr64, const __grid_constant__ CUtensorMap tr16, const AttnArgs a) {
  constexpr int TILE = AttnTile<HD>::bytes(128);  // 20480 (hd 80) or 16384 (hd 64)
  constexpr int BH_PITCH = 66;                    // floats per biasH row
  constexpr int ST_PITCH = 129;                   // floats per staging row (T_w, 128 columns)
  constexpr uint32_t IDESC_S = make_idesc(1u, 128, 128);
  static_assert(128 * ST_PITCH * 4 <= 2 * TILE + 128 * BH_PITCH * 4, "staging must fit over K + V + biasH");

  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t q64 = smem_base, q16 = q64 + 128 * 128;
  const uint32_t k64 = smem_base + TILE, k16 = k64 + 128 * 128;
  const uint32_t v64 = smem_base + 2 * TILE, v16 = v64 + 128 * 128;
  float* staging = reinterpret_cast<float*>(smem_gen + TILE);                 // over K, V, biasH (prologue only)
  float* biasH = reinterpret_cast<float*>(smem_gen + 3 * TILE);               // [128][66]
  const uint32_t bar_base = smem_base + 3 * TILE + 128 * BH_PITCH * 4;
  const uint32_t bar_q = bar_base, bar_k = bar_base + 8, bar_v = bar_base + 16, bar_s = bar_base + 24,
                 bar_o = bar_base + 32, tmem_slot = bar_base + 40;
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_gen + (tmem_slot - smem_base));

  const int tid = threadIdx.x, warp = tid >> 5;
  const int qt = blockIdx.x & 31;
  const int head = (blockIdx.x >> 5) % a.heads;
  const int b = (blockIdx.x >> 5) / a.heads;
  const int row_q = b * 4096 + qt * 128;
  const int col_q = head * HD, col_k = a.D + col_q, col_v = 2 * a.D + col_q;
  constexpr uint32_t TILE_TX = 128 * HD * 2;

  if (tid == 0) {
    tma_prefetch_desc(&tm64);
    tma_prefetch_desc(&tr64);
    if constexpr (HD == 80) { tma_prefetch_desc(&tm16); tma_prefetch_desc(&tr16); }
    mbar_init(bar_q, 1); mbar_init(bar_k, 1); mbar_init(bar_v, 1); mbar_init(bar_s, 1); mbar_init(bar_o, 1);
    fence_mbar_init();
  }
  if (warp == 0) {
    __syncwarp();
    tmem_alloc<1>(tmem_slot, 256);
    tmem_relinquish<1>();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;
  const uint32_t lane_base = tmem_base + ((uint32_t)(warp * 32) << 16);
  constexpr uint32_t S_COL = 0, O_COL = 128;

  // ---- prologue: Q tile, rel-pos tables T_h = Q rel_h^T (cols 0..127), T_w = Q rel_w^T (cols 128..255)
  if (tid == 0) {
    mbar_arrive_expect_tx(bar_q, TILE_TX);
    tma_load_2d(q64, &tm64, bar_q, col_q, row_q);
    if constexpr (HD == 80) tma_load_2d(q16, &tm16, bar_q, col_q + 64, row_q);
    mbar_arrive_expect_tx(bar_k, TILE_TX);
    tma_load_2d(k64, &tr64, bar_k, 0, 0);
    if constexpr (HD == 80) tma_load_2d(k16, &tr16, bar_k, 64, 0);
    mbar_arrive_expect_tx(bar_v, TILE_TX);
    tma_load_2d(v64, &tr64, bar_v, 0, 128);
    if constexpr (HD == 80) tma_load_2d(v16, &tr16, bar_v, 64, 128);
    mbar_wait(bar_q, 0); mbar_wait(bar_k, 0); mbar_wait(bar_v, 0);
    tc_fence_after();
    issue_qk<HD>(tmem_base + 0, q64, q16, k64, k16, IDESC_S);
    issue_qk<HD>(tmem_base + 128, q64, q16, v64, v16, IDESC_S);
    umma_commit<1>(bar_s);
  }
  __syncwarp();
  mbar_wait(bar_s, 0);
  tc_fence_after();

  const int qw = tid & 63;
  const int qh = 2 * qt + (tid >> 6);  // warp-uniform
  float bW[64];                        // bias_w[kw] * log2e for this query, kw = 0..63
  {
    uint32_t v[32];
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      tmem_ld_32x32(lane_base + 128 + c * 32, v);
      tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 32; ++i) staging[tid * ST_PITCH + c * 32 + i] = __uint_as_float(v[i]);
    }
#pragma unroll
    for (int i = 0; i < 64; ++i) bW[i] = staging[tid * ST_PITCH + qw + 63 - i] * kLog2e;  // T_w[q][qw - kw + 63]
    // T_h[q][qh - kh + 63]: qh is uniform over the warp, so the 64 needed columns start at column qh for all lanes
    uint32_t h0[32], h1[32];
    tmem_ld_32x32(lane_base + (uint32_t)qh, h0);
    tmem_ld_32x32(lane_base + (uint32_t)qh + 32, h1);
    tmem_ld_wait();
    __syncthreads();  // every thread has read its staging row (staging overlaps biasH)
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      biasH[tid * BH_PITCH + 63 - i] = __uint_as_float(h0[i]) * kLog2e;
      biasH[tid * BH_PITCH + 31 - i] = __uint_as_float(h1[i]) * kLog2e;
    }
  }
  tc_fence_before();
  fence_proxy_async_smem();  // generic writes over the K / V buffers before TMA refills them
  __syncthreads();

  if (tid == 0) {
    tc_fence_after();
    mbar_arrive_expect_tx(bar_k, TILE_TX);
    tma_load_2d(k64, &tm64, bar_k, col_k, b * 4096);
    if constexpr (HD == 80) tma_load_2d(k16, &tm16, bar_k, col_k + 64, b * 4096);
    mbar_arrive_expect_tx(bar_v, TILE_TX);
    tma_load_2d(v64, &tm64, bar_v, col_v, b * 4096);
    if constexpr (HD == 80) tma_load_2d(v16, &tm16, bar_v, col_v + 64, b * 4096);
  }

  float m_used = 0.f, l = 0.f;
  const float c1 = a.c1;
#pragma unroll 1
  for (int j = 0; j < 32; ++j) {
    const uint32_t par = (uint32_t)((j + 1) & 1);
    if (tid == 0) {
      mbar_wait(bar_k, par);
      tc_fence_after();
      issue_qk<HD>(tmem_base + S_COL, q64, q16, k64, k16, IDESC_S);
      umma_commit<1>(bar_s);
    }
    __syncwarp();
    mbar_wait(bar_s, par);
    tc_fence_after();
    if (tid == 0 && j + 1 < 32) {  // K buffer is free: prefetch the next key tile under the softmax
      mbar_arrive_expect_tx(bar_k, TILE_TX);
      tma_load_2d(k64, &tm64, bar_k, col_k, b * 4096 + (j + 1) * 128);
      if constexpr (HD == 80) tma_load_2d(k16, &tm16, bar_k, col_k + 64, b * 4096 + (j + 1) * 128);
    }
    __syncwarp();

    const float2 bh = *reinterpret_cast<const float2*>(&biasH[tid * BH_PITCH + 2 * j]);  // key rows 2j, 2j+1
    if (j == 0) {  // shift = max over the first 32 keys
      uint32_t v[32];
      tmem_ld_32x32(lane_base + S_COL, v);
      tmem_ld_wait();
      float mx = -INFINITY;
#pragma unroll
      for (int i = 0; i < 32; ++i) mx = fmaxf(mx, fmaf(__uint_as_float(v[i]), c1, bW[i] + bh.x));
      m_used = mx;
    }
    const float bhm0 = bh.x - m_used, bhm1 = bh.y - m_used;
    float tmax = -INFINITY;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      uint32_t v[32];
      tmem_ld_32x32(lane_base + S_COL + c * 32, v);
      tmem_ld_wait();
      uint32_t pk[16];
#pragma unroll
      for (int i = 0; i < 32; i += 2) {
        const float x0 = fmaf(__uint_as_float(v[i]), c1, bW[(c & 1) * 32 + i] + (c < 2 ? bhm0 : bhm1));
        const float x1 = fmaf(__uint_as_float(v[i + 1]), c1, bW[(c & 1) * 32 + i + 1] + (c < 2 ? bhm0 : bhm1));
        tmax = fmaxf(tmax, fmaxf(x0, x1));
        const float p0 = ex2_approx(fminf(x0, 120.f)), p1 = ex2_approx(fminf(x1, 120.f));
        l += p0 + p1;
        pk[i >> 1] = pack_bf16x2(p0, p1);
      }
      tmem_st_32x16(lane_base + S_COL + c * 16, pk);  // P over S: columns 16c..16c+15 are already consumed
    }
    tmem_st_wait();
    tc_fence_before();
    __syncthreads();
    if (tid == 0) {
      tc_fence_after();
      mbar_wait(bar_v, par);
      tc_fence_after();
      issue_pv<HD, 8>(tmem_base + O_COL, tmem_base + S_COL, v64, v16, j != 0);
      umma_commit<1>(bar_o);
      mbar_wait(bar_o, (uint32_t)(j & 1));  // P / S columns and the V buffer are free again
      if (j + 1 < 32) {
        mbar_arrive_expect_tx(bar_v, TILE_TX);
        tma_load_2d(v64, &tm64, bar_v, col_v, b * 4096 + (j + 1) * 128);
        if constexpr (HD == 80) tma_load_2d(v16, &tm16, bar_v, col_v + 64, b * 4096 + (j + 1) * 128);
      }
    }
    __syncwarp();
    const bool need = tmax > 32.f;
    if (__any_sync(0xffffffffu, need)) {  // exact but cold: move the shift up and rescale the accumulator
      mbar_wait(bar_o, (uint32_t)(j & 1));
      tc_fence_after();
      const float alpha = need ? ex2_approx(-tmax) : 1.f;
      if (need) m_used += tmax;
      l *= alpha;
      uint32_t v[16];
#pragma unroll
      for (int c = 0; c < HD / 16; ++c) {
        tmem_ld_32x16(lane_base + O_COL + c * 16, v);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = __float_as_uint(__uint_as_float(v[i]) * alpha);
        tmem_st_32x16(lane_base + O_COL + c * 16, v);
      }
      tmem_st_wait();
      tc_fence_before();
    }
  }

  mbar_wait(bar_o, 1u);  // PV of key tile 31
  tc_fence_after();
  store_o_row<HD>(lane_base + O_COL, 1.f / l, a.out + (size_t)(row_q + tid) * a.D + col_q, true);
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc<1>(tmem_base, 256);
}

// ------------------------------------------------------------------------------------------------ windowed
// The qkv matrix of a windowed block is stored in window order: [B*25 windows][196 tokens][3D], rows of padded
// tokens hold the qkv bias (the reference zero-pads AFTER norm1, so a padded token's q,k,v are the bias and it
// takes part in the softmax of the 9 border windows, image_encoder.py:170-174,258-259).
// Tensor maps over that matrix: tq64 box {64,128} / tq16 {16,128} for Q; tk64 {64,208} / tk16 {16,208} for K, V;
// over the rel table [64][HD] (rows 0..26 rel_pos_h, 32..58 rel_pos_w): tr64 {64,64}, tr16 {16,64}.
template <int HD>
__global__ void __launch_bounds__(128, 2)
attn_window_kernel(const __grid_constant__ CUtensorMap tq64, const __grid_constant__ CUtensorMap tq16,
                   const __grid_constant__ CUtensorMap tk64, const __grid_constant__ CUtensorMap tk16,
                   const __grid_constant__ CUtensorMap tr64, const __grid_constant__ CUtensorMap tr16, const AttnArgs a) {
  constexpr int QT = AttnTile<HD>::bytes(128);
  constexpr int KT = AttnTile<HD>::bytes(208);  // 33792 (hd 80) or 26624 (hd 64)
  constexpr int ST_PITCH = 65;
  constexpr int ST_BYTES = 128 * ST_PITCH * 4;  // 33280
  constexpr int VBUF = KT > ST_BYTES ? KT : ((ST_BYTES + 1023) / 1024) * 1024;
  constexpr uint32_t IDESC_S = make_idesc(1u, 128, 208);
  constexpr uint32_t IDESC_T = make_idesc(1u, 128, 64);

  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t q64 = smem_base, q16 = q64 + 128 * 128;
  const uint32_t k64 = smem_base + QT, k16 = k64 + 208 * 128;
  const uint32_t v64 = smem_base + QT + KT, v16 = v64 + 208 * 128;
  const uint32_t r64 = v64, r16 = v64 + 64 * 128;  // rel table lives in the V buffer until T is computed
  float* staging = reinterpret_cast<float*>(smem_gen + QT + KT);
  const uint32_t bar_base = smem_base + QT + KT + VBUF;
  const uint32_t bar_q = bar_base, bar_k = bar_base + 8, bar_v = bar_base + 16, bar_s = bar_base + 24,
                 bar_o = bar_base + 32, tmem_slot = bar_base + 40;
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_gen + (tmem_slot - smem_base));

  const int tid = threadIdx.x, warp = tid >> 5;
  const int mt = blockIdx.x & 1;
  const int head = (blockIdx.x >> 1) % a.heads;
  const int wb = (blockIdx.x >> 1) / a.heads;  // image * 25 + window
  const int row_w = wb * 196;
  const int col_q = head * HD, col_k = a.D + col_q, col_v = 2 * a.D + col_q;

  if (tid == 0) {
    tma_prefetch_desc(&tq64); tma_prefetch_desc(&tk64); tma_prefetch_desc(&tr64);
    if constexpr (HD == 80) { tma_prefetch_desc(&tq16); tma_prefetch_desc(&tk16); tma_prefetch_desc(&tr16); }
    mbar_init(bar_q, 1); mbar_init(bar_k, 1); mbar_init(bar_v, 1); mbar_init(bar_s, 1); mbar_init(bar_o, 1);
    fence_mbar_init();
  }
  if (warp == 0) {
    __syncwarp();
    tmem_alloc<1>(tmem_slot, 256);
    tmem_relinquish<1>();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;
  const uint32_t lane_base = tmem_base + ((uint32_t)(warp * 32) << 16);
  constexpr uint32_t O_COL = 112;  // S columns 104..207 are dead once P (columns 0..103) is written

  if (tid == 0) {
    mbar_arrive_expect_tx(bar_q, 128 * HD * 2);
    tma_load_2d(q64, &tq64, bar_q, col_q, row_w + mt * 128);
    if constexpr (HD == 80) tma_load_2d(q16, &tq16, bar_q, col_q + 64, row_w + mt * 128);
    mbar_arrive_expect_tx(bar_v, 64 * HD * 2);
    tma_load_2d(r64, &tr64, bar_v, 0, 0);
    if constexpr (HD == 80) tma_load_2d(r16, &tr16, bar_v, 64, 0);
    mbar_arrive_expect_tx(bar_k, 208 * HD * 2);
    tma_load_2d(k64, &tk64, bar_k, col_k, row_w);
    if constexpr (HD == 80) tma_load_2d(k16, &tk16, bar_k, col_k + 64, row_w);
    mbar_wait(bar_q, 0); mbar_wait(bar_v, 0);
    tc_fence_after();
    issue_qk<HD>(tmem_base + 0, q64, q16, r64, r16, IDESC_T);
    umma_commit<1>(bar_s);
  }
  __syncwarp();
  mbar_wait(bar_s, 0);
  tc_fence_after();

  // this thread's query: row r of the window (row-major 14x14); rows >= 196 are not tokens
  const int r = mt * 128 + tid;
  const bool is_tok = r < 196;
  const int qh = is_tok ? r / 14 : 0, qw = is_tok ? r % 14 : 0;
  float bH[14], bW[14];
  {
    uint32_t v[32];
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      tmem_ld_32x32(lane_base + c * 32, v);
      tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 32; ++i) staging[tid * ST_PITCH + c * 32 + i] = __uint_as_float(v[i]);
    }
#pragma unroll
    for (int k = 0; k < 14; ++k) {
      bH[k] = staging[tid * ST_PITCH + qh - k + 13] * kLog2e;       // q . rel_pos_h[qh - kh + 13]
      bW[k] = staging[tid * ST_PITCH + 32 + qw - k + 13] * kLog2e;  // q . rel_pos_w[qw - kw + 13]
    }
  }
  tc_fence_before();
  fence_proxy_async_smem();
  __syncthreads();
  if (tid == 0) {
    tc_fence_after();
    mbar_arrive_expect_tx(bar_v, 208 * HD * 2);
    tma_load_2d(v64, &tk64, bar_v, col_v, row_w);
    if constexpr (HD == 80) tma_load_2d(v16, &tk16, bar_v, col_v + 64, row_w);
    mbar_wait(bar_k, 0);
    tc_fence_after();
    issue_qk<HD>(tmem_base + 0, q64, q16, k64, k16, IDESC_S);
    umma_commit<1>(bar_s);
  }
  __syncwarp();
  mbar_wait(bar_s, 1);
  tc_fence_after();

  const float c1 = a.c1;
  float l = 0.f;
  {
    uint32_t v[32];
    tmem_ld_32x32(lane_base, v);
    tmem_ld_wait();
    float mx = -INFINITY;
#pragma unroll
    for (int i = 0; i < 32; ++i) mx = fmaxf(mx, fmaf(__uint_as_float(v[i]), c1, bH[i / 14] + bW[i % 14]));
#pragma unroll
    for (int k = 0; k < 14; ++k) bH[k] -= mx;  // fold the shift into the row bias
#pragma unroll
    for (int c = 0; c < 6; ++c) {
      if (c > 0) {
        tmem_ld_32x32(lane_base + c * 32, v);
        tmem_ld_wait();
      }
      uint32_t pk[16];
#pragma unroll
      for (int i = 0; i < 32; i += 2) {
        const int j0 = c * 32 + i, j1 = j0 + 1;
        const float x0 = fmaf(__uint_as_float(v[i]), c1, bH[j0 / 14] + bW[j0 % 14]);
        const float x1 = fmaf(__uint_as_float(v[i + 1]), c1, bH[j1 / 14] + bW[j1 % 14]);
        const float p0 = ex2_approx(fminf(x0, 120.f)), p1 = ex2_approx(fminf(x1, 120.f));
        l += p0 + p1;
        pk[i >> 1] = pack_bf16x2(p0, p1);
      }
      tmem_st_32x16(lane_base + c * 16, pk);
    }
    uint32_t w[16];
    tmem_ld_32x16(lane_base + 192, w);  // keys 192..207: only 192..195 exist
    tmem_ld_wait();
    uint32_t pk[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) pk[i] = 0u;
#pragma unroll
    for (int i = 0; i < 4; i += 2) {
      const int j0 = 192 + i, j1 = j0 + 1;
      const float x0 = fmaf(__uint_as_float(w[i]), c1, bH[j0 / 14] + bW[j0 % 14]);
      const float x1 = fmaf(__uint_as_float(w[i + 1]), c1, bH[j1 / 14] + bW[j1 % 14]);
      const float p0 = ex2_approx(fminf(x0, 120.f)), p1 = ex2_approx(fminf(x1, 120.f));
      l += p0 + p1;
      pk[i >> 1] = pack_bf16x2(p0, p1);
    }
    tmem_st_32x8(lane_base + 96, pk);
  }
  tmem_st_wait();
  tc_fence_before();
  __syncthreads();
  if (tid == 0) {
    tc_fence_after();
    mbar_wait(bar_v, 1);
    tc_fence_after();
    issue_pv<HD, 13>(tmem_base + O_COL, tmem_base + 0, v64, v16, false);
    umma_commit<1>(bar_o);
  }
  __syncwarp();
  mbar_wait(bar_o, 0);
  tc_fence_after();

  const int win = wb % 25, b = wb / 25;
  const int y = (win / 5) * 14 + qh, x = (win % 5) * 14 + qw;
  const bool valid = is_tok && y < 64 && x < 64;  // window_unpartition drops the padded rows / columns (:287-288)
  const size_t tok = (size_t)b * 4096 + (size_t)(valid ? y * 64 + x : 0);
  store_o_row<HD>(lane_base + O_COL, 1.f / l, a.out + tok * a.D + col_q, valid);
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc<1>(tmem_base, 256);
}

}  // namespace inksam
