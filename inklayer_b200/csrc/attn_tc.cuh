// Shared pieces of the encoder attention kernels for sm_100a (attn_global2.cuh, attn_window2..5.cuh): the argument block,
// the operand-tile geometry for head_dim 64 / 80, the tcgen05.mma issue helpers for S = Q K^T (both operands K-major from
// smem: a 128B-swizzled [rows][64] tile and, for head_dim 80, a 32B-swizzled [rows][16] tile) and O = P V (P from TMEM,
// V MN-major from smem), and the O-row store.  Reference arithmetic: Attention.forward + add_decomposed_rel_pos +
// get_rel_pos (SA/modeling/image_encoder.py:224-240, 292-361):
//   S = Q K^T          D -> TMEM fp32
//   T = Q REL^T        same MMA against the rel_pos table; T[q][r] = q . rel_pos[r] (the UNSCALED q, :234)
//   s = S*scale + T_h[q][qh-kh+S-1] + T_w[q][qw-kw+S-1]   (:354-359), all in log2 units
//   P = exp2(s - m)    bf16, stored back into TMEM over S (tcgen05.st), consumed as the A operand (TS MMA)
//   O += P V
// (The first-generation kernels that lived here -- 128 threads, one thread issuing TMA and MMA, 2 CTAs per SM -- were
// superseded in round 1 and removed in round 2.)
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

}  // namespace inksam
