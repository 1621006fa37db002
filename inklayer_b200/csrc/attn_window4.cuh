// Windowed (14x14) encoder attention, fourth version: ONE persistent CTA per SM in which every phase of an item runs on
// its own warps, so that the phases of consecutive items overlap.  Replaces Attention.forward + add_decomposed_rel_pos on
// window_partition'ed tokens + window_unpartition for the windowed blocks (SA/modeling/image_encoder.py:170-177,224-240,
// 243-289,325-361).
//
// What the profiles of the earlier versions showed (profiles/r02_window_attention.md): the kernel is neither bandwidth-
// nor pipe-bound but LATENCY-bound -- in attn_window2.cuh each CTA ran the chain  T = Q REL^T -> gather the bias terms
// -> S = Q K^T -> softmax -> O = P V -> store O  serially with 4 softmax warps (two such CTAs per SM; a third of the warp
// time went to the bias gather and the O store, a quarter to waiting for the tensor cores); attn_window3.cuh put 8 warps
// on one item at a time and lost the overlap of two independent chains.  Here:
//
//   warp 17      TMA producer: rel-pos table once; per (image, window, head) pair Q(mt=0), K, Q(mt=1), V into 2-slot rings
//   warp 16      MMA issuer (one thread); per item n (128 query rows: mt = n & 1 -> rows 0..127 / 128..195 of the window):
//                T(n) = Q REL^T -> TMEM columns [416,480);  S(n) = Q K^T -> S buffer n & 1;  then PV(n-1)
//   warps 8-11   bias warps (one per TMEM lane quarter): read T(n), gather this query's 14 + 14 decomposed rel-pos terms
//                (q . rel_pos_h[qh - kh + 13], q . rel_pos_w[qw - kw + 13], on the UNSCALED q: image_encoder.py:234,354-355)
//                into shared memory, in log2 units
//   warps 0-3    softmax group A: even items, S buffer 0      warps 4-7  softmax group B: odd items, S buffer 1
//                thread = query row = TMEM lane.  Two passes over the 196 logits of the row: the TRUE row maximum first
//                (as the reference softmax; exact for any logit range, nothing is clamped), then exp2(x - max), P (bf16)
//                over the first 104 columns of the S buffer, the row sum to shared memory
//   warps 12-15  epilogue warps: O(n) = P V (columns 112..191 of the S buffer) / row sum -> bf16 -> natural token order
//                (window_unpartition: padded rows / columns are dropped, image_encoder.py:287-288)
// The chain of a softmax group is  S(n) -> softmax -> PV(n) -> O read -> S(n+2);  T, the bias gather and the store of O
// run beside it, and the other group fills the tensor-core round trips.
// The qkv matrix is in window order with the qkv bias in the rows of padded tokens (the reference zero-pads AFTER norm1:
// padded tokens carry q = k = v = bias and are NOT masked, image_encoder.py:258-259).  Keys = the window's 196 tokens in
// one 208-row TMA box (rows 196..207 belong to the next window and get P = 0).
// TMEM (512 columns, the whole SM): S0 [0,208)  S1 [208,416)  T [416,480).
#pragma once
#include "attn_window3.cuh"

namespace inksam {

template <int HD>
struct AttnWindow4 {
  static constexpr int QT = AttnTile<HD>::bytes(128);   // 20480 / 16384
  static constexpr int KT = AttnTile<HD>::bytes(208);   // 33792 / 26624
  static constexpr int RT = 64 * 128 + (HD == 80 ? 64 * 32 : 0);  // rel table, 64 rows
  static constexpr int OFF_Q = 0, OFF_K = 2 * QT, OFF_V = 2 * QT + 2 * KT, OFF_R = 2 * QT + 4 * KT;
  static constexpr int BIAS_PITCH = 29;                  // floats per row: bias_h[14], bias_w[14] (+1: conflict-free)
  static constexpr int OFF_BIAS = ((OFF_R + RT + 1023) / 1024) * 1024;   // [2][128][29] floats
  static constexpr int OFF_L = OFF_BIAS + 2 * 128 * BIAS_PITCH * 4;      // row sums [2][128] floats
  static constexpr int OFF_BAR = ((OFF_L + 2 * 128 * 4 + 127) / 128) * 128;
  static constexpr int SMEM_BYTES = OFF_BAR + 256 + 1024;
  static constexpr int THREADS = 576;
};

template <int HD, int KPOLY>
__global__ void __launch_bounds__(576, 1)
attn_window4_kernel(const __grid_constant__ CUtensorMap tq64, const __grid_constant__ CUtensorMap tq16,
                    const __grid_constant__ CUtensorMap tk64, const __grid_constant__ CUtensorMap tk16,
                    const __grid_constant__ CUtensorMap tr64, const __grid_constant__ CUtensorMap tr16, const AttnArgs a,
                    const int num_pairs) {
  using L = AttnWindow4<HD>;
  constexpr uint32_t IDESC_S = make_idesc(1u, 128, 208);
  constexpr uint32_t IDESC_T = make_idesc(1u, 128, 64);
  constexpr uint32_t S_STRIDE = 208, T_COL = 416, O_COL = 112;
  static_assert(KPOLY % 2 == 0 && KPOLY >= 0 && KPOLY <= 32, "KPOLY: keys per 32-key chunk on the FMA-pipe exp2");

  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  auto q64 = [&](int s) { return smem_base + L::OFF_Q + s * L::QT; };
  auto k64 = [&](int s) { return smem_base + L::OFF_K + s * L::KT; };
  auto v64 = [&](int s) { return smem_base + L::OFF_V + s * L::KT; };
  const uint32_t r64 = smem_base + L::OFF_R, r16 = r64 + 64 * 128;
  float* bias_sm = reinterpret_cast<float*>(smem_gen + L::OFF_BIAS);
  float* lsum_sm = reinterpret_cast<float*>(smem_gen + L::OFF_L);
  const uint32_t bar0 = smem_base + L::OFF_BAR;
  auto bar_qfull = [&](int s) { return bar0 + 8 * s; };
  auto bar_qempty = [&](int s) { return bar0 + 16 + 8 * s; };
  auto bar_kfull = [&](int s) { return bar0 + 32 + 8 * s; };
  auto bar_vfull = [&](int s) { return bar0 + 48 + 8 * s; };
  auto bar_kvempty = [&](int s) { return bar0 + 64 + 8 * s; };
  const uint32_t bar_rel = bar0 + 80, bar_tfull = bar0 + 88, bar_tread = bar0 + 96;
  auto bar_sfull = [&](int s) { return bar0 + 112 + 8 * s; };
  auto bar_ofull = [&](int s) { return bar0 + 128 + 8 * s; };
  auto bar_oread = [&](int s) { return bar0 + 144 + 8 * s; };
  auto bar_pready = [&](int s) { return bar0 + 160 + 8 * s; };
  auto bar_bready = [&](int s) { return bar0 + 176 + 8 * s; };
  const uint32_t tmem_slot = bar0 + 200;
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_gen + L::OFF_BAR + 200);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) {
    for (int s = 0; s < 2; ++s) {
      mbar_init(bar_qfull(s), 1); mbar_init(bar_qempty(s), 1); mbar_init(bar_kfull(s), 1); mbar_init(bar_vfull(s), 1);
      mbar_init(bar_kvempty(s), 1); mbar_init(bar_sfull(s), 1); mbar_init(bar_ofull(s), 1);
      mbar_init(bar_oread(s), 4); mbar_init(bar_pready(s), 4); mbar_init(bar_bready(s), 4);
    }
    mbar_init(bar_rel, 1); mbar_init(bar_tfull, 1); mbar_init(bar_tread, 4);
    fence_mbar_init();
  }
  if (warp == 17) {
    if (lane == 0) {
      tma_prefetch_desc(&tq64); tma_prefetch_desc(&tk64); tma_prefetch_desc(&tr64);
      if constexpr (HD == 80) { tma_prefetch_desc(&tq16); tma_prefetch_desc(&tk16); tma_prefetch_desc(&tr16); }
    }
    __syncwarp();
    tmem_alloc<1>(tmem_slot, 512);
    tmem_relinquish<1>();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;
  pdl_launch_dependents();  // the proj GEMM (launched with PDL) may run its prologue under this kernel's tail

  // this CTA's pairs: blockIdx.x + j * gridDim.x, j = 0 .. my_pairs - 1; item n = 2 j + mt
  const int my_pairs = ((int)blockIdx.x < num_pairs) ? (num_pairs - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
  const int N = 2 * my_pairs;
  auto pair_of = [&](int j, int& wb, int& head) {
    const int p = blockIdx.x + j * gridDim.x;
    head = p % a.heads;
    wb = p / a.heads;  // image * 25 + window
  };
  // rows of the window handled by TMEM lane quarter wq of item n (row-major 14x14; rows >= 196 are not tokens)
  auto quarter_live = [](int n, int wq) { return (n & 1) * 128 + wq * 32 < 196; };

  if (warp == 17) {
    // ------------------------------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      mbar_arrive_expect_tx(bar_rel, 64 * HD * 2);
      tma_load_2d(r64, &tr64, bar_rel, 0, 0);
      if constexpr (HD == 80) tma_load_2d(r16, &tr16, bar_rel, 64, 0);
      for (int j = 0; j < my_pairs; ++j) {
        int wb, head;
        pair_of(j, wb, head);
        const int ks = j & 1;
        auto load_q = [&](int mt) {
          if (j >= 1) mbar_wait(bar_qempty(mt), (uint32_t)((j - 1) & 1));  // S of item 2 (j - 1) + mt has read the slot
          mbar_arrive_expect_tx(bar_qfull(mt), 128 * HD * 2);
          tma_load_2d(q64(mt), &tq64, bar_qfull(mt), head * HD, wb * 196 + mt * 128);
          if constexpr (HD == 80) tma_load_2d(q64(mt) + 128 * 128, &tq16, bar_qfull(mt), head * HD + 64, wb * 196 + mt * 128);
        };
        auto load_kv = [&](int which, uint32_t dst, uint32_t bar) {
          const int col = which * a.D + head * HD;
          mbar_arrive_expect_tx(bar, 208 * HD * 2);
          tma_load_2d(dst, &tk64, bar, col, wb * 196);
          if constexpr (HD == 80) tma_load_2d(dst + 208 * 128, &tk16, bar, col + 64, wb * 196);
        };
        load_q(0);
        if (j >= 2) mbar_wait(bar_kvempty(ks), (uint32_t)(((j >> 1) - 1) & 1));  // PV of pair j - 2 has retired
        load_kv(1, k64(ks), bar_kfull(ks));
        load_q(1);
        load_kv(2, v64(ks), bar_vfull(ks));
      }
    }
    __syncwarp();
  } else if (warp == 16) {
    // ------------------------------------------------------------------------------------------ MMA issuer
    if (elect_one()) {
      auto issue_pv_item = [&](int m) {
        const int jm = m >> 1, g = m & 1, ksm = jm & 1;
        mbar_wait(bar_pready(g), (uint32_t)(jm & 1));
        if (g == 0) mbar_wait(bar_vfull(ksm), (uint32_t)((jm >> 1) & 1));
        tc_fence_after();
        const uint32_t sbase = tmem_base + S_STRIDE * (uint32_t)g;
        issue_pv<HD, 13>(sbase + O_COL, sbase, v64(ksm), v64(ksm) + 208 * 128, false);
        umma_commit<1>(bar_ofull(g));
        if (g == 1) umma_commit<1>(bar_kvempty(ksm));  // K and V of the pair are free once its last PV has retired
      };
      mbar_wait(bar_rel, 0);
      for (int n = 0; n < N; ++n) {
        const int j = n >> 1, g = n & 1, ks = j & 1;
        mbar_wait(bar_qfull(g), (uint32_t)(j & 1));
        if (n > 0) mbar_wait(bar_tread, (uint32_t)((n - 1) & 1));  // T(n-1) has been read out of the T columns
        tc_fence_after();
        issue_qk<HD>(tmem_base + T_COL, q64(g), q64(g) + 128 * 128, r64, r16, IDESC_T);
        umma_commit<1>(bar_tfull);
        if (g == 0) mbar_wait(bar_kfull(ks), (uint32_t)((j >> 1) & 1));
        if (n >= 2) mbar_wait(bar_oread(g), (uint32_t)((j - 1) & 1));  // O(n-2) has been read out of this S buffer
        tc_fence_after();
        issue_qk<HD>(tmem_base + S_STRIDE * (uint32_t)g, q64(g), q64(g) + 128 * 128, k64(ks), k64(ks) + 208 * 128, IDESC_S);
        umma_commit<1>(bar_sfull(g));
        umma_commit<1>(bar_qempty(g));
        if (n > 0) issue_pv_item(n - 1);
      }
      if (N > 0) issue_pv_item(N - 1);
    }
    __syncwarp();
  } else if (warp >= 12) {
    // ------------------------------------------------------------------------------------------ epilogue warps: O -> global
    const int wq = warp & 3;
    const int row = wq * 32 + lane;
    const uint32_t lane_base = tmem_base + ((uint32_t)(wq * 32) << 16);
    for (int n = 0; n < N; ++n) {
      int wb, head;
      pair_of(n >> 1, wb, head);
      const int g = n & 1;
      mbar_wait(bar_pready(g), (uint32_t)((n >> 1) & 1));  // the row sums of item n are in shared memory
      mbar_wait(bar_ofull(g), (uint32_t)((n >> 1) & 1));
      tc_fence_after();
      if (quarter_live(n, wq)) {
        const int r = g * 128 + row;
        const bool is_tok = r < 196;
        const int qh = is_tok ? r / 14 : 0, qw = is_tok ? r % 14 : 0;
        const int win = wb % 25, b = wb / 25;
        const int y = (win / 5) * 14 + qh, x = (win % 5) * 14 + qw;
        const bool valid = is_tok && y < 64 && x < 64;  // window_unpartition drops the padded rows / columns (:287-288)
        const size_t tok = (size_t)b * 4096 + (size_t)(valid ? y * 64 + x : 0);
        const float inv = 1.f / lsum_sm[g * 128 + row];
        store_o_row<HD>(lane_base + S_STRIDE * (uint32_t)g + O_COL, inv, a.out + tok * a.D + head * HD, valid);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_oread(g));
    }
  } else if (warp >= 8) {
    // ------------------------------------------------------------------------------------------ bias warps: T -> rel-pos terms
    const int wq = warp & 3;
    const int row = wq * 32 + lane;
    const uint32_t lane_base = tmem_base + ((uint32_t)(wq * 32) << 16);
    for (int n = 0; n < N; ++n) {
      const int g = n & 1;
      mbar_wait(bar_tfull, (uint32_t)(n & 1));
      tc_fence_after();
      if (quarter_live(n, wq)) {
        const int r = g * 128 + row;
        const bool is_tok = r < 196;
        const int qh = is_tok ? r / 14 : 0, qw = is_tok ? r % 14 : 0;
        float tl[64];  // this query's T row; indexed per thread below -> local memory
        uint32_t v[64];
        tmem_ld_32x32p(lane_base + T_COL, v);
        tmem_ld_32x32p(lane_base + T_COL + 32, v + 32);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 64; ++i) tl[i] = __uint_as_float(v[i]);
        const float* th = tl + qh + 13;       // q . rel_pos_h[qh - kh + 13]
        const float* tw = tl + 32 + qw + 13;  // q . rel_pos_w[qw - kw + 13]
        float* dst = bias_sm + (size_t)(g * 128 + row) * L::BIAS_PITCH;
#pragma unroll
        for (int k = 0; k < 14; ++k) {
          dst[k] = th[-k] * kLog2e;
          dst[14 + k] = tw[-k] * kLog2e;
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(bar_tread);
        mbar_arrive(bar_bready(g));
      }
    }
  } else {
    // ------------------------------------------------------------------------------------------ softmax groups
    const int g = warp >> 2, wq = warp & 3;
    const int row = wq * 32 + lane;  // row of the 128-row tile == TMEM lane
    const uint32_t sbase = tmem_base + ((uint32_t)(wq * 32) << 16) + S_STRIDE * (uint32_t)g;
    const uint64_t c12 = f2_pack(a.c1, a.c1);
    const bool live = quarter_live(g, wq);  // group g always works on items with mt == g
    for (int n = g; n < N; n += 2) {
      const uint32_t ph = (uint32_t)((n >> 1) & 1);
      uint64_t bH2[14], bW2[7];
      mbar_wait(bar_bready(g), ph);
      if (live) {
        const float* src = bias_sm + (size_t)(g * 128 + row) * L::BIAS_PITCH;
#pragma unroll
        for (int k = 0; k < 14; ++k) bH2[k] = f2_pack(src[k], src[k]);
#pragma unroll
        for (int k = 0; k < 7; ++k) bW2[k] = f2_pack(src[14 + 2 * k], src[15 + 2 * k]);
      }
      mbar_wait(bar_sfull(g), ph);
      tc_fence_after();
      float lsum = 1.f;
      if (live) {
        // x(j) = s(j) * scale * log2e + bias_h[j / 14] + bias_w[j % 14], two keys per instruction (j even: same key row)
#define INKSAM_WIN_X2(v, j, i) \
  f2_fma(f2_pack(__uint_as_float((v)[i]), __uint_as_float((v)[(i) + 1])), c12, f2_add(bH2[(j) / 14], bW2[((j) % 14) / 2]))
        // ---- pass 1: the row maximum over the 196 keys (keys 196..207 of the box belong to the next window)
        float m = -INFINITY;
        {
          uint32_t va[32], vb[32];
          tmem_ld_32x32(sbase, va);
#pragma unroll
          for (int c = 0; c < 7; ++c) {
            uint32_t (&cur)[32] = (c & 1) ? vb : va;
            uint32_t (&nxt)[32] = (c & 1) ? va : vb;
            tmem_ld_wait();
            if (c < 6) tmem_ld_32x32(sbase + (c + 1) * 32, nxt);  // chunk 6: columns 192..223, only 192..195 are keys
#pragma unroll
            for (int i = 0; i < (c < 6 ? 32 : 4); i += 2) {
              float x0, x1;
              f2_unpack(INKSAM_WIN_X2(cur, c * 32 + i, i), x0, x1);
              m = fmaxf(m, fmaxf(x0, x1));
            }
          }
        }
        // ---- pass 2: p = exp2(x - max), P (bf16) over S, row sum
        const uint64_t nm2 = f2_pack(-m, -m);
#pragma unroll
        for (int k = 0; k < 14; ++k) bH2[k] = f2_add(bH2[k], nm2);  // fold the shift into the row bias
        uint64_t l2 = f2_pack(0.f, 0.f);
        {
          uint32_t va[32], vb[32];
          tmem_ld_32x32(sbase, va);
#pragma unroll
          for (int c = 0; c < 7; ++c) {
            uint32_t (&cur)[32] = (c & 1) ? vb : va;
            uint32_t (&nxt)[32] = (c & 1) ? va : vb;
            tmem_ld_wait();
            if (c < 6) tmem_ld_32x32(sbase + (c + 1) * 32, nxt);
            uint32_t pk[16];
#pragma unroll
            for (int i = 0; i < 32; i += 2) {
              if (i < (c < 6 ? 32 : 4)) {
                float x0, x1, p0, p1;
                f2_unpack(INKSAM_WIN_X2(cur, c * 32 + i, i), x0, x1);
                if (i >= 32 - KPOLY && c < 6) {
                  f2_unpack(f2_exp2_poly(f2_pack(fmaxf(x0, -126.f), fmaxf(x1, -126.f))), p0, p1);
                } else {
                  p0 = ex2_approx(x0); p1 = ex2_approx(x1);
                }
                l2 = f2_add(l2, f2_pack(p0, p1));
                pk[i >> 1] = pack_bf16x2(p0, p1);
              } else {
                pk[i >> 1] = 0u;
              }
            }
            // P columns 16c .. 16c+15 <- keys 32c .. 32c+31: only S chunks <= c (all read) live there
            if (c < 6) tmem_st_32x16(sbase + c * 16, pk);
            else tmem_st_32x8(sbase + 96, *reinterpret_cast<uint32_t(*)[8]>(&pk[0]));  // keys 192..207: 4 real, 12 zero
          }
        }
#undef INKSAM_WIN_X2
        tmem_st_wait();
        float la, lb;
        f2_unpack(l2, la, lb);
        lsum = la + lb;
      }
      lsum_sm[g * 128 + row] = lsum;
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_pready(g));
      // S buffer g is rewritten by S(n+2) only after the epilogue warps have read O(n) (bar_oread): nothing to wait for here
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 17) tmem_dealloc<1>(tmem_base, 512);
}

}  // namespace inksam
