// Windowed (14x14) encoder attention, persistent and warp-specialised: replaces Attention.forward +
// add_decomposed_rel_pos on window_partition'ed tokens + window_unpartition for the 28 windowed blocks of ViT-H
// (SA/modeling/image_encoder.py:170-177,224-240,243-289,325-361).
//
// Work item = 128 query rows (mt = 0: rows 0..127, mt = 1: rows 128..195) of one (image, window, head); the two items
// of a (window, head) run back to back on one CTA and share one load of K and V.  The keys are the
// window's 196 tokens (one 208-row TMA box; rows 196..207 belong to the next window and get P = 0).  The qkv matrix is
// in window order with the qkv bias in the rows of padded tokens (the reference zero-pads AFTER norm1, so padded tokens
// carry q = k = v = bias and are NOT masked, image_encoder.py:258-259).
//
// Persistent CTAs (2 per SM, 256 TMEM columns each), 160 threads:
//   warps 0-3  softmax: thread = query row = TMEM lane.  Warps whose 32 rows are all beyond row 195 only keep the
//              barrier protocol alive.
//   warp  4    one elected thread issues TMA and tcgen05.mma.  Q / K / V live in single smem buffers that are refilled
//              as soon as the MMA that last reads them has retired (Q after every S = Q K^T; K after the S and V after
//              the O = P V of the pair's second tile), so the loads of the next item fly under this item's softmax.  The rel-pos table stays resident.
// TMEM columns: T = Q REL^T [0,64) -> S [0,208) -> P (bf16) over [0,104) -> O [112,192).
#pragma once
#include "attn_tc.cuh"

namespace inksam {

template <int HD>
struct AttnWindow2 {
  static constexpr int QT = AttnTile<HD>::bytes(128);   // 20480 / 16384
  static constexpr int KT = AttnTile<HD>::bytes(208);   // 33792 / 26624
  static constexpr int RT = 64 * 128 + (HD == 80 ? 64 * 32 : 0);  // rel table, 64 rows
  static constexpr int OFF_Q = 0, OFF_K = QT, OFF_V = QT + KT, OFF_R = QT + 2 * KT;
  static constexpr int OFF_BAR = ((OFF_R + RT + 1023) / 1024) * 1024;
  static constexpr int SMEM_BYTES = OFF_BAR + 128 + 1024;
  static constexpr int THREADS = 160;
};

template <int HD>
__global__ void __launch_bounds__(160, 2)
attn_window2_kernel(const __grid_constant__ CUtensorMap tq64, const __grid_constant__ CUtensorMap tq16,
                    const __grid_constant__ CUtensorMap tk64, const __grid_constant__ CUtensorMap tk16,
                    const __grid_constant__ CUtensorMap tr64, const __grid_constant__ CUtensorMap tr16, const AttnArgs a,
                    const int num_items) {
  using L = AttnWindow2<HD>;
  constexpr uint32_t IDESC_S = make_idesc(1u, 128, 208);
  constexpr uint32_t IDESC_T = make_idesc(1u, 128, 64);
  constexpr uint32_t O_COL = 112;  // S columns 104..207 are dead once P (columns 0..103) is written

  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t q64 = smem_base + L::OFF_Q, q16 = q64 + 128 * 128;
  const uint32_t k64 = smem_base + L::OFF_K, k16 = k64 + 208 * 128;
  const uint32_t v64 = smem_base + L::OFF_V, v16 = v64 + 208 * 128;
  const uint32_t r64 = smem_base + L::OFF_R, r16 = r64 + 64 * 128;
  const uint32_t bar0 = smem_base + L::OFF_BAR;
  const uint32_t bar_q = bar0, bar_k = bar0 + 8, bar_v = bar0 + 16, bar_rel = bar0 + 24, bar_t = bar0 + 32,
                 bar_tread = bar0 + 40, bar_s = bar0 + 48, bar_p = bar0 + 56, bar_o = bar0 + 64, bar_ofree = bar0 + 72,
                 tmem_slot = bar0 + 80;
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_gen + L::OFF_BAR + 80);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) {
    mbar_init(bar_q, 1); mbar_init(bar_k, 1); mbar_init(bar_v, 1); mbar_init(bar_rel, 1);
    mbar_init(bar_t, 1); mbar_init(bar_s, 1); mbar_init(bar_o, 1);
    mbar_init(bar_tread, 4); mbar_init(bar_p, 4); mbar_init(bar_ofree, 4);
    fence_mbar_init();
  }
  if (warp == 4) {
    if (lane == 0) {
      tma_prefetch_desc(&tq64); tma_prefetch_desc(&tk64); tma_prefetch_desc(&tr64);
      if constexpr (HD == 80) { tma_prefetch_desc(&tq16); tma_prefetch_desc(&tk16); tma_prefetch_desc(&tr16); }
    }
    __syncwarp();
    tmem_alloc<1>(tmem_slot, 256);
    tmem_relinquish<1>();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;
  pdl_launch_dependents();  // the proj GEMM (launched with PDL) may run its prologue under this kernel's tail

  // The CTA walks (window, head) pairs blockIdx.x, blockIdx.x + gridDim.x, ...; each pair is two consecutive items
  // (row tiles mt = 0, 1) that share ONE load of the window's K and V.  num_items = 2 * pairs.
  const int num_pairs = num_items >> 1;
  auto decode = [&](int n, int& wb, int& head, int& mt) -> bool {  // n-th item of this CTA
    const int pair = blockIdx.x + (n >> 1) * gridDim.x;
    mt = n & 1;
    head = pair % a.heads;
    wb = pair / a.heads;
    return pair < num_pairs;
  };

  if (warp == 4) {
    // ------------------------------------------------------------------------------------------ issuer (TMA + MMA)
    if (elect_one()) {
      auto load_q = [&](int n) {
        int wb, head, mt;
        decode(n, wb, head, mt);
        mbar_arrive_expect_tx(bar_q, 128 * HD * 2);
        tma_load_2d(q64, &tq64, bar_q, head * HD, wb * 196 + mt * 128);
        if constexpr (HD == 80) tma_load_2d(q16, &tq16, bar_q, head * HD + 64, wb * 196 + mt * 128);
      };
      auto load_kv = [&](int n, int which, uint32_t d64, uint32_t d16, uint32_t bar) {
        int wb, head, mt;
        decode(n, wb, head, mt);
        const int col = which * a.D + head * HD;
        mbar_arrive_expect_tx(bar, 208 * HD * 2);
        tma_load_2d(d64, &tk64, bar, col, wb * 196);
        if constexpr (HD == 80) tma_load_2d(d16, &tk16, bar, col + 64, wb * 196);
      };
      mbar_arrive_expect_tx(bar_rel, 64 * HD * 2);
      tma_load_2d(r64, &tr64, bar_rel, 0, 0);
      if constexpr (HD == 80) tma_load_2d(r16, &tr16, bar_rel, 64, 0);
      int wb, head, mt;
      if (decode(0, wb, head, mt)) {
        load_q(0);
        load_kv(0, 1, k64, k16, bar_k);
        load_kv(0, 2, v64, v16, bar_v);
      }
      mbar_wait(bar_rel, 0);
      bool refill_v = false;
      for (int n = 0; decode(n, wb, head, mt); ++n) {
        const uint32_t ph = n & 1, kph = (n >> 1) & 1;   // item phase (Q, T, S, P, O), pair phase (K, V)
        int wb2, head2, mt2;
        const bool has_next = decode(n + 1, wb2, head2, mt2);
        mbar_wait(bar_q, ph);
        tc_fence_after();
        issue_qk<HD>(tmem_base, q64, q16, r64, r16, IDESC_T);  // T = Q REL^T; runs after PV of the previous item
        umma_commit<1>(bar_t);
        if (refill_v) {  // first item of a pair: the previous pair's last PV has to retire before V is overwritten
          mbar_wait(bar_o, ph ^ 1u);
          load_kv(n, 2, v64, v16, bar_v);
        }
        if (mt == 0) mbar_wait(bar_k, kph);
        mbar_wait(bar_tread, ph);                 // T has been read out of columns 0..63
        if (n > 0) mbar_wait(bar_ofree, ph ^ 1u);  // O of the previous item has been read out of columns 112..191
        tc_fence_after();
        issue_qk<HD>(tmem_base, q64, q16, k64, k16, IDESC_S);
        umma_commit<1>(bar_s);
        mbar_wait(bar_s, ph);  // Q (and, after the pair's second tile, K) is free: refill under this item's softmax
        if (has_next) {
          load_q(n + 1);
          if (mt == 1) load_kv(n + 1, 1, k64, k16, bar_k);
        }
        if (mt == 0) mbar_wait(bar_v, kph);
        mbar_wait(bar_p, ph);
        tc_fence_after();
        issue_pv<HD, 13>(tmem_base + O_COL, tmem_base, v64, v16, false);
        umma_commit<1>(bar_o);
        refill_v = (mt == 1 && has_next);
      }
    }
    __syncwarp();
  } else {
    // ------------------------------------------------------------------------------------------ softmax warps
    const uint32_t lane_base = tmem_base + ((uint32_t)(warp * 32) << 16);
    const float c1 = a.c1;
    int wb, head, mt;
    for (int n = 0; decode(n, wb, head, mt); ++n) {
      const uint32_t ph = n & 1;
      const int r = mt * 128 + tid;  // row of the window (row-major 14x14); rows >= 196 are not tokens
      const bool is_tok = r < 196;
      const bool warp_live = mt * 128 + warp * 32 < 196;
      const int qh = is_tok ? r / 14 : 0, qw = is_tok ? r % 14 : 0;
      float bH[14], bW[14];

      mbar_wait(bar_t, ph);
      tc_fence_after();
      if (warp_live) {
        float tl[64];  // this query's T row; indexed per thread below -> local memory
        uint32_t v[32], v2[32];
        tmem_ld_32x32(lane_base, v);
        tmem_ld_32x32(lane_base + 32, v2);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) { tl[i] = __uint_as_float(v[i]); tl[32 + i] = __uint_as_float(v2[i]); }
        const float* th = tl + qh + 13;
        const float* tw = tl + 32 + qw + 13;
#pragma unroll
        for (int k = 0; k < 14; ++k) {
          bH[k] = th[-k] * kLog2e;  // q . rel_pos_h[qh - kh + 13]
          bW[k] = tw[-k] * kLog2e;  // q . rel_pos_w[qw - kw + 13]
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_tread);

      mbar_wait(bar_s, ph);
      tc_fence_after();
      float l = 0.f;
      if (warp_live) {
        // S is read in 32-column chunks; the tcgen05.ld of chunk c+1 is in flight while chunk c is exponentiated
        uint32_t va[32], vb[32], w[16];
        tmem_ld_32x32(lane_base, va);
        tmem_ld_wait();
        float mx = -INFINITY;
#pragma unroll
        for (int i = 0; i < 32; ++i) mx = fmaxf(mx, fmaf(__uint_as_float(va[i]), c1, bH[i / 14] + bW[i % 14]));
#pragma unroll
        for (int k = 0; k < 14; ++k) bH[k] -= mx;  // fold the shift into the row bias
#pragma unroll
        for (int c = 0; c < 6; ++c) {
          uint32_t (&cur)[32] = (c & 1) ? vb : va;
          uint32_t (&nxt)[32] = (c & 1) ? va : vb;
          if (c < 5) tmem_ld_32x32(lane_base + (c + 1) * 32, nxt);
          else tmem_ld_32x16(lane_base + 192, w);  // keys 192..207: only 192..195 exist
          uint32_t pk[16];
#pragma unroll
          for (int i = 0; i < 32; i += 2) {
            const int j0 = c * 32 + i, j1 = j0 + 1;
            const float x0 = fmaf(__uint_as_float(cur[i]), c1, bH[j0 / 14] + bW[j0 % 14]);
            const float x1 = fmaf(__uint_as_float(cur[i + 1]), c1, bH[j1 / 14] + bW[j1 % 14]);
            const float p0 = ex2_approx(fminf(x0, 120.f)), p1 = ex2_approx(fminf(x1, 120.f));
            l += p0 + p1;
            pk[i >> 1] = pack_bf16x2(p0, p1);
          }
          tmem_st_32x16(lane_base + c * 16, pk);
          tmem_ld_wait();
        }
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
        tmem_st_wait();
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_p);

      mbar_wait(bar_o, ph);
      tc_fence_after();
      if (warp_live) {
        const int win = wb % 25, b = wb / 25;
        const int y = (win / 5) * 14 + qh, x = (win % 5) * 14 + qw;
        const bool valid = is_tok && y < 64 && x < 64;  // window_unpartition drops the padded rows / columns (:287-288)
        const size_t tok = (size_t)b * 4096 + (size_t)(valid ? y * 64 + x : 0);
        store_o_row_once<HD>(lane_base + O_COL, 1.f / l, a.out + tok * a.D + head * HD, valid);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_ofree);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 4) tmem_dealloc<1>(tmem_base, 256);
}

}  // namespace inksam
