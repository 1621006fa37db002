// Windowed (14x14) encoder attention, fifth version = the role-specialised pipeline of attn_window4.cuh (TMA producer, two MMA
// issuers, bias warps that hand the decomposed rel-pos terms to the tensor cores as an MMA operand, epilogue warps) with the
// SCHEDULE of attn_window3.cuh: all eight softmax warps work on ONE item, two warps per TMEM lane quarter splitting the
// keys, while the tensor cores produce the next item's S in the other S buffer.  tools/window_trace.py on version 4 showed
// each of its two softmax groups idle for ~45 % of the time, waiting for the chain  P ready -> PV -> O read -> next S  of
// its own S buffer, and alone on the MUFU pipe for most of the rest; here the next S is (mostly) ready when an item ends
// and the two warps of every SM sub-partition exponentiate at the same time.
//
// Replaces Attention.forward + add_decomposed_rel_pos on window_partition'ed tokens + window_unpartition for the windowed
// blocks (SA/modeling/image_encoder.py:170-177,224-240,243-289,325-361).  Item n = 128 query rows (mt = n & 1: rows 0..127 /
// 128..195) of one (image, window, head) pair; S buffer n & 1.
//   warp 17      TMA producer            warp 16  MMA issuer 1: S(n) = Q K^T + [bias] E^T and T(n+1) = Q REL^T
//   warp 18      MMA issuer 2: O(n) = P V (P in two pieces, see below)
//   warps 8-11   bias warps: T(n) -> this query's 14 + 14 rel-pos terms -> bf16 high / low halves -> TMEM operand [480,512)
//   warps 0-7    softmax: thread = (query row, key half).  Half 0 owns keys 0..111 (S columns [0,112)), walks them upwards
//                and writes P (bf16) over columns [0,56); half 1 owns keys 112..207 (196..207 are not keys), walks them
//                DOWNWARDS and writes P over columns [160,208): each half overwrites only S columns it has already read
//                itself, so the halves never wait for each other except to exchange their row maxima (one 64-thread named
//                barrier per lane quarter).  TRUE row maximum, as in the reference softmax: exact for any logit range.
//   warps 12-15  epilogue warps: O(n) (columns [64,144) of the S buffer) / (sum of the two halves' row sums) -> bf16 -> token order
// TMEM (512 columns): S0 [0,208)  S1 [208,416)  T [416,480)  bias operand [480,512).
#pragma once
#include "attn_window4.cuh"

namespace inksam {

template <int HD>
struct AttnWindow5 {
  static constexpr int QT = AttnTile<HD>::bytes(128);   // 20480 / 16384
  static constexpr int KT = AttnTile<HD>::bytes(208);   // 33792 / 26624
  static constexpr int RT = 64 * 128 + (HD == 80 ? 64 * 32 : 0);  // rel table, 64 rows
  static constexpr int ET = 208 * 128;                   // selection matrix E: 208 key rows x 64 bf16, 128B swizzle
  static constexpr int OFF_Q = 0, OFF_K = 2 * QT, OFF_V = 2 * QT + 2 * KT, OFF_R = 2 * QT + 4 * KT;
  static constexpr int OFF_E = ((OFF_R + RT + 1023) / 1024) * 1024;
  static constexpr int OFF_L = OFF_E + ET;               // row sums [2 buffers][2 halves][128], then row maxima [2][2][128] floats
  static constexpr int OFF_BAR = OFF_L + 2 * 2 * 2 * 128 * 4;
  static constexpr int SMEM_BYTES = OFF_BAR + 256 + 1024;
  static constexpr int THREADS = 608;
};

template <int HD, int KPOLY>
__global__ void __launch_bounds__(608, 1)
attn_window5_kernel(const __grid_constant__ CUtensorMap tq64, const __grid_constant__ CUtensorMap tq16,
                    const __grid_constant__ CUtensorMap tk64, const __grid_constant__ CUtensorMap tk16,
                    const __grid_constant__ CUtensorMap tr64, const __grid_constant__ CUtensorMap tr16, const AttnArgs a,
                    const int num_pairs) {
  using L = AttnWindow5<HD>;
  constexpr uint32_t IDESC_S = make_idesc(1u, 128, 208);
  constexpr uint32_t IDESC_T = make_idesc(1u, 128, 64);
  constexpr uint32_t S_STRIDE = 208, T_COL = 416, B_COL = 480, O_COL = 64, P1_COL = 160;
  static_assert(KPOLY % 2 == 0 && KPOLY >= 0 && KPOLY <= 32, "KPOLY: keys per 32-key chunk on the FMA-pipe exp2");

  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  auto q64 = [&](int s) { return smem_base + L::OFF_Q + s * L::QT; };
  auto k64 = [&](int s) { return smem_base + L::OFF_K + s * L::KT; };
  auto v64 = [&](int s) { return smem_base + L::OFF_V + s * L::KT; };
  const uint32_t r64 = smem_base + L::OFF_R, r16 = r64 + 64 * 128;
  const uint32_t e64 = smem_base + L::OFF_E;
  float* lsum_sm = reinterpret_cast<float*>(smem_gen + L::OFF_L);
  const uint32_t bar0 = smem_base + L::OFF_BAR;
  auto bar_qfull = [&](int s) { return bar0 + 8 * s; };
  auto bar_qempty = [&](int s) { return bar0 + 16 + 8 * s; };
  auto bar_kfull = [&](int s) { return bar0 + 32 + 8 * s; };
  auto bar_vfull = [&](int s) { return bar0 + 48 + 8 * s; };
  auto bar_kvempty = [&](int s) { return bar0 + 64 + 8 * s; };
  const uint32_t bar_rel = bar0 + 80, bar_tfull = bar0 + 88, bar_tread = bar0 + 96, bar_bready = bar0 + 104;
  auto bar_sfull = [&](int s) { return bar0 + 112 + 8 * s; };
  auto bar_ofull = [&](int s) { return bar0 + 128 + 8 * s; };
  auto bar_oread = [&](int s) { return bar0 + 144 + 8 * s; };
  auto bar_pready = [&](int s) { return bar0 + 160 + 8 * s; };
  const uint32_t bar_bfree = bar0 + 176;
  auto bar_done = [&](int i) { return bar0 + 184 + 8 * i; };  // teardown: the issuers' last commit-armed arrivals have landed
  const uint32_t tmem_slot = bar0 + 200;
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_gen + L::OFF_BAR + 200);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) {
    for (int s = 0; s < 2; ++s) {
      mbar_init(bar_qfull(s), 1); mbar_init(bar_qempty(s), 1); mbar_init(bar_kfull(s), 1); mbar_init(bar_vfull(s), 1);
      mbar_init(bar_kvempty(s), 1); mbar_init(bar_sfull(s), 1); mbar_init(bar_ofull(s), 1);
      mbar_init(bar_oread(s), 4); mbar_init(bar_pready(s), 8);
    }
    mbar_init(bar_rel, 1); mbar_init(bar_tfull, 1); mbar_init(bar_tread, 4); mbar_init(bar_bready, 4); mbar_init(bar_bfree, 1);
    mbar_init(bar_done(0), 1); mbar_init(bar_done(1), 1);
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
  // The selection matrix E (bf16, the B operand of the bias MMA: 208 key rows x 64, K-major, 128-byte swizzle):
  // E[key][c] = 1 for c = kh, 14 + kh (high and low half of bias_h[kh]) and c = 28 + kw, 42 + kw (bias_w[kw]), key = 14 kh + kw.
  for (int i = tid; i < L::ET / 16; i += L::THREADS)
    reinterpret_cast<uint4*>(smem_gen + L::OFF_E)[i] = make_uint4(0u, 0u, 0u, 0u);
  __syncthreads();
  for (int i = tid; i < 196 * 4; i += L::THREADS) {
    const int key = i >> 2, which = i & 3;
    const int kh = key / 14, kw = key - 14 * kh;
    const int c = which == 0 ? kh : which == 1 ? 14 + kh : which == 2 ? 28 + kw : 42 + kw;
    const uint32_t off = (uint32_t)(key >> 3) * 1024u + (uint32_t)(key & 7) * 128u + ((((uint32_t)c >> 3) ^ (uint32_t)(key & 7)) << 4) +
                         (uint32_t)(c & 7) * 2u;
    *reinterpret_cast<uint16_t*>(smem_gen + L::OFF_E + off) = 0x3F80;  // bf16 1.0
  }
  fence_proxy_async_smem();  // generic writes of E -> read by tcgen05.mma (async proxy)
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
  // debug trace of CTA 0 (AttnArgs::trace): event e of item n <- clock64()
  auto trace = [&](int n, int e) {
    if (a.trace != nullptr && blockIdx.x == 0) a.trace[n * 16 + e] = clock64();
  };

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
        if constexpr (HD == 80) {
          // V as FIVE [208 keys][16 columns] tiles (32-byte rows, 32-byte swizzle), 6656 bytes apart: one MN-major operand
          // of N = 80 for a single tcgen05.mma per 16 keys.  (Issuing an MMA costs the tensor pipe ~46 cycles however small
          // it is -- tools/micro/mma_issue.cu -- so the 64 + 16 column split of the K tile would double the cost of P V.)
          mbar_arrive_expect_tx(bar_vfull(ks), 208 * HD * 2);
#pragma unroll
          for (int t = 0; t < 5; ++t)
            tma_load_2d(v64(ks) + t * (208 * 32), &tk16, bar_vfull(ks), 2 * a.D + head * HD + 16 * t, wb * 196);
        } else {
          load_kv(2, v64(ks), bar_vfull(ks));
        }
      }
    }
    __syncwarp();
  } else if (warp == 18) {
    // ------------------------------------------------------------------------------------------ MMA issuer 2: O(n) = P V
    // (Issuing a tcgen05.mma costs the issuing thread ~50 cycles of descriptor moves into uniform registers; at 40 MMAs
    // per item one thread was the bottleneck of the whole pipeline, tools/window_trace.py.  PV is 26 of the 40.)
    if (elect_one()) {
      for (int m = 0; m < N; ++m) {
        const int jm = m >> 1, g = m & 1, ksm = jm & 1;
        mbar_wait(bar_pready(g), (uint32_t)(jm & 1));
        if (g == 0) mbar_wait(bar_vfull(ksm), (uint32_t)((jm >> 1) & 1));
        tc_fence_after();
        const uint32_t sbase = tmem_base + S_STRIDE * (uint32_t)g;
        // keys 0..111 (7 k-steps of 16): P at columns [0,56), V rows 0..111; keys 112..207 (6 k-steps): P at [160,208)
        if constexpr (HD == 80) {
          // B operand: MN-major, 32-byte swizzle, five 16-column atoms along N (leading byte offset 6656), 8-key groups 256
          // bytes apart along K (stride byte offset); 16 keys per MMA = 512 bytes
          constexpr uint32_t ID80 = make_idesc_bmn(1u, 128, 80);
          uint64_t vd = make_sw32_desc(v64(ksm));
          vd = (vd & ~(uint64_t)0x3FFF0000ull) | ((uint64_t)((208 * 32) >> 4) << 16);
#pragma unroll
          for (int ks2 = 0; ks2 < 13; ++ks2) {
            const uint32_t p_tmem = ks2 < 7 ? sbase + 8u * ks2 : sbase + P1_COL + 8u * (ks2 - 7);
            umma_ts(sbase + O_COL, p_tmem, vd + (uint64_t)(ks2 * (16 * 32 / 16)), ID80, (uint32_t)(ks2 != 0));
          }
        } else {
          issue_pv<HD, 7>(sbase + O_COL, sbase, v64(ksm), v64(ksm) + 208 * 128, false);
          issue_pv<HD, 6>(sbase + O_COL, sbase + P1_COL, v64(ksm) + 112 * 128, v64(ksm) + 208 * 128 + 112 * 32, true);
        }
        umma_commit<1>(bar_ofull(g));
        if (g == 1) umma_commit<1>(bar_kvempty(ksm));  // V (and K: its S MMAs retired long ago) of the pair are free
        trace(m, 2);
      }
      // Teardown: nobody waits for this thread's last kv_empty arrival.  One more commit, waited for here: when it has
      // landed every earlier commit-armed arrival of this thread has landed too, so none can hit shared memory that
      // already belongs to the next CTA.
      umma_commit<1>(bar_done(1));
      mbar_wait(bar_done(1), 0);
    }
    __syncwarp();
  } else if (warp == 16) {
    // ------------------------------------------------------------------------------------------ MMA issuer 1: T and S
    if (elect_one()) {
      auto issue_t = [&](int n) {  // T(n) = Q(n) REL^T; the T columns are free once the bias warps have read T(n-1)
        const int g = n & 1;
        tc_fence_after();
        issue_qk<HD>(tmem_base + T_COL, q64(g), q64(g) + 128 * 128, r64, r16, IDESC_T);
        umma_commit<1>(bar_tfull);
        trace(n, 0);
      };
      auto ready_t = [&](int n) {
        return mbar_test(bar_qfull(n & 1), (uint32_t)((n >> 1) & 1)) && (n == 0 || mbar_test(bar_tread, (uint32_t)((n - 1) & 1)));
      };
      auto ready_s = [&](int n) {
        const int j = n >> 1, g = n & 1;
        return mbar_test(bar_qfull(g), (uint32_t)(j & 1)) && mbar_test(bar_kfull(j & 1), (uint32_t)((j >> 1) & 1)) &&
               (n < 2 || mbar_test(bar_oread(g), (uint32_t)((j - 1) & 1))) && mbar_test(bar_bready, (uint32_t)(n & 1));
      };
      auto issue_s = [&](int n) {
        const int j = n >> 1, g = n & 1, ks = j & 1;
        tc_fence_after();
        const uint32_t s_tmem = tmem_base + S_STRIDE * (uint32_t)g;
        issue_qk<HD>(s_tmem, q64(g), q64(g) + 128 * 128, k64(ks), k64(ks) + 208 * 128, IDESC_S);
        umma_commit<1>(bar_qempty(g));  // (T(n), the other reader of this Q slot, was issued earlier by this thread)
        // += [bias_h | bias_w] E^T: the bias terms (bf16 high + low halves, written into TMEM by the bias warps) times the
        // 0/1 selection matrix -> every logit gets bias_h[kh] + bias_w[kw] in fp32 on the tensor cores
        const uint64_t ed = make_kmajor_sw128_desc(e64);
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_ts(s_tmem, tmem_base + B_COL + 8u * k, ed + 2u * k, IDESC_S, 1u);
        umma_commit<1>(bar_sfull(g));
        umma_commit<1>(bar_bfree);
        trace(n, 1);
      };
      mbar_wait(bar_rel, 0);
      // Two streams with their own preconditions: S(s) (needs its bias operand and its S buffer back from the epilogue
      // warps) and T(t), which runs up to two items ahead so that the bias gather is off the critical path.
      int t = 0, sN = 0;
      while (sN < N) {
        if (sN < t && ready_s(sN)) { issue_s(sN); ++sN; continue; }
        if (t < N && t <= sN + 1 && ready_t(t)) { issue_t(t); ++t; continue; }
      }
      umma_commit<1>(bar_done(0));  // teardown: the last q_empty / b_free arrivals (see issuer 2)
      mbar_wait(bar_done(0), 0);
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
      if (warp == 12 && lane == 0) trace(n, 8);
      if (quarter_live(n, wq)) {
        const int r = g * 128 + row;
        const bool is_tok = r < 196;
        const int qh = is_tok ? r / 14 : 0, qw = is_tok ? r % 14 : 0;
        const int win = wb % 25, b = wb / 25;
        const int y = (win / 5) * 14 + qh, x = (win % 5) * 14 + qw;
        const bool valid = is_tok && y < 64 && x < 64;  // window_unpartition drops the padded rows / columns (:287-288)
        const size_t tok = (size_t)b * 4096 + (size_t)(valid ? y * 64 + x : 0);
        const float inv = 1.f / (lsum_sm[(g * 2 + 0) * 128 + row] + lsum_sm[(g * 2 + 1) * 128 + row]);
        // O is read into registers in one go and the S buffer handed back BEFORE the row is scaled and stored
        uint32_t o0[32], o1[32], o2[16];
        const uint32_t o_taddr = lane_base + S_STRIDE * (uint32_t)g + O_COL;
        tmem_ld_32x32(o_taddr, o0);
        tmem_ld_32x32(o_taddr + 32, o1);
        if constexpr (HD == 80) tmem_ld_32x16(o_taddr + 64, o2);
        tmem_ld_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_oread(g));
        if (warp == 12 && lane == 0) trace(n, 9);
        if (valid) {
          __nv_bfloat16* dst = a.out + tok * a.D + head * HD;
#pragma unroll
          for (int c = 0; c < 2; ++c) {
            const uint32_t (&o)[32] = c ? o1 : o0;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              uint32_t pk[8];
#pragma unroll
              for (int j = 0; j < 8; ++j)
                pk[j] = pack_bf16x2(__uint_as_float(o[h * 16 + 2 * j]) * inv, __uint_as_float(o[h * 16 + 2 * j + 1]) * inv);
              st_global_v8(dst + c * 32 + h * 16, pk);
            }
          }
          if constexpr (HD == 80) {
            uint32_t pk[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) pk[j] = pack_bf16x2(__uint_as_float(o2[2 * j]) * inv, __uint_as_float(o2[2 * j + 1]) * inv);
            st_global_v8(dst + 64, pk);
          }
        }
      } else {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_oread(g));
      }
    }
  } else if (warp >= 8) {
    // ------------------------------------------------------------------------------------------ bias warps: T -> bias operand
    const int wq = warp & 3;
    const int row = wq * 32 + lane;
    const uint32_t lane_base = tmem_base + ((uint32_t)(wq * 32) << 16);
    const float to_s = kLog2e / a.c1;  // bias (natural units) -> units of q . k, so that x = c1 * (q . k + bias')
    for (int n = 0; n < N; ++n) {
      const int g = n & 1;
      const bool live = quarter_live(n, wq);
      uint32_t pk[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) pk[i] = 0u;
      mbar_wait(bar_tfull, (uint32_t)(n & 1));
      tc_fence_after();
      if (warp == 8 && lane == 0) trace(n, 3);
      if (live) {
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
        // operand columns (two bf16 each): K index 0..13 bias_h high halves, 14..27 low halves, 28..41 bias_w high, 42..55 low
        float hi[28], lo[28];
#pragma unroll
        for (int k = 0; k < 14; ++k) {
          const float bh = th[-k] * to_s, bw = tw[-k] * to_s;
          hi[k] = __bfloat162float(__float2bfloat16_rn(bh));
          lo[k] = bh - hi[k];
          hi[14 + k] = __bfloat162float(__float2bfloat16_rn(bw));
          lo[14 + k] = bw - hi[14 + k];
        }
#pragma unroll
        for (int k = 0; k < 7; ++k) {
          pk[k] = pack_bf16x2(hi[2 * k], hi[2 * k + 1]);
          pk[7 + k] = pack_bf16x2(lo[2 * k], lo[2 * k + 1]);
          pk[14 + k] = pack_bf16x2(hi[14 + 2 * k], hi[15 + 2 * k]);
          pk[21 + k] = pack_bf16x2(lo[14 + 2 * k], lo[15 + 2 * k]);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_tread);  // T(n) is in registers: the issuer may overwrite the T columns
      if (warp == 8 && lane == 0) trace(n, 4);
      if (n > 0) {
        mbar_wait(bar_bfree, (uint32_t)((n - 1) & 1));  // the bias MMA of item n-1 has read the operand columns
        tc_fence_after();
      }
      tmem_st_32x32(lane_base + B_COL, pk);  // dead quarters store zeros (their rows are never read back)
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_bready);
      if (warp == 8 && lane == 0) trace(n, 5);
    }
  } else {
    // ------------------------------------------------------------------------------------------ softmax warps
    const int h = warp >> 2, wq = warp & 3;
    const int row = wq * 32 + lane;  // row of the 128-row tile == TMEM lane
    const uint32_t lane_base = tmem_base + ((uint32_t)(wq * 32) << 16);
    float* mx_sm = lsum_sm + 2 * 2 * 128;  // row maxima [2 buffers][2 halves][128]
    const float c1 = a.c1;
    const uint64_t c12 = f2_pack(c1, c1);
    for (int n = 0; n < N; ++n) {
      const int g = n & 1;
      const uint32_t sbase = lane_base + S_STRIDE * (uint32_t)g;
      const bool live = quarter_live(n, wq);
      mbar_wait(bar_sfull(g), (uint32_t)((n >> 1) & 1));
      tc_fence_after();
      if (h == 0 && wq == 0 && lane == 0) trace(n, 6);
      // chunk c of this half: 32 S columns (half 1's first chunk: columns 176..207, of which 192..195 are keys and 196..207
      // are not).  Half 0 walks columns 0, 32, 64, 96(16); half 1 walks 176, 144, 112.
      // (The half is a run-time value of the warp on purpose: compile-time specialisation of the two halves into
      // straight-line code measured 4 % SLOWER -- the exponentials then leave in one burst and jam the MUFU queue.)
      float m = -INFINITY;
      if (live) {
        // ---- pass 1: the maximum of this half's logits (S already holds q . k + bias / scale)
        uint32_t v[32];
        if (h == 0) {
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            if (c < 3) tmem_ld_32x32(sbase + c * 32, v);
            else tmem_ld_32x16(sbase + 96, *reinterpret_cast<uint32_t(*)[16]>(&v[0]));
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < (c < 3 ? 32 : 16); i += 2) m = fmaxf(m, fmaxf(__uint_as_float(v[i]), __uint_as_float(v[i + 1])));
          }
        } else {
#pragma unroll
          for (int c = 0; c < 3; ++c) {
            tmem_ld_32x32(sbase + 176 - c * 32, v);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; i += 2)
              if (c > 0 || i < 20) m = fmaxf(m, fmaxf(__uint_as_float(v[i]), __uint_as_float(v[i + 1])));  // columns 196..207: not keys
          }
        }
      }
      mx_sm[(g * 2 + h) * 128 + row] = m;
      named_bar_sync(1 + wq, 64);  // the two warps of this lane quarter
      float lsum = 0.f;
      if (live) {
        m = fmaxf(m, mx_sm[(g * 2 + (h ^ 1)) * 128 + row]);
        if (h == 0 && wq == 0 && lane == 0) trace(n, 7);
        // ---- pass 2: p = exp2(c1 * (s - max)); P (bf16) over S columns this half has already read
        const float nm = -m * c1;
        const uint64_t nm2 = f2_pack(nm, nm);
        uint64_t l2 = f2_pack(0.f, 0.f);
        uint32_t v[32];
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          if (h == 1 && c == 3) break;
          const int ncol = (h == 0 && c == 3) ? 16 : 32;
          const uint32_t scol = h == 0 ? (uint32_t)(c * 32) : (uint32_t)(176 - c * 32);
          if (ncol == 32) tmem_ld_32x32(sbase + scol, v);
          else tmem_ld_32x16(sbase + scol, *reinterpret_cast<uint32_t(*)[16]>(&v[0]));
          tmem_ld_wait();
          uint32_t pk[16];
#pragma unroll
          for (int i = 0; i < 32; i += 2) {
            if (i < ncol && !(h == 1 && c == 0 && i >= 20)) {
              float x0, x1;
              f2_unpack(f2_fma(f2_pack(__uint_as_float(v[i]), __uint_as_float(v[i + 1])), c12, nm2), x0, x1);
              float p0, p1;
              if (KPOLY > 0 && i >= 32 - KPOLY) {
                f2_unpack(f2_exp2_poly(f2_pack(fmaxf(x0, -126.f), fmaxf(x1, -126.f))), p0, p1);
              } else {
                p0 = ex2_approx(x0); p1 = ex2_approx(x1);
              }
              l2 = f2_add(l2, f2_pack(p0, p1));
              pk[i >> 1] = pack_bf16x2(p0, p1);
            } else {
              pk[i >> 1] = 0u;  // keys 196..207 of the box belong to the next window: P = 0
            }
          }
          // keys of S columns [scol, scol + ncol) -> P columns: half 0 at scol / 2, half 1 at 160 + (scol - 112) / 2
          const uint32_t pcol = h == 0 ? scol / 2 : P1_COL + (scol - 112) / 2;
          if (ncol == 32) tmem_st_32x16(sbase + pcol, pk);
          else tmem_st_32x8(sbase + pcol, *reinterpret_cast<uint32_t(*)[8]>(&pk[0]));
        }
        tmem_st_wait();
        float la, lb;
        f2_unpack(l2, la, lb);
        lsum = la + lb;
      }
      lsum_sm[(g * 2 + h) * 128 + row] = lsum;
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_pready(g));
      if (h == 0 && wq == 0 && lane == 0) trace(n, 10);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 17) tmem_dealloc<1>(tmem_base, 512);
}

}  // namespace inksam
