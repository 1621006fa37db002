// Windowed (14x14) encoder attention, third version: ONE persistent CTA per SM that keeps the tensor cores, the TMA engine
// and eight softmax warps busy at the same time.  Replaces Attention.forward + add_decomposed_rel_pos on
// window_partition'ed tokens + window_unpartition for the windowed blocks (SA/modeling/image_encoder.py:170-177,224-240,
// 243-289,325-361), like attn_window2.cuh, whose per-item chain  T -> read T -> S -> softmax -> PV -> store O  ran
// serially inside each of two CTAs per SM (ncu: 40 % of warp samples on the hand-off barriers, 120 us per layer at batch 8).
//
// Work item n = 128 query rows (mt = n & 1: rows 0..127 / 128..195) of one (image, window, head) pair; a CTA walks pairs
// blockIdx.x, blockIdx.x + gridDim.x, ... and both items of a pair share one load of the window's K and V.  The qkv
// matrix is in window order with the qkv bias in the rows of padded tokens (the reference zero-pads AFTER norm1: padded
// tokens carry q = k = v = bias and are NOT masked, image_encoder.py:258-259).  Keys = the window's 196 tokens in one
// 208-row TMA box (rows 196..207 belong to the next window and get P = 0).
//
//   warp 9      TMA producer: rel-pos table once; per pair Q(mt=0), K, Q(mt=1), V into 2-slot rings (a pair ahead)
//   warp 8      MMA issuer (one thread).  Per item n, in this order:  T(n) = Q REL^T -> its own TMEM columns;
//               S(n) = Q K^T -> S buffer n & 1;  then PV(n-1).  T and S of item n therefore execute while the softmax warps
//               are still busy with item n-1, and the MMA latency leaves the softmax warps' critical path.
//   warps 0-7   softmax, 2 warps per TMEM lane quarter: thread = (query row, key half).  Half h owns keys [98h, 98h+98)
//               = key rows 7h..7h+6 of the 14x14 window: it reads its 98 S columns into registers in one go, adds the
//               decomposed rel-pos bias, the two halves exchange their row maxima through shared memory (one named
//               barrier, which also orders every S read before any P write), then exp2(x - max), P (bf16) back into the S
//               buffer's first 104 columns, partial row sums to shared memory.  The TRUE row maximum is subtracted, as in
//               the reference softmax: exact for any logit range, nothing is clamped.  O(n-1) = P V (columns 112..191 of
//               the other S buffer) is normalised and stored in the middle of item n, between the max exchange and the
//               exponentials, i.e. after PV(n-1) has had a few hundred cycles to finish; S(n+1) needs that buffer.
// TMEM (512 columns, the whole SM): S0 [0,208)  S1 [208,416)  T [416,480).
#pragma once
#include "attn_tc.cuh"

namespace inksam {

template <int HD>
struct AttnWindow3 {
  static constexpr int QT = AttnTile<HD>::bytes(128);   // 20480 / 16384
  static constexpr int KT = AttnTile<HD>::bytes(208);   // 33792 / 26624
  static constexpr int RT = 64 * 128 + (HD == 80 ? 64 * 32 : 0);  // rel table, 64 rows
  static constexpr int OFF_Q = 0, OFF_K = 2 * QT, OFF_V = 2 * QT + 2 * KT, OFF_R = 2 * QT + 4 * KT;
  static constexpr int OFF_X = ((OFF_R + RT + 1023) / 1024) * 1024;  // row maxima and row sums: [2][2][2][128] floats
  static constexpr int OFF_BAR = OFF_X + 4096;
  static constexpr int SMEM_BYTES = OFF_BAR + 256 + 1024;
  static constexpr int THREADS = 320;
};

__device__ __forceinline__ void tmem_ld_32x32p(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_st_32x32p(uint32_t taddr, const uint32_t* r) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]),
      "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]),
      "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
}
__device__ __forceinline__ void tmem_st_32x16p(uint32_t taddr, const uint32_t* r) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}

// This thread's half of an O row (HD / 2 fp32 columns at o_taddr) * inv -> bf16 -> 16-byte stores.
template <int HD>
__device__ __forceinline__ void store_o_half(uint32_t o_taddr, float inv, __nv_bfloat16* dst, bool valid) {
  constexpr int NC = HD / 2;  // 40 or 32
  uint32_t v[40];
  tmem_ld_32x32p(o_taddr, v);
  if constexpr (HD == 80) {
    uint32_t w[8];
    tmem_ld_32x8(o_taddr + 32, w);
    tmem_ld_wait();
#pragma unroll
    for (int i = 0; i < 8; ++i) v[32 + i] = w[i];
  } else {
    tmem_ld_wait();
  }
  if (!valid) return;
#pragma unroll
  for (int c = 0; c < NC / 8; ++c) {
    uint32_t pk[4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
      pk[i] = pack_bf16x2(__uint_as_float(v[8 * c + 2 * i]) * inv, __uint_as_float(v[8 * c + 2 * i + 1]) * inv);
    st_global_v4(dst + 8 * c, pk[0], pk[1], pk[2], pk[3]);
  }
}

// KPOLY: how many of a thread's 98 keys take exp2 from the FMA-pipe cubic instead of the MUFU pipe (even, <= 98).
template <int HD, int KPOLY>
__global__ void __launch_bounds__(320, 1)
attn_window3_kernel(const __grid_constant__ CUtensorMap tq64, const __grid_constant__ CUtensorMap tq16,
                    const __grid_constant__ CUtensorMap tk64, const __grid_constant__ CUtensorMap tk16,
                    const __grid_constant__ CUtensorMap tr64, const __grid_constant__ CUtensorMap tr16, const AttnArgs a,
                    const int num_pairs) {
  using L = AttnWindow3<HD>;
  constexpr uint32_t IDESC_S = make_idesc(1u, 128, 208);
  constexpr uint32_t IDESC_T = make_idesc(1u, 128, 64);
  constexpr uint32_t S_STRIDE = 208, T_COL = 416, O_COL = 112;
  static_assert(KPOLY % 2 == 0 && KPOLY >= 0 && KPOLY <= 98, "KPOLY");

  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  auto q64 = [&](int s) { return smem_base + L::OFF_Q + s * L::QT; };
  auto k64 = [&](int s) { return smem_base + L::OFF_K + s * L::KT; };
  auto v64 = [&](int s) { return smem_base + L::OFF_V + s * L::KT; };
  const uint32_t r64 = smem_base + L::OFF_R, r16 = r64 + 64 * 128;
  float* xch = reinterpret_cast<float*>(smem_gen + L::OFF_X);  // [buf 2][kind 2: max, sum][half 2][128]
  auto xmax = [&](int buf, int half) { return xch + ((buf * 2 + 0) * 2 + half) * 128; };
  auto xsum = [&](int buf, int half) { return xch + ((buf * 2 + 1) * 2 + half) * 128; };
  const uint32_t bar0 = smem_base + L::OFF_BAR;
  auto bar_qfull = [&](int s) { return bar0 + 8 * s; };
  auto bar_qempty = [&](int s) { return bar0 + 16 + 8 * s; };
  auto bar_kfull = [&](int s) { return bar0 + 32 + 8 * s; };
  auto bar_vfull = [&](int s) { return bar0 + 48 + 8 * s; };
  auto bar_kvempty = [&](int s) { return bar0 + 64 + 8 * s; };
  const uint32_t bar_rel = bar0 + 80, bar_tfull = bar0 + 88, bar_tread = bar0 + 96, bar_pready = bar0 + 104;
  auto bar_sfull = [&](int s) { return bar0 + 112 + 8 * s; };
  auto bar_ofull = [&](int s) { return bar0 + 128 + 8 * s; };
  auto bar_oread = [&](int s) { return bar0 + 144 + 8 * s; };
  const uint32_t tmem_slot = bar0 + 192;
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_gen + L::OFF_BAR + 192);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) {
    for (int s = 0; s < 2; ++s) {
      mbar_init(bar_qfull(s), 1); mbar_init(bar_qempty(s), 1); mbar_init(bar_kfull(s), 1); mbar_init(bar_vfull(s), 1);
      mbar_init(bar_kvempty(s), 1); mbar_init(bar_sfull(s), 1); mbar_init(bar_ofull(s), 1); mbar_init(bar_oread(s), 8);
    }
    mbar_init(bar_rel, 1); mbar_init(bar_tfull, 1); mbar_init(bar_tread, 8); mbar_init(bar_pready, 8);
    fence_mbar_init();
  }
  if (warp == 9) {
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

  if (warp == 9) {
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
  } else if (warp == 8) {
    // ------------------------------------------------------------------------------------------ MMA issuer
    if (elect_one()) {
      auto issue_pv_item = [&](int m) {
        const int jm = m >> 1, mtm = m & 1, ksm = jm & 1;
        mbar_wait(bar_pready, (uint32_t)(m & 1));
        if (mtm == 0) mbar_wait(bar_vfull(ksm), (uint32_t)((jm >> 1) & 1));
        tc_fence_after();
        const uint32_t sbase = tmem_base + S_STRIDE * (uint32_t)mtm;
        issue_pv<HD, 13>(sbase + O_COL, sbase, v64(ksm), v64(ksm) + 208 * 128, false);
        umma_commit<1>(bar_ofull(mtm));
        if (mtm == 1) umma_commit<1>(bar_kvempty(ksm));  // K and V of the pair are free once its last PV has retired
      };
      mbar_wait(bar_rel, 0);
      for (int n = 0; n < N; ++n) {
        const int j = n >> 1, mt = n & 1, ks = j & 1;
        mbar_wait(bar_qfull(mt), (uint32_t)(j & 1));
        if (n > 0) mbar_wait(bar_tread, (uint32_t)((n - 1) & 1));  // T(n-1) has been read out of the T columns
        tc_fence_after();
        issue_qk<HD>(tmem_base + T_COL, q64(mt), q64(mt) + 128 * 128, r64, r16, IDESC_T);
        umma_commit<1>(bar_tfull);
        if (mt == 0) mbar_wait(bar_kfull(ks), (uint32_t)((j >> 1) & 1));
        if (n >= 2) mbar_wait(bar_oread(mt), (uint32_t)((j - 1) & 1));  // O(n-2) has been read out of this S buffer
        tc_fence_after();
        issue_qk<HD>(tmem_base + S_STRIDE * (uint32_t)mt, q64(mt), q64(mt) + 128 * 128, k64(ks), k64(ks) + 208 * 128, IDESC_S);
        umma_commit<1>(bar_sfull(mt));
        umma_commit<1>(bar_qempty(mt));
        if (n > 0) issue_pv_item(n - 1);
      }
      if (N > 0) issue_pv_item(N - 1);
    }
    __syncwarp();
  } else {
    // ------------------------------------------------------------------------------------------ softmax warps
    const int h = warp >> 2, wq = warp & 3;
    const int row = wq * 32 + lane;  // row of the 128-row tile == TMEM lane
    const uint32_t lane_base = tmem_base + ((uint32_t)(wq * 32) << 16);
    const uint64_t c12 = f2_pack(a.c1, a.c1);
    bool prev_live = false, prev_valid = false;
    __nv_bfloat16* prev_dst = a.out;

    auto store_prev = [&](int m) {  // O(m): normalise by the row sum of both key halves and store this thread's half row
      const int pb = m & 1;
      mbar_wait(bar_ofull(pb), (uint32_t)((m >> 1) & 1));
      tc_fence_after();
      if (prev_live) {
        const float inv = 1.f / (xsum(pb, 0)[row] + xsum(pb, 1)[row]);
        store_o_half<HD>(lane_base + S_STRIDE * (uint32_t)pb + O_COL + (uint32_t)((HD / 2) * h), inv, prev_dst, prev_valid);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_oread(pb));
    };

    for (int n = 0; n < N; ++n) {
      int wb, head;
      pair_of(n >> 1, wb, head);
      const int mt = n & 1;
      const int r = mt * 128 + row;  // row of the window (row-major 14x14); rows >= 196 are not tokens
      const bool is_tok = r < 196;
      const bool warp_live = mt * 128 + wq * 32 < 196;
      const int qh = is_tok ? r / 14 : 0, qw = is_tok ? r % 14 : 0;
      const uint32_t sbase = lane_base + S_STRIDE * (uint32_t)mt;

      // ---- 1. decomposed rel-pos terms of this query: bias_h for key rows 7h..7h+6, bias_w for the 14 key columns
      uint64_t bH2[7], bW2[7];
      mbar_wait(bar_tfull, (uint32_t)(n & 1));
      tc_fence_after();
      if (warp_live) {
        float tl[64];  // this query's T row; indexed per thread below -> local memory
        uint32_t v[64];
        tmem_ld_32x32p(lane_base + T_COL, v);
        tmem_ld_32x32p(lane_base + T_COL + 32, v + 32);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 64; ++i) tl[i] = __uint_as_float(v[i]);
        const float* th = tl + qh + 13 - 7 * h;  // q . rel_pos_h[qh - kh + 13], kh = 7h + i
        const float* tw = tl + 32 + qw + 13;     // q . rel_pos_w[qw - kw + 13]
#pragma unroll
        for (int i = 0; i < 7; ++i) {
          const float bh = th[-i] * kLog2e;
          bH2[i] = f2_pack(bh, bh);
          bW2[i] = f2_pack(tw[-2 * i] * kLog2e, tw[-2 * i - 1] * kLog2e);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_tread);

      // ---- 2. this thread's 98 logits (log2 units) and their maximum
      uint32_t s[98];
      float mloc = -INFINITY;
      mbar_wait(bar_sfull(mt), (uint32_t)((n >> 1) & 1));
      tc_fence_after();
      if (warp_live) {
        const uint32_t sa = sbase + (uint32_t)(98 * h);
        tmem_ld_32x32p(sa, s);
        tmem_ld_32x32p(sa + 32, s + 32);
        tmem_ld_32x32p(sa + 64, s + 64);
        uint32_t tail[2];
        tmem_ld_32x2(sa + 96, tail);
        tmem_ld_wait();
        s[96] = tail[0]; s[97] = tail[1];
#pragma unroll
        for (int j = 0; j < 98; j += 2) {
          float x0, x1;
          f2_unpack(f2_fma(f2_pack(__uint_as_float(s[j]), __uint_as_float(s[j + 1])), c12, f2_add(bH2[j / 14], bW2[(j % 14) / 2])), x0, x1);
          mloc = fmaxf(mloc, fmaxf(x0, x1));
          s[j] = __float_as_uint(x0); s[j + 1] = __float_as_uint(x1);
        }
      }
      xmax(n & 1, h)[row] = mloc;
      tc_fence_before();
      named_bar_sync(1, 256);  // every S column of this item has been read; both halves' maxima (and the sums of item n-1) are visible
      tc_fence_after();

      // ---- 3. O of the previous item
      if (n > 0) store_prev(n - 1);

      // ---- 4. probabilities
      float lsum = 0.f;
      if (warp_live) {
        const float m = fmaxf(xmax(n & 1, 0)[row], xmax(n & 1, 1)[row]);
        const uint64_t nm2 = f2_pack(-m, -m);
        uint64_t l2 = f2_pack(0.f, 0.f);
        uint32_t pk[49];
#pragma unroll
        for (int j = 0; j < 98; j += 2) {
          const uint64_t y2 = f2_add(f2_pack(__uint_as_float(s[j]), __uint_as_float(s[j + 1])), nm2);
          uint64_t p2;
          float p0, p1;
          if (j >= 98 - KPOLY) {
            float y0, y1;
            f2_unpack(y2, y0, y1);
            p2 = f2_exp2_poly(f2_pack(fmaxf(y0, -126.f), fmaxf(y1, -126.f)));
            f2_unpack(p2, p0, p1);
          } else {
            float y0, y1;
            f2_unpack(y2, y0, y1);
            p0 = ex2_approx(y0); p1 = ex2_approx(y1);
            p2 = f2_pack(p0, p1);
          }
          l2 = f2_add(l2, p2);
          pk[j >> 1] = pack_bf16x2(p0, p1);
        }
        const uint32_t pa = sbase + (uint32_t)(49 * h);  // P (bf16 pairs): key 98h + j -> column 49h + j / 2
        tmem_st_32x32p(pa, pk);
        tmem_st_32x16p(pa + 32, pk + 32);
        tmem_st_32x1(pa + 48, pk[48]);
        if (h == 1) {  // keys 196..207 of the 208-row box belong to the next window: P = 0
          tmem_st_32x4(sbase + 98, 0u, 0u, 0u, 0u);
          tmem_st_32x2(sbase + 102, 0u, 0u);
        }
        tmem_st_wait();
        float la, lb;
        f2_unpack(l2, la, lb);
        lsum = la + lb;
      }
      xsum(n & 1, h)[row] = lsum;
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_pready);

      const int win = wb % 25, b = wb / 25;
      const int y = (win / 5) * 14 + qh, x = (win % 5) * 14 + qw;
      prev_live = warp_live;
      prev_valid = is_tok && y < 64 && x < 64;  // window_unpartition drops the padded rows / columns (:287-288)
      prev_dst = a.out + ((size_t)b * 4096 + (size_t)(prev_valid ? y * 64 + x : 0)) * a.D + head * HD + (HD / 2) * h;
    }
    if (N > 0) {
      named_bar_sync(1, 256);  // the row sums of the last item
      store_prev(N - 1);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 9) tmem_dealloc<1>(tmem_base, 512);
}

}  // namespace inksam
