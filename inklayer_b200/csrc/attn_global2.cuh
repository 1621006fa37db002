// Global (64x64-token) encoder attention, warp-specialised: replaces Attention.forward + add_decomposed_rel_pos for
// the global blocks (SA/modeling/image_encoder.py:224-240,325-361; blocks 7/15/23/31 of ViT-H, build_sam.py:19).
//
// One CTA = 256 queries of one (image, head) = two 128-row tiles A and B that ping-pong through the tensor cores:
//   warps 0-3   softmax group A (thread = query row = TMEM lane)       warps 4-7   softmax group B
//   warp 10     TMA producer: Q tiles, rel-pos tables, then K and V tiles of 128 keys through two 2-slot rings
//   warps 8, 9  MMA issuers (one thread each) of group A and group B: issuing ~13 tcgen05.mma per sub-tile is itself a
//               serial cost (descriptor moves to uniform registers), so each group gets its own issuing thread.  The unit of work is a SUB-TILE t of 64 keys (= one row kh = t of the 64x64
//               key grid, so the rel-pos row bias is a single scalar per query and sub-tile); S is double buffered
//               per group, so S(t+1) is already in TMEM when softmax(t) ends and the MMAs never sit on a group's
//               critical path.  Per group:   S(0) S(1), then for t = 0..63:  PV(t) S(t+2)
// TMEM (512 columns): S_A0 [0,64) S_A1 [64,128) S_B0 [128,192) S_B1 [192,256)  O_A [256,336)  O_B [384,464);
// P(t) (bf16) is written back over the first 32 columns of S(t).  The prologue borrows all 512 columns for the
// rel-pos products T_h / T_w of both tiles (T[q][r] = q . rel_pos[r] on the UNSCALED q, image_encoder.py:234,354-355)
// and the K ring for the tables.
// Softmax: exp2 with a shift taken from the first 32 keys and no per-element maximum in the hot loop; two cold paths keep
// it exact for any logit range: a sub-tile whose sum reaches 2^100 (a key that may have hit the 2^120 exponent clamp) is
// redone with the shift raised to its true maximum BEFORE its P is used, and once the running sum passes 2^32 the shift
// moves up by floor(log2 l) -- both are exact power-of-two rescales of l and O.
#pragma once
#include "attn_tc.cuh"

namespace inksam {

template <int HD>
struct AttnGlobal2 {
  static constexpr int TILE = AttnTile<HD>::bytes(128);  // 20480 (hd 80) or 16384 (hd 64)
  static constexpr int BH_PITCH = 66;                     // floats per biasH row
  static constexpr int OFF_Q = 0;                         // Q_A, Q_B
  static constexpr int OFF_K = 2 * TILE;                  // K ring, 2 slots
  static constexpr int OFF_V = 4 * TILE;                  // V ring, 2 slots
  static constexpr int OFF_BH = 6 * TILE;                 // biasH [256][66] fp32
  static constexpr int TW_PITCH = 97;                     // floats per row of a warp's T_w scratch (odd: conflict-free writes)
  static constexpr int TW_WARP = 32 * TW_PITCH * 4;       // prologue scratch of one softmax warp; the 8 of them overlay biasH
  static constexpr int OFF_BAR = OFF_BH + (8 * TW_WARP > 256 * BH_PITCH * 4 ? 8 * TW_WARP : 256 * BH_PITCH * 4);
  static constexpr int SMEM_BYTES = OFF_BAR + 256 + 1024;
  static constexpr int THREADS = 352;
};

template <int HD, int KPOLY>
__global__ void __launch_bounds__(352, 1)
attn_global2_kernel(const __grid_constant__ CUtensorMap tm64, const __grid_constant__ CUtensorMap tm16,
                    const __grid_constant__ CUtensorMap tr64, const __grid_constant__ CUtensorMap tr16, const AttnArgs a) {
  using L = AttnGlobal2<HD>;
  constexpr int TILE = L::TILE;
  constexpr uint32_t TILE_TX = 128 * HD * 2;
  constexpr uint32_t IDESC_T = make_idesc(1u, 128, 128);
  constexpr uint32_t IDESC_S = make_idesc(1u, 128, 64);
  constexpr int NT = 32;   // key tiles of 128 (TMA / smem ring granularity)
  constexpr int NS = 64;   // sub-tiles of 64 keys (MMA / softmax granularity)
  constexpr int kPoly = KPOLY;  // keys per sub-tile whose exp2 runs on the FMA pipe (even, <= 32)

  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  auto q_tile = [&](int g) { return smem_base + L::OFF_Q + g * TILE; };
  auto k_tile = [&](int s) { return smem_base + L::OFF_K + s * TILE; };
  auto v_tile = [&](int s) { return smem_base + L::OFF_V + s * TILE; };
  float* biasH = reinterpret_cast<float*>(smem_gen + L::OFF_BH);
  const uint32_t bar0 = smem_base + L::OFF_BAR;
  // barrier map (8 bytes each)
  const uint32_t bar_q = bar0, bar_rel = bar0 + 8, bar_tdone = bar0 + 16, bar_tread = bar0 + 24;
  auto bar_kfull = [&](int s) { return bar0 + 32 + 8 * s; };
  auto bar_kempty = [&](int s) { return bar0 + 48 + 8 * s; };
  auto bar_vfull = [&](int s) { return bar0 + 64 + 8 * s; };
  auto bar_vempty = [&](int s) { return bar0 + 80 + 8 * s; };
  auto bar_sfull = [&](int g, int buf) { return bar0 + 96 + 8 * (2 * g + buf); };
  auto bar_pready = [&](int g, int buf) { return bar0 + 128 + 8 * (2 * g + buf); };
  auto bar_pvdone = [&](int g, int buf) { return bar0 + 160 + 8 * (2 * g + buf); };
  const uint32_t tmem_slot = bar0 + 192;
  auto bar_done = [&](int g) { return bar0 + 200 + 8 * g; };  // teardown: an issuer's last commit-armed arrivals have landed
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_gen + L::OFF_BAR + 192);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int qp = blockIdx.x & 15;
  const int head = (blockIdx.x >> 4) % a.heads;
  const int b = (blockIdx.x >> 4) / a.heads;
  const int row_q = b * 4096 + qp * 256;
  const int col_q = head * HD, col_k = a.D + col_q, col_v = 2 * a.D + col_q;

  if (tid == 0) {
    mbar_init(bar_q, 1); mbar_init(bar_rel, 1); mbar_init(bar_tdone, 2); mbar_init(bar_tread, 8);
    mbar_init(bar_done(0), 1); mbar_init(bar_done(1), 1);
    for (int s = 0; s < 2; ++s) {
      mbar_init(bar_kfull(s), 1); mbar_init(bar_kempty(s), 2); mbar_init(bar_vfull(s), 1); mbar_init(bar_vempty(s), 2);
      for (int buf = 0; buf < 2; ++buf) {
        mbar_init(bar_sfull(s, buf), 1); mbar_init(bar_pready(s, buf), 4); mbar_init(bar_pvdone(s, buf), 1);
      }
    }
    fence_mbar_init();
  }
  if (warp == 10) {
    if (lane == 0) {
      tma_prefetch_desc(&tm64); tma_prefetch_desc(&tr64);
      if constexpr (HD == 80) { tma_prefetch_desc(&tm16); tma_prefetch_desc(&tr16); }
    }
    __syncwarp();
    tmem_alloc<1>(tmem_slot, 512);
    tmem_relinquish<1>();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;
  // Debug trace of CTA 0 (AttnArgs::trace, tools/global_trace.py): event e of sub-tile t <- clock64().  Compiled in with
  // -DINKSAM_ATTN_TRACE only: even disabled at run time the hooks cost ~10 % of this kernel's time.
  auto trace = [&](int t, int e) {
#ifdef INKSAM_ATTN_TRACE
    if (a.trace != nullptr && blockIdx.x == 0) a.trace[t * 16 + e] = clock64();
#endif
  };
  pdl_launch_dependents();  // the proj GEMM (launched with PDL) may run its prologue under this kernel's tail
  auto S_COL = [](int g, int buf) { return (uint32_t)(128 * g + 64 * buf); };
  auto O_COL = [](int g) { return (uint32_t)(256 + 128 * g); };

  if (warp == 10) {
    // ---------------------------------------------------------------------------------------- TMA producer
    if (lane == 0) {
      auto load_tile = [&](uint32_t dst, const CUtensorMap* m64, const CUtensorMap* m16, uint32_t bar, int col, int row) {
        tma_load_2d(dst, m64, bar, col, row);
        if constexpr (HD == 80) tma_load_2d(dst + 128 * 128, m16, bar, col + 64, row);
      };
      mbar_arrive_expect_tx(bar_q, 2 * TILE_TX);
      load_tile(q_tile(0), &tm64, &tm16, bar_q, col_q, row_q);
      load_tile(q_tile(1), &tm64, &tm16, bar_q, col_q, row_q + 128);
      mbar_arrive_expect_tx(bar_rel, 2 * TILE_TX);
      load_tile(k_tile(0), &tr64, &tr16, bar_rel, 0, 0);     // rel_pos_h rows
      load_tile(k_tile(1), &tr64, &tr16, bar_rel, 0, 128);   // rel_pos_w rows
      // V (HD == 80): FIVE [128 keys][16 columns] tiles with 32-byte rows / 32-byte swizzle, 4096 bytes apart = one MN-major
      // operand of N = 80, so that P V takes ONE tcgen05.mma per 16 keys instead of a 64-column and a 16-column one: the
      // tensor pipe spends >= 46 cycles on an MMA however small it is (tools/micro/mma_issue.cu), and with 13 MMAs per
      // 64-key sub-tile and group that issue cost, not the MUFU pipe, bounded this kernel.
      auto load_v = [&](uint32_t dst, uint32_t bar, int row) {
        if constexpr (HD == 80) {
#pragma unroll
          for (int t = 0; t < 5; ++t) tma_load_2d(dst + t * 4096, &tm16, bar, col_v + 16 * t, row);
        } else {
          tma_load_2d(dst, &tm64, bar, col_v, row);
        }
      };
      for (int s = 0; s < 2; ++s) {
        mbar_arrive_expect_tx(bar_vfull(s), TILE_TX);
        load_v(v_tile(s), bar_vfull(s), b * 4096 + s * 128);
      }
      mbar_wait(bar_tdone, 0);  // the T MMAs have read the rel tables: the K ring is free
      for (int j = 0; j < NT; ++j) {
        const int s = j & 1;
        const uint32_t par = (uint32_t)(((j >> 1) & 1) ^ 1);
        if (j >= 2) mbar_wait(bar_kempty(s), par);
        mbar_arrive_expect_tx(bar_kfull(s), TILE_TX);
        load_tile(k_tile(s), &tm64, &tm16, bar_kfull(s), col_k, b * 4096 + j * 128);
        if (j >= 2) {
          mbar_wait(bar_vempty(s), par);
          mbar_arrive_expect_tx(bar_vfull(s), TILE_TX);
          load_v(v_tile(s), bar_vfull(s), b * 4096 + j * 128);
        }
      }
    }
    __syncwarp();
  } else if (warp >= 8) {
    // ---------------------------------------------------------------------------------------- MMA issuer of group g
    if (elect_one()) {
      const int g = warp - 8;
      const uint32_t qa = q_tile(g), qb = q_tile(g) + 128 * 128;
      // S(t) = Q_g . K[keys 64t .. 64t+63]^T into S buffer t & 1; K tile t >> 1 sits in ring slot (t >> 1) & 1
      auto issue_s = [&](int t) {
        const uint32_t kt = k_tile((t >> 1) & 1);
        issue_qk<HD>(tmem_base + S_COL(g, t & 1), qa, qb, kt + (t & 1) * 64 * 128, kt + 128 * 128 + (t & 1) * 64 * 32, IDESC_S);
      };
      mbar_wait(bar_q, 0);
      mbar_wait(bar_rel, 0);
      tc_fence_after();
      // T_h -> columns 256g, T_w -> columns 256g + 128
      issue_qk<HD>(tmem_base + 256 * g, qa, qb, k_tile(0), k_tile(0) + 128 * 128, IDESC_T);
      issue_qk<HD>(tmem_base + 256 * g + 128, qa, qb, k_tile(1), k_tile(1) + 128 * 128, IDESC_T);
      umma_commit<1>(bar_tdone);
      mbar_wait(bar_tread, 0);  // both groups have pulled their bias terms out of TMEM (T_B overlaps O_A)
      tc_fence_after();
      mbar_wait(bar_kfull(0), 0);
      tc_fence_after();
      for (int t = 0; t < 2; ++t) {
        issue_s(t);
        umma_commit<1>(bar_sfull(g, t));
      }
      umma_commit<1>(bar_kempty(0));
      for (int t = 0; t < NS; ++t) {
        const int j = t >> 1, buf = t & 1, vs = j & 1;
        const uint32_t par = (uint32_t)((t >> 1) & 1);
        if (buf == 0) {
          mbar_wait(bar_vfull(vs), (uint32_t)((j >> 1) & 1));
          if (j + 1 < NT) mbar_wait(bar_kfull((j + 1) & 1), (uint32_t)(((j + 1) >> 1) & 1));
        }
        if (g == 0) trace(t, 9);
        mbar_wait(bar_pready(g, buf), par);
        tc_fence_after();
        if (g == 0) trace(t, 10);
        const uint32_t vt = v_tile(vs);
        if constexpr (HD == 80) {
          // MN-major B operand, 32-byte swizzle: five 16-column atoms along N (leading byte offset 4096), 8-key groups 256
          // bytes apart along K; the sub-tile's 64 keys start 64 * 32 bytes into the tile; 16 keys per MMA = 512 bytes
          constexpr uint32_t ID80 = make_idesc_bmn(1u, 128, 80);
          uint64_t vd = make_sw32_desc(vt + buf * 64 * 32);
          vd = (vd & ~(uint64_t)0x3FFF0000ull) | ((uint64_t)(4096 >> 4) << 16);
#pragma unroll
          for (int ks = 0; ks < 4; ++ks)
            umma_ts(tmem_base + O_COL(g), tmem_base + S_COL(g, buf) + 8u * ks, vd + (uint64_t)(ks * (16 * 32 / 16)), ID80,
                    (uint32_t)(t != 0 || ks != 0));
        } else {
          issue_pv<HD, 4>(tmem_base + O_COL(g), tmem_base + S_COL(g, buf), vt + buf * 64 * 128, vt + 128 * 128 + buf * 64 * 32,
                          t != 0);
        }
        umma_commit<1>(bar_pvdone(g, buf));
        if (t + 2 < NS) {
          issue_s(t + 2);  // executes after the PV above (in order), the last reader of P(t) in this S buffer
          umma_commit<1>(bar_sfull(g, buf));
        }
        if (g == 0) trace(t, 11);
        if (buf == 1) {
          umma_commit<1>(bar_vempty(vs));
          if (j + 1 < NT) umma_commit<1>(bar_kempty((j + 1) & 1));
        }
      }
      // Teardown: nobody waits for this thread's last v_empty / k_empty arrivals.  One more commit, waited for here: when it
      // has landed, every earlier commit-armed arrival of this thread has landed too.
      umma_commit<1>(bar_done(g));
      mbar_wait(bar_done(g), 0);
    }
    __syncwarp();
  } else {
    // ---------------------------------------------------------------------------------------- softmax groups
    const int g = warp >> 2;          // 0: tile A, 1: tile B
    const int r = tid & 127;          // row inside the tile == TMEM lane
    const uint32_t lane_base = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
    const int qt = 2 * qp + g;
    const int qw = r & 63;
    const int qh = 2 * qt + (r >> 6);  // warp-uniform
    float* bh_row = biasH + (size_t)(g * 128 + r) * L::BH_PITCH;

    mbar_wait(bar_tdone, 0);
    tc_fence_after();
    uint64_t bW2[32];  // (bias_w[2j], bias_w[2j+1]) * log2e for this query, packed for f32x2 arithmetic
    {
      // T_w[q][qw - kw + 63], kw = 0..63: a 64-column window of the thread's T_w row that starts at column qw, i.e. at a
      // different column in every lane.  TMEM addresses are warp-uniform, so the 95 columns the warp's 32 queries need
      // (qw = qw0 + lane) go through a shared-memory scratch: row pitch 97 floats -> the writes are conflict-free, the
      // skewed reads 2-way.  (Round 1 did this transpose in a per-thread local array: 64 loads x 32 lines per warp, 4 %
      // of the kernel.)  The scratch of the 8 warps overlays biasH, hence the 256-thread barrier before biasH is written.
      float* tw = reinterpret_cast<float*>(smem_gen + L::OFF_BH + warp * L::TW_WARP) + lane * L::TW_PITCH;
      const uint32_t qw0 = (uint32_t)(qw & 32);
      uint32_t v[32];
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        tmem_ld_32x32(lane_base + 256 * g + 128 + qw0 + c * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) tw[c * 32 + i] = __uint_as_float(v[i]);
      }
      __syncwarp();
      const float* tp = tw + lane + 63;  // column qw + 63 of the row
#pragma unroll
      for (int i = 0; i < 64; i += 2) bW2[i >> 1] = f2_pack(tp[-i] * kLog2e, tp[-i - 1] * kLog2e);
      // T_h[q][qh - kh + 63]: qh is uniform over the warp, so the 64 needed columns start at column qh for all lanes
      uint32_t h0[32], h1[32];
      tmem_ld_32x32(lane_base + 256 * g + (uint32_t)qh, h0);
      tmem_ld_32x32(lane_base + 256 * g + (uint32_t)qh + 32, h1);
      tmem_ld_wait();
      named_bar_sync(1, 256);  // every softmax warp has read its scratch
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        bh_row[63 - i] = __uint_as_float(h0[i]) * kLog2e;
        bh_row[31 - i] = __uint_as_float(h1[i]) * kLog2e;
      }
    }
    tc_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive(bar_tread);

    float m_used = 0.f, l = 0.f;
    uint64_t l2 = f2_pack(0.f, 0.f);  // two partial row sums (even / odd keys), folded into l after every sub-tile
    const float c1 = a.c1;
    const uint64_t c12 = f2_pack(c1, c1);
    const uint32_t o_addr = lane_base + O_COL(g);
#pragma unroll 1
    for (int t = 0; t < NS; ++t) {
      const int buf = t & 1;
      const uint32_t par = (uint32_t)((t >> 1) & 1);
      const uint32_t s_addr = lane_base + S_COL(g, buf);
      const float bh = bh_row[t];  // key row kh = t
      mbar_wait(bar_sfull(g, buf), par);
      tc_fence_after();
      if (g == 0 && lane == 0) trace(t, 4 + (warp & 3));
      uint32_t v0[32], v1[32];
      tmem_ld_32x32(s_addr, v0);
      tmem_ld_32x32(s_addr + 32, v1);
      tmem_ld_wait();
      if (t == 0) {  // shift = max over the first 32 keys
        float mx = -INFINITY;
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          float b0, b1;
          f2_unpack(bW2[i >> 1], b0, b1);
          mx = fmaxf(mx, fmaxf(fmaf(__uint_as_float(v0[i]), c1, b0 + bh), fmaf(__uint_as_float(v0[i + 1]), c1, b1 + bh)));
        }
        m_used = mx;
      }
      const float bb = bh - m_used;
      const uint64_t bb2 = f2_pack(bb, bb);
      uint32_t pk[32];
      // two keys per instruction: x = s * c1 + (bias_w + bias_h - shift) as fma.rn.f32x2 / add.rn.f32x2
#pragma unroll
      for (int i = 0; i < 32; i += 2) {
        float x0, x1;
        f2_unpack(f2_fma(f2_pack(__uint_as_float(v0[i]), __uint_as_float(v0[i + 1])), c12, f2_add(bW2[i >> 1], bb2)), x0, x1);
        // no clamp: a key more than 128 log2 units above the shift gives +inf, which makes the sub-tile's sum +inf and
        // sends it through the cold path below, where the shift is raised and the sub-tile is exponentiated again
        const float p0 = ex2_approx(x0), p1 = ex2_approx(x1);
        l2 = f2_add(l2, f2_pack(p0, p1));
        pk[i >> 1] = pack_bf16x2(p0, p1);
      }
      // The MUFU pipe (16 ex2 per clock and SM) is the busiest unit of this kernel: kPoly of the 64 keys of a sub-tile
      // take their exp2 from an FMA-pipe polynomial instead.
#pragma unroll
      for (int i = 0; i < 32; i += 2) {
        float x0, x1;
        f2_unpack(f2_fma(f2_pack(__uint_as_float(v1[i]), __uint_as_float(v1[i + 1])), c12, f2_add(bW2[16 + (i >> 1)], bb2)), x0, x1);
        float p0, p1;
        if (i >= 32 - kPoly) {  // the polynomial needs its argument in [-126, 126]; 126 still trips the cold path (sum >= 2^100)
          f2_unpack(f2_exp2_poly(f2_pack(fmaxf(fminf(x0, 126.f), -126.f), fmaxf(fminf(x1, 126.f), -126.f))), p0, p1);
        } else {
          p0 = ex2_approx(x0); p1 = ex2_approx(x1);
        }
        l2 = f2_add(l2, f2_pack(p0, p1));
        pk[16 + (i >> 1)] = pack_bf16x2(p0, p1);
      }
      float lsub;
      {
        float la, lb;
        f2_unpack(l2, la, lb);
        lsub = la + lb;
        l2 = f2_pack(0.f, 0.f);
      }
      // A sub-tile whose sum reaches 2^100 (or overflowed to +inf; !(x <= c) also catches NaN) holds a key more than 94
      // log2 units above the shift: before P(t) is handed to the tensor cores the shift is raised to this sub-tile's true
      // maximum, everything accumulated so far is rescaled by the exact power of two, and the sub-tile is exponentiated
      // again.  Exact for any logit range; cold.
      if (__any_sync(0xffffffffu, !(lsub <= 1.2676506e30f))) {
        const bool hot = !(lsub <= 1.2676506e30f);
        float mx = -INFINITY;
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          float b0, b1, d0, d1;
          f2_unpack(bW2[i >> 1], b0, b1);
          f2_unpack(bW2[16 + (i >> 1)], d0, d1);
          mx = fmaxf(mx, fmaxf(fmaf(__uint_as_float(v0[i]), c1, b0), fmaf(__uint_as_float(v0[i + 1]), c1, b1)));
          mx = fmaxf(mx, fmaxf(fmaf(__uint_as_float(v1[i]), c1, d0), fmaf(__uint_as_float(v1[i + 1]), c1, d1)));
        }
        const float delta = hot ? ceilf(mx + bb) : 0.f;  // integer, > 94: an exact power-of-two rescale
        const float alpha = hot ? ex2_approx(-delta) : 1.f;  // flushes to 0 when the old terms no longer matter
        if (t > 0) {
          mbar_wait(bar_pvdone(g, buf ^ 1), (uint32_t)(((t - 1) >> 1) & 1));  // PV(t-1): O holds every earlier sub-tile
          tc_fence_after();
          uint32_t v[16];
#pragma unroll
          for (int c = 0; c < HD / 16; ++c) {
            tmem_ld_32x16(o_addr + c * 16, v);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 16; ++i) v[i] = __float_as_uint(__uint_as_float(v[i]) * alpha);
            tmem_st_32x16(o_addr + c * 16, v);
          }
          tmem_st_wait();
          tc_fence_before();
        }
        m_used += delta;
        l *= alpha;
        const float bn = bb - delta;
        lsub = 0.f;
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          float b0, b1, d0, d1;
          f2_unpack(bW2[i >> 1], b0, b1);
          f2_unpack(bW2[16 + (i >> 1)], d0, d1);
          const float p0 = ex2_approx(fmaf(__uint_as_float(v0[i]), c1, b0 + bn));
          const float p1 = ex2_approx(fmaf(__uint_as_float(v0[i + 1]), c1, b1 + bn));
          const float p2 = ex2_approx(fmaf(__uint_as_float(v1[i]), c1, d0 + bn));
          const float p3 = ex2_approx(fmaf(__uint_as_float(v1[i + 1]), c1, d1 + bn));
          lsub += (p0 + p1) + (p2 + p3);
          pk[i >> 1] = pack_bf16x2(p0, p1);
          pk[16 + (i >> 1)] = pack_bf16x2(p2, p3);
        }
      }
      l += lsub;
      tmem_st_32x32(s_addr, pk);  // P(t) over the first 32 columns of S(t), all of which have been read
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_pready(g, buf));
      if (g == 0 && lane == 0) trace(t, warp & 3);
      // The running sum bounds every probability so far: once it passes 2^32 the shift moves up by log2(l) (an exact
      // power-of-two rescale of l and O), so no per-element maximum has to be tracked in the hot loop.
      const bool need = l > 4294967296.f;
      if (__any_sync(0xffffffffu, need)) {  // exact but cold: move the shift up and rescale the accumulator
        mbar_wait(bar_pvdone(g, buf), par);  // PV(t) has been accumulated with the old shift
        tc_fence_after();
        const float tmax = need ? (float)((int)((__float_as_uint(l) >> 23) & 0xffu) - 127) : 0.f;  // floor(log2 l)
        const float alpha = need ? ex2_approx(-tmax) : 1.f;
        if (need) m_used += tmax;
        l *= alpha;
        uint32_t v[16];
#pragma unroll
        for (int c = 0; c < HD / 16; ++c) {
          tmem_ld_32x16(o_addr + c * 16, v);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = __float_as_uint(__uint_as_float(v[i]) * alpha);
          tmem_st_32x16(o_addr + c * 16, v);
        }
        tmem_st_wait();
        tc_fence_before();
      }
    }
    mbar_wait(bar_pvdone(g, 1), 1u);  // PV of sub-tile 63 (the 32nd completion on buffer 1)
    tc_fence_after();
    store_o_row<HD>(o_addr, 1.f / l, a.out + (size_t)(row_q + g * 128 + r) * a.D + col_q, true);
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 10) tmem_dealloc<1>(tmem_base, 512);
}

}  // namespace inksam
