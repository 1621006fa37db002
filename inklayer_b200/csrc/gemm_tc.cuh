// Persistent, warp-specialised tcgen05 GEMM for sm_100a.
//
//   out[M,N] = epilogue( A[M,K] * W[N,K]^T )      A, W K-major; bf16 (kind::f16) or fp32-as-tf32 (kind::tf32)
//
// Structure (one CTA per SM, or one CTA pair per 2 SMs when CG == 2):
//   warp 0      TMA producer: A/W tiles -> 128B-swizzled smem ring (STAGES deep), mbarrier complete_tx
//   warp 1      MMA issuer: one thread issues tcgen05.mma (UMMA M = 128*CG, N = BN, K = 32 bytes),
//               accumulators live in TMEM, double buffered (2 x BN columns)
//   warp 2      TMEM allocator
//   warps 4..11 epilogue: tcgen05.ld (thread == output row; warp w reads TMEM lanes 32*(w%4).., the two warps of a
//               lane quarter split the tile's columns), fused bias / GELU / residual / LayerNorm, direct 32-byte
//               vector stores.  Residual / row-add operands are prefetched one 32-column chunk ahead so that their
//               DRAM latency is off the critical path.  The row-wise LayerNorm epilogues use warps 4..7 only.
//
// This one kernel replaces every nn.Linear / 1x1 conv / 3x3 conv / conv-transpose GEMM on the hot path:
//   reference SA/modeling/image_encoder.py:227 (qkv), :238 (proj), common.py:21-26 (MLP),
//   image_encoder.py:88-104 (neck), :391-395 (patch embed), and the image-side linears of
//   SA/modeling/transformer.py:218-221 and mask_decoder.py:53-59.
#pragma once
#include <cuda_fp16.h>
#include "ptx.cuh"
#include "gemm_args.h"

namespace inksam {

// Exact (erf-form) GELU, nn.GELU() default (SA/modeling/common.py:18, image_encoder.py:30):
//   gelu(x) = x/2 + |x|/2 * erf(|x|/sqrt2), erf by Abramowitz-Stegun 7.1.26 (|error| <= 1.5e-7, below fp32 erff's
// own error after the 0.5*x scaling), branch-free: 2 MUFU (rcp, ex2) + 12 FMA-pipe instructions per element.
__device__ __forceinline__ float gelu_erf(float x) {
  const float z = fabsf(x) * 0.70710678118654752f;
  float t;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(0.3275911f, z, 1.0f)));
  float p = fmaf(1.061405429f, t, -1.453152027f);
  p = fmaf(p, t, 1.421413741f);
  p = fmaf(p, t, -0.284496736f);
  p = fmaf(p, t, 0.254829592f);
  p *= t;
  const float e = ex2_approx(z * z * -1.4426950408889634f);
  const float erf_abs = fmaf(-p, e, 1.0f);
  const float hx = 0.5f * x;
  return fmaf(fabsf(hx), erf_abs, hx);
}

// The same on two values per instruction (fma.rn.f32x2 / mul.rn.f32x2): 7 FMA-pipe instructions per element instead
// of 12.  The epilogue warps share their schedulers with the thread that issues tcgen05.mma, so every instruction
// saved here is an issue slot returned to the MMA stream.  Same operation order as gelu_erf: identical results.
__device__ __forceinline__ void gelu_erf2(float& x0, float& x1) {
  const uint64_t z = f2_pack(fabsf(x0) * 0.70710678118654752f, fabsf(x1) * 0.70710678118654752f);
  float d0, d1;
  f2_unpack(f2_fma(f2_pack(0.3275911f, 0.3275911f), z, f2_pack(1.0f, 1.0f)), d0, d1);
  float t0, t1;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t0) : "f"(d0));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t1) : "f"(d1));
  const uint64_t t = f2_pack(t0, t1);
  uint64_t p = f2_fma(f2_pack(1.061405429f, 1.061405429f), t, f2_pack(-1.453152027f, -1.453152027f));
  p = f2_fma(p, t, f2_pack(1.421413741f, 1.421413741f));
  p = f2_fma(p, t, f2_pack(-0.284496736f, -0.284496736f));
  p = f2_fma(p, t, f2_pack(0.254829592f, 0.254829592f));
  p = f2_mul(p, t);
  float a0, a1;
  f2_unpack(f2_mul(f2_mul(z, z), f2_pack(-1.4426950408889634f, -1.4426950408889634f)), a0, a1);
  const uint64_t e = f2_pack(ex2_approx(a0), ex2_approx(a1));
  float p0, p1, e0, e1;
  f2_unpack(p, p0, p1);
  f2_unpack(e, e0, e1);
  const float h0 = 0.5f * x0, h1 = 0.5f * x1;
  x0 = fmaf(fabsf(h0), fmaf(-p0, e0, 1.0f), h0);
  x1 = fmaf(fabsf(h1), fmaf(-p1, e1, 1.0f), h1);
}

// epilogue warps per CTA: 8 (two per TMEM lane quarter, splitting the columns) except for the encoder-neck epilogues
// with whole-row LayerNorm statistics, which keep one warp per lane quarter (EPI_RESID_LN_F32 splits the row and
// exchanges the partial statistics through shared memory)
__host__ __device__ constexpr int gemm_epi_warps(int epi) {
  return (epi == EPI_LN_BF16_PAD || epi == EPI_LN_F32_NCHW) ? 4 : 8;
}

// ST != 0 (the epilogues with fp32 row-major output): every epilogue warp owns two 2 KB staging buffers from which its
// 32-row x 16-column fp32 sub-tiles leave by TMA store (32 KB in all; the CTA-pair ring keeps its 6 stages of 32 KB).
// ST == 2 (EPI_RESID_LN_F32) also keeps the LayerNorm partials of the unstaged layout.
template <int CG, int TF32, int BN, int ST = 0>
struct GemmCfg {
  static constexpr int BM = 128;                 // rows per CTA (UMMA M = BM * CG)
  static constexpr int BNL = BN / CG;            // W rows loaded by each CTA
  static constexpr int A_BYTES = BM * 128;       // one stage of A: 128 rows x 128 bytes
  static constexpr int B_BYTES = BNL * 128;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int STAGING_BYTES = ST == 3 ? 8 * 2048 : (ST ? 8 * 2 * 2048 : 0);  // ST == 3: one 2 KB buffer per epilogue warp
  static constexpr int AUX_BYTES = ST == 1 ? 256 /*barriers*/ + 1024 /*the tile's bias slice (fp32 residual epilogue)*/
                                           : 256 /*barriers*/ + 4096 /*LN partials*/ + 3072 /*epilogue constants*/;
  static constexpr int RING_BUDGET = ST ? (227 * 1024 - 1024 - AUX_BYTES - STAGING_BYTES) : 192 * 1024;
  static constexpr int STAGES = RING_BUDGET / STAGE_BYTES > 8 ? 8 : RING_BUDGET / STAGE_BYTES;
  static constexpr int KELEMS = TF32 ? 32 : 64;  // elements per 128-byte smem row
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + STAGING_BYTES + 1024 /*align slack*/ + AUX_BYTES;
  static constexpr uint32_t IDESC = make_idesc(TF32 ? 2u : 1u, BM * CG, BN);
  static constexpr int TMEM_COLS = 2 * BN;       // double-buffered accumulators (power of two: 256 or 512)
};

__host__ __device__ constexpr bool epi_is_resid(int epi) { return epi == EPI_BIAS_RESID_F32 || epi == EPI_BIAS_RESID_STATS_F32; }
__host__ __device__ constexpr bool epi_is_fold(int epi) { return epi == EPI_FOLD_BF16 || epi == EPI_FOLD_GELU_BF16; }
// fp32 row-major outputs leave through staging buffers and TMA stores: thread-per-row 32-byte stores put 32 sectors of 32
// different lines into every instruction and held the decoder's image-side GEMMs at 70 % L1/LSU throughput and 38-49 % of
// the HBM peak (ncu, round 2).
__host__ __device__ constexpr int gemm_staged(int epi) {
  return (epi_is_resid(epi) || epi == EPI_BIAS_ROWADD_F32 || epi == EPI_BIAS_F32) ? 1
         : epi == EPI_RESID_LN_F32 ? 2
         // bf16 row-major outputs (qkv of the global blocks, mlp.lin1): one staging buffer per warp (the 6-stage ring and the
         // epilogue constants leave room for 16 KB).  In the step, mlp.lin1 with its thread-per-row stores removed ran 110 us
         // per launch faster (tools/encoder_ab.py): 32 sectors of 32 lines per store instruction again.
         : (epi == EPI_BIAS_BF16 || epi == EPI_BIAS_GELU_BF16 || epi_is_fold(epi)) ? 3 : 0;
}

template <int CG, int TF32, int BN, int EPI>
__global__ void __launch_bounds__(384, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const __grid_constant__ CUtensorMap tmO, const __grid_constant__ CUtensorMap tmX, const GemmArgs g) {
  using C = GemmCfg<CG, TF32, BN, gemm_staged(EPI)>;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t staging_base = smem_base + C::STAGES * C::STAGE_BYTES;  // 1024-byte aligned (stages are 16 KB multiples)
  const uint32_t bar_base = staging_base + C::STAGING_BYTES;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (C::STAGES + s); };
  auto tfull_bar = [&](int s) { return bar_base + 8u * (2 * C::STAGES + s); };
  auto tempty_bar = [&](int s) { return bar_base + 8u * (2 * C::STAGES + 2 + s); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * C::STAGES + 4);
  volatile uint32_t* tmem_slot_ptr =
      reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = (CG == 2) ? cluster_ctarank() : 0u;
  const bool leader = (rank == 0);

  const int tiles_m = (g.M + C::BM * CG - 1) / (C::BM * CG);
  const int tiles_n = (g.N + BN - 1) / BN;
  const int num_tiles = tiles_m * tiles_n;
  const int num_kb = g.a_conv3x3 ? 9 * (g.conv_c / C::KELEMS) : (g.K + C::KELEMS - 1) / C::KELEMS;
  const int cluster_id = blockIdx.x / CG;
  const int num_clusters = gridDim.x / CG;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < C::STAGES; ++s) {
      mbar_init(full_bar(s), 1);   // one expect_tx arrive (CG == 2: the leader's covers the bytes of both CTAs)
      mbar_init(empty_bar(s), 1);  // one tcgen05.commit
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(tfull_bar(s), 1);          // one tcgen05.commit
      mbar_init(tempty_bar(s), gemm_epi_warps(EPI) * CG);  // one arrive per epilogue warp of the pair
    }
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc<CG>(tmem_slot, C::TMEM_COLS);
    tmem_relinquish<CG>();
  }
  // per-channel constants of the conv-transpose epilogue, read back with volatile broadcast LDS.128 so that the
  // compiler does not hoist 192 loop-invariant values into registers
  const uint32_t econst = bar_base + 256 + 4096;  // (not present in the staged layout, which has no such epilogue)
  if constexpr (EPI == EPI_UPSCALE1) {
    float* ec = reinterpret_cast<float*>(smem_raw + (econst - smem_u32(smem_raw)));
    if (threadIdx.x >= 128 && threadIdx.x < 192) {
      const int c = threadIdx.x - 128;
      ec[c] = g.bias[c]; ec[64 + c] = g.ln_w[c]; ec[128 + c] = g.ln_b[c];
    }
  }
  // EPI_RESID_LN_F32: bias, LayerNorm weight and bias of the 256 columns.  As __ldg loads in the epilogue they sat on its
  // critical path behind the streaming residual loads (L1 misses: a fifth of the epilogue warps' stall samples, ncu).
  if constexpr (EPI == EPI_RESID_LN_F32) {
    float* ec = reinterpret_cast<float*>(smem_raw + (econst - smem_u32(smem_raw)));
    if (threadIdx.x >= 128) {
      const int c = threadIdx.x - 128;  // 0..255
      ec[c] = g.bias[c]; ec[256 + c] = g.ln_w[c]; ec[512 + c] = g.ln_b[c];
    }
  }
  tc_fence_before();
  if constexpr (CG == 2) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;
  // PDL: everything above (barrier init, TMEM allocation, descriptor prefetch) touched no global data and may have
  // run under the tail of the preceding kernel; from here on its results are needed (and its inputs may be overwritten)
  pdl_grid_dep_wait();
  pdl_launch_dependents();

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    if (elect_one()) {  // elect (not lane == 0): the compiler then emits straight-line uniform-datapath code
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
        const int m_blk = tile / tiles_n, n_blk = tile % tiles_n;
        const int m0 = m_blk * C::BM * CG + (int)rank * C::BM;
        const int n0 = n_blk * BN + (int)rank * C::BNL;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1u);
          const uint32_t a_dst = smem_base + stage * C::STAGE_BYTES;
          const uint32_t b_dst = a_dst + C::A_BYTES;
          const int k0 = kb * C::KELEMS;
          if constexpr (CG == 1) {
            mbar_arrive_expect_tx(full_bar(stage), C::STAGE_BYTES);
            if (g.a_conv3x3) {
              const int cpt = g.conv_c / C::KELEMS;  // k-blocks per filter tap
              const int tap = kb / cpt, c0 = (kb % cpt) * C::KELEMS;
              const int b = m0 >> 12, y0 = (m0 & 4095) >> 6;
              tma_load_3d(a_dst, &tmA, full_bar(stage), c0, tap % 3, b * 66 + y0 + tap / 3);
            } else {
              tma_load_2d(a_dst, &tmA, full_bar(stage), k0, m0);
            }
            tma_load_2d(b_dst, &tmB, full_bar(stage), k0, n0);
          } else {
            const uint32_t lead_full = mapa_u32(full_bar(stage), 0);
            // The peer's bytes may land on the leader's barrier before this expect_tx; the phase still cannot complete
            // without the leader's arrive, and the peer cannot run a phase ahead (it waits for the MMA's commit).
            if (leader) mbar_arrive_expect_tx(full_bar(stage), 2 * C::STAGE_BYTES);
            if (g.a_conv3x3) {
              const int cpt = g.conv_c / C::KELEMS;
              const int tap = kb / cpt, c0 = (kb % cpt) * C::KELEMS;
              const int b = m0 >> 12, y0 = (m0 & 4095) >> 6;
              tma_load_3d_cg2(a_dst, &tmA, lead_full, c0, tap % 3, b * 66 + y0 + tap / 3);
            } else {
              tma_load_2d_cg2(a_dst, &tmA, lead_full, k0, m0);
            }
            tma_load_2d_cg2(b_dst, &tmB, lead_full, k0, n0);
          }
          if (++stage == C::STAGES) { stage = 0; phase ^= 1u; }
        }
      }
      // Teardown: the last STAGES tcgen05.commit arrivals on this CTA's empty barriers (for a CTA pair: multicast from the
      // leader's MMA thread, possibly in the other SM) have not been waited for by anyone.  Walk the ring once more without
      // loading: when this returns every commit-armed arrival has landed in this CTA's shared memory, so nothing can
      // arrive after the CTA has exited and its shared memory has been handed to the next CTA.
      for (int i = 0; i < C::STAGES; ++i) {
        mbar_wait(empty_bar(stage), phase ^ 1u);
        if (++stage == C::STAGES) { stage = 0; phase ^= 1u; }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer (pair leader only)
    if (leader && elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      int as = 0;
      uint32_t aphase = 0;
      const uint32_t idesc = (!TF32 && g.ab_f16) ? make_idesc(0u, C::BM * CG, BN) : C::IDESC;  // fp16 operands: formats 0
      for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
        mbar_wait_cluster(tempty_bar(as), aphase ^ 1u);  // epilogue has drained this accumulator
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(as * BN);
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(full_bar(stage), phase);
          tc_fence_after();
          const uint32_t a_addr = smem_base + stage * C::STAGE_BYTES;
          const uint64_t adesc = make_kmajor_sw128_desc(a_addr);
          const uint64_t bdesc = make_kmajor_sw128_desc(a_addr + C::A_BYTES);
#pragma unroll
          for (int k = 0; k < 4; ++k)  // 4 x 32 bytes of K per 128-byte swizzled row
            umma<CG, TF32>(d_tmem, adesc + 2u * k, bdesc + 2u * k, idesc, (uint32_t)((kb | k) != 0));
          umma_commit<CG>(empty_bar(stage));  // smem slot free once these MMAs retire
          if (++stage == C::STAGES) { stage = 0; phase ^= 1u; }
        }
        umma_commit<CG>(tfull_bar(as));  // accumulator complete -> epilogue
        if (++as == 2) { as = 0; aphase ^= 1u; }
      }
    }
    __syncwarp();
  } else if (warp >= 4 && warp < 4 + gemm_epi_warps(EPI)) {
    // ------------------------------------------------------------ epilogue (thread == output row)
    const int q = warp & 3;  // TMEM sub-partition this warp may read: lanes [32q, 32q+32)
    constexpr int CSPLIT = gemm_epi_warps(EPI) / 4;  // warps sharing a lane quarter
    const int half = (warp - 4) >> 2;                // which share of the columns is this warp's
    int as = 0;
    uint32_t aphase = 0;
    uint32_t sk = 0;  // EPI_BIAS_RESID_STATS_F32 / fp16 output: running count of this warp's TMA stores (staging buffer = sk & 1)
    // Windowed encoder blocks: q / k / v of a padded window token are the bare bias (the reference pads AFTER norm1 with zeros,
    // image_encoder.py:169-173,243-264).  Those 804 rows per image used to be written by a kernel of their own between this GEMM
    // and the attention (10 us x 28 blocks); the epilogue warps are idle until the first tile's accumulator is complete
    // (~5 us), so they write them here: 16-byte stores, consecutive threads on consecutive columns of a row.
    if constexpr (EPI == EPI_FOLD_BF16 || EPI == EPI_BIAS_BF16) {
      if (g.win_remap && g.pad_bias != nullptr) {
        const int groups = g.N >> 3;  // 16-byte column groups per row
        const long long items = (long long)g.pad_images * 804 * groups;
        for (long long it = (long long)blockIdx.x * 256 + ((int)threadIdx.x - 128); it < items; it += (long long)gridDim.x * 256) {
          const int cg = (int)(it % groups);
          const int rp = (int)(it / groups);
          const int p = rp % 804, b = rp / 804;
          // p: the 6 bottom rows of the padded 70x70 grid (420 tokens), then the 6 right columns of the other 64 rows
          const int y = p < 420 ? 64 + p / 70 : (p - 420) / 6, x = p < 420 ? p % 70 : 64 + (p - 420) % 6;
          const size_t row = (size_t)(b * 25 + (y / 14) * 5 + x / 14) * 196 + (y % 14) * 14 + x % 14;
          const float4 f0 = __ldg(reinterpret_cast<const float4*>(g.pad_bias + cg * 8));
          const float4 f1 = __ldg(reinterpret_cast<const float4*>(g.pad_bias + cg * 8 + 4));
          st_global_v4(reinterpret_cast<__nv_bfloat16*>(g.out) + row * g.ldo + cg * 8, pack_bf16x2(f0.x, f0.y), pack_bf16x2(f0.z, f0.w),
                       pack_bf16x2(f1.x, f1.y), pack_bf16x2(f1.z, f1.w));
        }
      }
    }
    for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
      const int m_blk = tile / tiles_n, n_blk = tile % tiles_n;
      const int m = m_blk * C::BM * CG + (int)rank * C::BM + q * 32 + lane;
      const int n0 = n_blk * BN;
      const bool row_ok = m < g.M;
      // qkv of a windowed block: token (b,y,x) is stored at its place in the 14x14-window partition of the 64x64
      // grid (SA/modeling/image_encoder.py:243-264), so the attention kernel reads whole windows with one TMA box
      int mo = m;
      if (g.win_remap) {
        const int b = m >> 12, y = (m >> 6) & 63, x = m & 63;
        mo = (b * 25 + (y / 14) * 5 + x / 14) * 196 + (y % 14) * 14 + x % 14;
      }
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * BN);
      constexpr int NCH = BN / 32;
      constexpr int NCW = NCH / CSPLIT;  // 32-column chunks handled by this warp
      const int c_first = half * NCW;
      constexpr bool kHasAdd = (epi_is_resid(EPI) || EPI == EPI_BIAS_ROWADD_F32);
      constexpr bool kSimple = (EPI == EPI_BIAS_BF16 || EPI == EPI_BIAS_GELU_BF16 || epi_is_fold(EPI) || epi_is_resid(EPI) ||
                                EPI == EPI_BIAS_ROWADD_F32 || EPI == EPI_BIAS_F32);
      // residual / row-add operand of this thread's row, fetched one chunk ahead of its use (the first chunk even
      // before the accumulator is complete).  In-place use (out == resid) is safe: a chunk is loaded before the
      // same thread stores it, and no other thread touches it.
      uint32_t radd[kHasAdd ? 2 : 1][32];
      const float* add_row = nullptr;
      constexpr bool kRowAdd = (EPI == EPI_BIAS_ROWADD_F32);
      [[maybe_unused]] const __half* add_row16 = nullptr;  // GemmArgs::rowadd_f16
      if constexpr (kHasAdd) {
        const int mm = row_ok ? m : 0;
        if constexpr (epi_is_resid(EPI)) add_row = g.resid + (size_t)mm * g.ldo + n0;
        else add_row = g.rowadd + (size_t)(mm % g.rowadd_mod) * g.N + n0;
        if constexpr (kRowAdd) add_row16 = reinterpret_cast<const __half*>(g.rowadd) + (size_t)(mm % g.rowadd_mod) * g.N + n0;
        if (n0 + c_first * 32 < g.N) {
#pragma unroll
          for (int h = 0; h < 4; ++h) {
            if (kRowAdd && g.rowadd_f16) {  // fp16 table: 64 bytes per 32-column chunk
              if (h < 2) ld_global_v8_stream(add_row16 + c_first * 32 + h * 16, *reinterpret_cast<uint32_t(*)[8]>(&radd[0][h * 8]));
            } else {
              ld_global_v8_stream(add_row + c_first * 32 + h * 8, *reinterpret_cast<uint32_t(*)[8]>(&radd[0][h * 8]));
            }
          }
        }
      }
      // folded LayerNorm (gemm_args.h): mean and rstd of this thread's row from the partial sums the producer of A left
      // in stats_in ([np][M][2]: lanes read consecutive rows), added in fixed order; fetched before the accumulator wait
      uint64_t rstd2 = 0, nrm2 = 0;
      if constexpr (epi_is_fold(EPI)) {
        const float2* sp = reinterpret_cast<const float2*>(g.stats_in) + (row_ok ? m : 0);
        float s = 0.f, ss = 0.f;
        for (int i = 0; i < g.stats_np; ++i) {
          const float2 t = sp[(size_t)i * g.M];
          s += t.x; ss += t.y;
        }
        const float inv_k = 1.0f / (float)g.K;
        const float mean = s * inv_k;
        const float rstd = rsqrtf(fmaxf(ss * inv_k - mean * mean, 0.f) + g.fold_eps);
        rstd2 = f2_pack(rstd, rstd);
        nrm2 = f2_pack(-rstd * mean, -rstd * mean);
      }
      // fp32 residual epilogue: the BN bias values of this tile go through shared memory.  As __ldg loads inside the chunk loop
      // they queued behind the streaming residual loads (32 sectors per instruction) and the add that consumes them was the
      // hottest stall of the short-K attn.proj GEMM (17 % of the epilogue warps' samples, ncu).  One value per epilogue thread,
      // fetched before the accumulator wait; two named barriers per tile keep the single 1 KB slice consistent.
      [[maybe_unused]] const uint32_t bias_smem = bar_base + 256;
      if constexpr (epi_is_fold(EPI)) {  // the same for the folded-LayerNorm epilogues: bias and column sums of the tile
        const int bi = (int)threadIdx.x - 128;
        float bv = 0.f, cv = 0.f;
        if (n0 + bi < g.N) { bv = __ldg(g.bias + n0 + bi); cv = __ldg(g.colsum + n0 + bi); }
        named_bar_sync(9, 256);
        asm volatile("st.shared.f32 [%0], %1;" ::"r"(econst + 4u * (uint32_t)bi), "f"(bv) : "memory");
        asm volatile("st.shared.f32 [%0], %1;" ::"r"(econst + 4u * (uint32_t)(256 + bi)), "f"(cv) : "memory");
        named_bar_sync(9, 256);
      }
      if constexpr (epi_is_resid(EPI)) {
        const int bi = (int)threadIdx.x - 128;  // 0..255: column of the tile
        float bv = 0.f;
        if (g.bias != nullptr && n0 + bi < g.N) bv = __ldg(g.bias + n0 + bi);
        named_bar_sync(9, 256);  // every epilogue warp is done with the previous tile's slice
        asm volatile("st.shared.f32 [%0], %1;" ::"r"(bias_smem + 4u * (uint32_t)bi), "f"(bv) : "memory");
        named_bar_sync(9, 256);
      }
      mbar_wait(tfull_bar(as), aphase);
      tc_fence_after();

      [[maybe_unused]] auto lds4c = [](uint32_t addr) {  // broadcast read of epilogue constants (volatile: not hoisted into registers)
        float4 r;
        asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "r"(addr));
        return r;
      };
      // 32 rows x 32 fp32 columns of this warp (x = this thread's row) -> out[row0 .. row0+31][col .. col+31] through the warp's
      // two 64B-swizzled staging halves and two TMA stores (rows >= M are clipped by the tensor map)
      [[maybe_unused]] auto stage_chunk = [&](const float* x, int col) {
        const uint32_t sbuf = staging_base + (uint32_t)(warp - 4) * 4096u;
        const uint32_t srow = (uint32_t)lane * 64u, sx = (uint32_t)((lane >> 1) & 3);
#pragma unroll
        for (int hf = 0; hf < 2; ++hf) {
          if (lane == 0) bulk_wait_group_read<1>();
          __syncwarp();
          const uint32_t sb = sbuf + (uint32_t)hf * 2048u;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int e = hf * 16 + 4 * j;
            st_shared_v4(sb + srow + (((uint32_t)j ^ sx) << 4), x[e], x[e + 1], x[e + 2], x[e + 3]);
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            tma_store_2d(&tmO, sb, col + hf * 16, m - lane);
            bulk_commit_group();
          }
        }
      };

      // the same for an fp16 output: the chunk's 32 rows x 64 bytes are the geometry of ONE staging half (as the bf16 copy of
      // EPI_BIAS_RESID_STATS_F32, tensor map tmX); the two halves of the warp's staging area rotate (sk)
      [[maybe_unused]] auto stage_chunk_f16 = [&](const float* x, int col) {
        uint32_t pk[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) pk[j] = pack_f16x2_sat(x[2 * j], x[2 * j + 1]);
        const uint32_t sbuf = staging_base + (uint32_t)(warp - 4) * 4096u;
        const uint32_t srow = (uint32_t)lane * 64u, sx = (uint32_t)((lane >> 1) & 3);
        if (lane == 0) bulk_wait_group_read<1>();
        __syncwarp();
        const uint32_t sb = sbuf + (sk++ & 1u) * 2048u;
#pragma unroll
        for (int j = 0; j < 4; ++j)
          st_shared_v4(sb + srow + (((uint32_t)j ^ sx) << 4), __uint_as_float(pk[4 * j]), __uint_as_float(pk[4 * j + 1]),
                       __uint_as_float(pk[4 * j + 2]), __uint_as_float(pk[4 * j + 3]));
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
          tma_store_2d(&tmX, sb, col, m - lane);
          bulk_commit_group();
        }
      };

      // bf16 output through the warp's single 2 KB staging buffer (ST == 3): 32 rows x 64 bytes, one TMA store per chunk
      [[maybe_unused]] auto stage_chunk_bf16 = [&](const float* x, int col) {
        uint32_t pk[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) pk[j] = pack_bf16x2(x[2 * j], x[2 * j + 1]);
        const uint32_t sb = staging_base + (uint32_t)(warp - 4) * 2048u;
        const uint32_t srow = (uint32_t)lane * 64u, sx = (uint32_t)((lane >> 1) & 3);
        if (lane == 0) bulk_wait_group_read<0>();  // the previous chunk's store has read the buffer
        __syncwarp();
#pragma unroll
        for (int j = 0; j < 4; ++j)
          st_shared_v4(sb + srow + (((uint32_t)j ^ sx) << 4), __uint_as_float(pk[4 * j]), __uint_as_float(pk[4 * j + 1]),
                       __uint_as_float(pk[4 * j + 2]), __uint_as_float(pk[4 * j + 3]));
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
          tma_store_2d(&tmX, sb, col, m - lane);
          bulk_commit_group();
        }
      };

      if constexpr (epi_is_resid(EPI)) {
        // out = resid + acc + bias, fp32.  Thread-per-row global stores of a 32 x 32 fp32 sub-tile touch 32 lines per
        // instruction; instead each warp writes 32 x 16 halves into 64B-swizzled staging buffers (conflict-free 16-byte
        // stores) and one lane sends them out with TMA stores.  Two buffers per warp: a buffer is rewritten once the
        // store issued from it two halves earlier has read it (bulk async-group wait).
        uint32_t v[32];
        float* f = reinterpret_cast<float*>(v);
        const uint32_t sbuf = staging_base + (uint32_t)(warp - 4) * 4096u;
        const int row0 = m - lane;
        const uint32_t srow = (uint32_t)lane * 64u, sx = (uint32_t)((lane >> 1) & 3);
        // folded LayerNorm (gemm_args.h): this thread's share of its row's sum and sum of squares, and a bf16 copy
        uint64_t rs2 = f2_pack(0.f, 0.f), rq2 = rs2;
        constexpr bool emit = (EPI == EPI_BIAS_RESID_STATS_F32);
#pragma unroll
        for (int ci = 0; ci < NCW; ++ci) {
          const int c = c_first + ci;
          const int col = n0 + c * 32;
          if (col < g.N) {  // warp-uniform
            tmem_ld_32x32(taddr + c * 32, v);
            if (ci + 1 < NCW && col + 32 < g.N) {
#pragma unroll
              for (int h = 0; h < 4; ++h)
                ld_global_v8_stream(add_row + (c + 1) * 32 + h * 8, *reinterpret_cast<uint32_t(*)[8]>(&radd[(ci + 1) & 1][h * 8]));
            }
            tmem_ld_wait();
            uint32_t pk[emit ? 16 : 1];  // bf16 copy of this thread's 32 columns of the chunk
#pragma unroll
            for (int hf = 0; hf < 2; ++hf) {
              if (lane == 0) bulk_wait_group_read<1>();
              __syncwarp();
              uint32_t sb = sbuf + (uint32_t)hf * 2048u;
              if constexpr (emit) sb = sbuf + (sk++ & 1u) * 2048u;  // three stores per chunk: rotate instead
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const int e = hf * 16 + 4 * j;          // column of the chunk
                // bias: broadcast 16-byte loads that stay in L1 (the residual stream bypasses it); the two adds run on
                // pairs (add.rn.f32x2): (acc + bias) + resid, the order of the unfused reference
                const float4 b4 = lds4c(bias_smem + 4u * (uint32_t)(c * 32 + e));  // the tile's bias slice (staged above)
                float x0, x1, x2, x3;
                f2_unpack(f2_add(f2_add(f2_pack(f[e], f[e + 1]), f2_pack(b4.x, b4.y)),
                                 f2_pack(__uint_as_float(radd[ci & 1][e]), __uint_as_float(radd[ci & 1][e + 1]))), x0, x1);
                f2_unpack(f2_add(f2_add(f2_pack(f[e + 2], f[e + 3]), f2_pack(b4.z, b4.w)),
                                 f2_pack(__uint_as_float(radd[ci & 1][e + 2]), __uint_as_float(radd[ci & 1][e + 3]))), x2, x3);
                st_shared_v4(sb + srow + (((uint32_t)j ^ sx) << 4), x0, x1, x2, x3);
                if constexpr (emit) {
                  const uint64_t a2 = f2_pack(x0, x1), b2 = f2_pack(x2, x3);
                  rs2 = f2_add(rs2, f2_add(a2, b2));
                  rq2 = f2_fma(a2, a2, f2_fma(b2, b2, rq2));
                  pk[hf * 8 + 2 * j] = pack_bf16x2(x0, x1);
                  pk[hf * 8 + 2 * j + 1] = pack_bf16x2(x2, x3);
                }
              }
              fence_proxy_async_smem();
              __syncwarp();
              if (lane == 0) {
                tma_store_2d(&tmO, sb, col + hf * 16, row0);
                bulk_commit_group();
              }
            }
            if constexpr (emit) {
              // the chunk's bf16 copy (32 rows x 64 bytes: the geometry of an fp32 half) leaves the same way; thread-per-
              // row 32-byte global stores cost the short-K attn.proj GEMM as much as the fp32 ones once did
              if (lane == 0) bulk_wait_group_read<1>();
              __syncwarp();
              const uint32_t sb = sbuf + (sk++ & 1u) * 2048u;
#pragma unroll
              for (int j = 0; j < 4; ++j)
                st_shared_v4(sb + srow + (((uint32_t)j ^ sx) << 4), __uint_as_float(pk[4 * j]), __uint_as_float(pk[4 * j + 1]),
                             __uint_as_float(pk[4 * j + 2]), __uint_as_float(pk[4 * j + 3]));
              fence_proxy_async_smem();
              __syncwarp();
              if (lane == 0) {
                tma_store_2d(&tmX, sb, col, row0);
                bulk_commit_group();
              }
            }
          }
        }
        if (emit && row_ok) {
          float s0, s1, q0, q1;
          f2_unpack(rs2, s0, s1);
          f2_unpack(rq2, q0, q1);
          reinterpret_cast<float2*>(g.stats_out)[(size_t)(n_blk * CSPLIT + half) * g.M + m] = make_float2(s0 + s1, q0 + q1);
        }
      } else if constexpr (kSimple) {
        uint32_t v[32];
        float* f = reinterpret_cast<float*>(v);
        // folded LayerNorm: out = rstd * acc + (-rstd * mean) * colsum + bias
        constexpr bool fold = epi_is_fold(EPI);
#pragma unroll
        for (int ci = 0; ci < NCW; ++ci) {
          const int c = c_first + ci;
          const int col = n0 + c * 32;
          if (col < g.N) {  // warp-uniform
            tmem_ld_32x32(taddr + c * 32, v);
            if constexpr (kHasAdd) {
              if (ci + 1 < NCW && col + 32 < g.N) {
#pragma unroll
                for (int h = 0; h < 4; ++h) {
                  if (kRowAdd && g.rowadd_f16) {
                    if (h < 2)
                      ld_global_v8_stream(add_row16 + (c + 1) * 32 + h * 16, *reinterpret_cast<uint32_t(*)[8]>(&radd[(ci + 1) & 1][h * 8]));
                  } else {
                    ld_global_v8_stream(add_row + (c + 1) * 32 + h * 8, *reinterpret_cast<uint32_t(*)[8]>(&radd[(ci + 1) & 1][h * 8]));
                  }
                }
              }
            }
            tmem_ld_wait();
            if constexpr (fold) {
#pragma unroll
              for (int j = 0; j < 32; j += 4) {
                const float4 b4 = lds4c(econst + 4u * (uint32_t)(c * 32 + j));          // the tile's bias / column-sum slices
                const float4 c4 = lds4c(econst + 4u * (uint32_t)(256 + c * 32 + j));    // (staged before the accumulator wait)
                f2_unpack(f2_fma(rstd2, f2_pack(f[j], f[j + 1]), f2_fma(nrm2, f2_pack(c4.x, c4.y), f2_pack(b4.x, b4.y))), f[j], f[j + 1]);
                f2_unpack(f2_fma(rstd2, f2_pack(f[j + 2], f[j + 3]), f2_fma(nrm2, f2_pack(c4.z, c4.w), f2_pack(b4.z, b4.w))), f[j + 2],
                          f[j + 3]);
              }
            } else if (g.bias != nullptr) {
#pragma unroll
              for (int j = 0; j < 32; j += 4) {
                const float4 b4 = __ldg(reinterpret_cast<const float4*>(g.bias + col + j));
                f2_unpack(f2_add(f2_pack(f[j], f[j + 1]), f2_pack(b4.x, b4.y)), f[j], f[j + 1]);
                f2_unpack(f2_add(f2_pack(f[j + 2], f[j + 3]), f2_pack(b4.z, b4.w)), f[j + 2], f[j + 3]);
              }
            }
            if constexpr (EPI == EPI_BIAS_GELU_BF16 || EPI == EPI_FOLD_GELU_BF16) {
#pragma unroll
              for (int j = 0; j < 32; j += 2) gelu_erf2(f[j], f[j + 1]);
            }
            if constexpr (EPI == EPI_BIAS_BF16 || EPI == EPI_BIAS_GELU_BF16 || fold) {
              if (!g.win_remap) {  // (window-order rows are not contiguous: thread-per-row stores below)
                stage_chunk_bf16(f, col);
              } else if (row_ok) {
                __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(g.out) + (size_t)mo * g.ldo + col;
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                  uint32_t pk[8];
#pragma unroll
                  for (int j = 0; j < 8; ++j) pk[j] = pack_bf16x2(f[h * 16 + 2 * j], f[h * 16 + 2 * j + 1]);
                  st_global_v8(o + h * 16, pk);
                }
              }
            } else {
              if constexpr (kHasAdd) {
#pragma unroll
                if (kRowAdd && g.rowadd_f16) {
#pragma unroll
                  for (int j = 0; j < 16; ++j) {
                    const float2 r2 = __half22float2(*reinterpret_cast<const __half2*>(&radd[ci & 1][j]));
                    f[2 * j] += r2.x; f[2 * j + 1] += r2.y;
                  }
                } else {
#pragma unroll
                  for (int j = 0; j < 32; ++j) f[j] += __uint_as_float(radd[ci & 1][j]);
                }
              }
              if (EPI == EPI_BIAS_ROWADD_F32 && g.out_f16) {
                stage_chunk_f16(f, col);
              } else {
                stage_chunk(f, col);
              }
            }
          }
        }
      } else if constexpr (EPI == EPI_RESID_LN_F32) {
        // out = LayerNorm_N(resid + acc + bias), N == BN: the two warps of a lane quarter take half a row each.
        // Pass 1 adds residual (prefetched a chunk ahead) and bias, keeps the sums in TMEM and accumulates sum and sum
        // of squares; the halves meet through shared memory (one named barrier per lane quarter); pass 2 normalises.
        uint32_t v[32];
        float* f = reinterpret_cast<float*>(v);
        uint32_t rr[2][32];
        int mm = row_ok ? (g.resid_mod ? m % g.resid_mod : m) : 0;
        if (g.row_img != nullptr && row_ok) mm += __ldg(g.row_img + m / g.resid_mod) * g.resid_mod;
        // the residual rows are fp32, or fp16 (GemmArgs::resid_f16: 64 bytes per 32-column chunk)
        const char* rrow = reinterpret_cast<const char*>(g.resid) + (size_t)mm * g.ldo * (g.resid_f16 ? 2 : 4);
        auto load_resid = [&](int b, int c) {
          if (g.resid_f16) {
#pragma unroll
            for (int h = 0; h < 2; ++h)
              ld_global_v8_stream(rrow + c * 64 + h * 32, *reinterpret_cast<uint32_t(*)[8]>(&rr[b][h * 8]));
          } else {
#pragma unroll
            for (int h = 0; h < 4; ++h)
              ld_global_v8_stream(rrow + c * 128 + h * 32, *reinterpret_cast<uint32_t(*)[8]>(&rr[b][h * 8]));
          }
        };
        load_resid(0, c_first);
        float sum = 0.f, sq = 0.f;
#pragma unroll
        for (int ci = 0; ci < NCW; ++ci) {
          const int c = c_first + ci;
          tmem_ld_32x32(taddr + c * 32, v);
          if (ci + 1 < NCW) load_resid((ci + 1) & 1, c + 1);
          tmem_ld_wait();
          if (g.resid_f16) {
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              const float4 b4 = lds4c(econst + 4 * (c * 32 + j));
              const float2 r0 = __half22float2(*reinterpret_cast<const __half2*>(&rr[ci & 1][j >> 1]));
              const float2 r1 = __half22float2(*reinterpret_cast<const __half2*>(&rr[ci & 1][(j >> 1) + 1]));
              f[j] += b4.x + r0.x; f[j + 1] += b4.y + r0.y; f[j + 2] += b4.z + r1.x; f[j + 3] += b4.w + r1.y;
            }
          } else {
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              const float4 b4 = lds4c(econst + 4 * (c * 32 + j));
              f[j] += b4.x + __uint_as_float(rr[ci & 1][j]);
              f[j + 1] += b4.y + __uint_as_float(rr[ci & 1][j + 1]);
              f[j + 2] += b4.z + __uint_as_float(rr[ci & 1][j + 2]);
              f[j + 3] += b4.w + __uint_as_float(rr[ci & 1][j + 3]);
            }
          }
          tmem_st_32x32(taddr + c * 32, v);
#pragma unroll
          for (int j = 0; j < 32; ++j) { sum += f[j]; sq += f[j] * f[j]; }
        }
        tmem_st_wait();
        float2* part = reinterpret_cast<float2*>(smem_raw + (bar_base - smem_u32(smem_raw)) + 256) + as * 256;
        part[half * 128 + q * 32 + lane] = make_float2(sum, sq);
        asm volatile("bar.sync %0, 64;" ::"r"(1 + q) : "memory");
        const float2 other = part[(half ^ 1) * 128 + q * 32 + lane];
        const float mean = (sum + other.x) * (1.0f / BN);
        const float var = fmaxf((sq + other.y) * (1.0f / BN) - mean * mean, 0.f);
        const float rstd = rsqrtf(var + g.ln_eps);
#pragma unroll
        for (int ci = 0; ci < NCW; ++ci) {
          const int c = c_first + ci;
          tmem_ld_32x32(taddr + c * 32, v);
          tmem_ld_wait();
#pragma unroll
          for (int h = 0; h < 8; ++h) {
            const float4 w0 = lds4c(econst + 4 * (256 + c * 32 + h * 4));
            const float4 b0 = lds4c(econst + 4 * (512 + c * 32 + h * 4));
            f[h * 4] = (f[h * 4] - mean) * rstd * w0.x + b0.x;
            f[h * 4 + 1] = (f[h * 4 + 1] - mean) * rstd * w0.y + b0.y;
            f[h * 4 + 2] = (f[h * 4 + 2] - mean) * rstd * w0.z + b0.z;
            f[h * 4 + 3] = (f[h * 4 + 3] - mean) * rstd * w0.w + b0.w;
          }
          if (g.out_f16) stage_chunk_f16(f, c * 32); else stage_chunk(f, c * 32);
        }
      } else if constexpr (EPI == EPI_LN_BF16_PAD || EPI == EPI_LN_F32_NCHW) {
        // LayerNorm over the N == BN columns of this thread's row; the row lives in TMEM, so the three
        // passes (mean, variance, normalise) re-read TMEM instead of holding 256 registers.
        uint32_t v[32];
        float* f = reinterpret_cast<float*>(v);
        float sum = 0.f;
#pragma unroll 1
        for (int c = 0; c < NCH; ++c) {
          tmem_ld_32x32(taddr + c * 32, v);
          tmem_ld_wait();
          if constexpr (EPI == EPI_RESID_LN_F32) {
            int mm = row_ok ? (g.resid_mod ? m % g.resid_mod : m) : 0;
            if (g.row_img != nullptr && row_ok) mm += __ldg(g.row_img + m / g.resid_mod) * g.resid_mod;
            const float* r = g.resid + (size_t)mm * g.ldo + c * 32;
#pragma unroll
            for (int h = 0; h < 4; ++h) {
              uint32_t pk[8];
              ld_global_v8(r + h * 8, pk);
#pragma unroll
              for (int j = 0; j < 8; ++j) f[h * 8 + j] += __uint_as_float(pk[j]) + __ldg(g.bias + c * 32 + h * 8 + j);
            }
            tmem_st_32x32(taddr + c * 32, v);  // keep resid + acc + bias in TMEM for the next passes
          }
#pragma unroll
          for (int j = 0; j < 32; ++j) sum += f[j];
        }
        if constexpr (EPI == EPI_RESID_LN_F32) tmem_st_wait();
        const float mean = sum * (1.0f / BN);
        float sq = 0.f;
#pragma unroll 1
        for (int c = 0; c < NCH; ++c) {
          tmem_ld_32x32(taddr + c * 32, v);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 32; ++j) { const float d = f[j] - mean; sq += d * d; }
        }
        const float rstd = rsqrtf(sq * (1.0f / BN) + g.ln_eps);
#pragma unroll 1
        for (int c = 0; c < NCH; ++c) {
          tmem_ld_32x32(taddr + c * 32, v);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 32; ++j)
            f[j] = (f[j] - mean) * rstd * __ldg(g.ln_w + c * 32 + j) + __ldg(g.ln_b + c * 32 + j);
          if (row_ok) {
            if constexpr (EPI == EPI_LN_BF16_PAD) {
              // token m = (b, y, x) -> row (b*66 + y+1)*66 + x+1 of the zero-bordered map
              const int b = m >> 12, y = (m >> 6) & 63, x = m & 63;
              __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(g.out) +
                                 ((size_t)(b * 66 + y + 1) * 66 + x + 1) * g.ldo + c * 32;
#pragma unroll
              for (int h = 0; h < 2; ++h) {
                uint32_t pk[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) pk[j] = pack_bf16x2(f[h * 16 + 2 * j], f[h * 16 + 2 * j + 1]);
                st_global_v8(o + h * 16, pk);
              }
            } else if constexpr (EPI == EPI_LN_F32_NCHW) {
              // out[b][n][tok]: consecutive lanes are consecutive tokens -> coalesced per channel
              const int b = m >> 12, tok = m & 4095;
              float* o = reinterpret_cast<float*>(g.out) + ((size_t)b * BN + c * 32) * 4096 + tok;
#pragma unroll
              for (int j = 0; j < 32; ++j) o[(size_t)j * 4096] = f[j];
            } else {
              float* o = reinterpret_cast<float*>(g.out) + (size_t)m * g.ldo + c * 32;
#pragma unroll
              for (int h = 0; h < 4; ++h) {
                uint32_t pk[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) pk[j] = __float_as_uint(f[h * 8 + j]);
                st_global_v8(o + h * 8, pk);
              }
            }
          }
        }
      } else if constexpr (EPI == EPI_UPSCALE1) {
        // Columns are (dy,dx,o): 4 sub-pixels x 64 channels. LayerNorm2d over o per sub-pixel, then GELU.
        // Row m = (box, y, x) of the 64x64 grid; output pixel (2y+dy, 2x+dx) of a [box][128][128][64] map.
        static_assert(EPI != EPI_UPSCALE1 || BN == 256, "upscale1 needs BN == 256");
        uint32_t v[2][32];
        const int bx = m >> 12, y = (m >> 6) & 63, x = m & 63;
#pragma unroll 1
        for (int spi = 0; spi < 2; ++spi) {
          const int sp = half * 2 + spi;  // the two warps of a lane quarter take the sub-pixel rows dy = 0 / 1
          float* f0 = reinterpret_cast<float*>(v[0]);
          float* f1 = reinterpret_cast<float*>(v[1]);
          tmem_ld_32x32(taddr + sp * 64, v[0]);
          tmem_ld_32x32(taddr + sp * 64 + 32, v[1]);
          tmem_ld_wait();
          auto lds4 = [](uint32_t addr) {
            float4 r;
            asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "r"(addr));
            return r;
          };
          float sum = 0.f;
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            const float4 b0 = lds4(econst + 4 * j), b1 = lds4(econst + 4 * (32 + j));
            f0[j] += b0.x; f0[j + 1] += b0.y; f0[j + 2] += b0.z; f0[j + 3] += b0.w;
            f1[j] += b1.x; f1[j + 1] += b1.y; f1[j + 2] += b1.z; f1[j + 3] += b1.w;
            sum += (f0[j] + f0[j + 1]) + (f0[j + 2] + f0[j + 3]) + (f1[j] + f1[j + 1]) + (f1[j + 2] + f1[j + 3]);
          }
          const float mean = sum * (1.0f / 64);
          float sq = 0.f;
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const float d0 = f0[j] - mean, d1 = f1[j] - mean;
            sq += d0 * d0 + d1 * d1;
          }
          const float rstd = rsqrtf(sq * (1.0f / 64) + g.ln_eps);
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            const float4 w0 = lds4(econst + 4 * (64 + j)), w1 = lds4(econst + 4 * (96 + j));
            const float4 c0 = lds4(econst + 4 * (128 + j)), c1 = lds4(econst + 4 * (160 + j));
            f0[j] = gelu_erf((f0[j] - mean) * rstd * w0.x + c0.x);
            f0[j + 1] = gelu_erf((f0[j + 1] - mean) * rstd * w0.y + c0.y);
            f0[j + 2] = gelu_erf((f0[j + 2] - mean) * rstd * w0.z + c0.z);
            f0[j + 3] = gelu_erf((f0[j + 3] - mean) * rstd * w0.w + c0.w);
            f1[j] = gelu_erf((f1[j] - mean) * rstd * w1.x + c1.x);
            f1[j + 1] = gelu_erf((f1[j + 1] - mean) * rstd * w1.y + c1.y);
            f1[j + 2] = gelu_erf((f1[j + 2] - mean) * rstd * w1.z + c1.z);
            f1[j + 3] = gelu_erf((f1[j + 3] - mean) * rstd * w1.w + c1.w);
          }
          if (row_ok && g.out_f16) {  // 64 channels = 128 contiguous bytes per sub-pixel
            const int Y = 2 * y + (sp >> 1), X = 2 * x + (sp & 1);
            uint32_t* o = reinterpret_cast<uint32_t*>(g.out) + (((size_t)bx * 128 + Y) * 128 + X) * 32;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              uint32_t pk[8];
#pragma unroll
              for (int j = 0; j < 8; ++j) pk[j] = pack_f16x2_sat(f0[h * 16 + 2 * j], f0[h * 16 + 2 * j + 1]);
              st_global_v8(o + h * 8, pk);
#pragma unroll
              for (int j = 0; j < 8; ++j) pk[j] = pack_f16x2_sat(f1[h * 16 + 2 * j], f1[h * 16 + 2 * j + 1]);
              st_global_v8(o + 16 + h * 8, pk);
            }
          } else if (row_ok) {
            const int Y = 2 * y + (sp >> 1), X = 2 * x + (sp & 1);
            float* o = reinterpret_cast<float*>(g.out) + (((size_t)bx * 128 + Y) * 128 + X) * 64;
#pragma unroll
            for (int h = 0; h < 4; ++h) {
              uint32_t pk[8];
#pragma unroll
              for (int j = 0; j < 8; ++j) pk[j] = v[0][h * 8 + j];
              st_global_v8(o + h * 8, pk);
#pragma unroll
              for (int j = 0; j < 8; ++j) pk[j] = v[1][h * 8 + j];
              st_global_v8(o + 32 + h * 8, pk);
            }
          }
        }
      } else if constexpr (EPI == EPI_UPSCALE2_MASK) {
        // Columns are (dy,dx,o): 4 sub-pixels x 32 channels; GELU then dot with the box's hypernetwork
        // vector (mask token 0) -> one low-res logit per sub-pixel. Row m = (box, Y, X) of the 128x128 map.
        static_assert(EPI != EPI_UPSCALE2_MASK || BN == 128, "upscale2 needs BN == 128");
        uint32_t v[32];
        float* f = reinterpret_cast<float*>(v);
        const int mm = row_ok ? m : 0;
        const int bx = mm >> 14, Y = (mm >> 7) & 127, X = mm & 127;
        const float* hy = g.hyper + (size_t)bx * 32;
        float logit[2];
#pragma unroll
        for (int spi = 0; spi < 2; ++spi) {
          const int sp = half * 2 + spi;  // this warp's sub-pixel row dy = half
          tmem_ld_32x32(taddr + sp * 32, v);
          tmem_ld_wait();
          float acc = 0.f;
#pragma unroll
          for (int j = 0; j < 32; j += 4) {  // GELU on pairs (gelu_erf2: 7 instead of 12 FMA-pipe instructions per element)
            const float4 b4 = __ldg(reinterpret_cast<const float4*>(g.bias + j));
            const float4 h4 = __ldg(reinterpret_cast<const float4*>(hy + j));
            float x0 = f[j] + b4.x, x1 = f[j + 1] + b4.y, x2 = f[j + 2] + b4.z, x3 = f[j + 3] + b4.w;
            gelu_erf2(x0, x1);
            gelu_erf2(x2, x3);
            acc += x0 * h4.x; acc += x1 * h4.y; acc += x2 * h4.z; acc += x3 * h4.w;
          }
          logit[spi] = acc;
        }
        if (row_ok) {
          float* o = reinterpret_cast<float*>(g.out) + ((size_t)bx * 256 + 2 * Y + half) * 256 + 2 * X;
          *reinterpret_cast<float2*>(o) = make_float2(logit[0], logit[1]);
        }
      }

      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if constexpr (CG == 1) mbar_arrive(tempty_bar(as));
        else mbar_arrive_cluster(mapa_u32(tempty_bar(as), 0));
      }
      if (++as == 2) { as = 0; aphase ^= 1u; }
    }
  }

  // ---------------------------------------------------------------- teardown
  if constexpr (gemm_staged(EPI) != 0) {
    if (warp >= 4 && lane == 0) bulk_wait_group_all();  // this warp's TMA stores have left shared memory and are complete
  }
  tc_fence_before();
  if constexpr (CG == 2) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  if (warp == 2) tmem_dealloc<CG>(tmem_base, C::TMEM_COLS);
}

}  // namespace inksam
