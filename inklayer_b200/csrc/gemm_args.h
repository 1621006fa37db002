// Epilogue selectors and argument block of the tcgen05 GEMM (host-visible, no device code).
#pragma once
#include <stdint.h>

namespace inksam {

enum GemmEpi : int {
  EPI_BIAS_BF16 = 0,       // out_bf16 = acc + bias
  EPI_BIAS_GELU_BF16 = 1,  // out_bf16 = gelu_erf(acc + bias)
  EPI_BIAS_RESID_F32 = 2,  // out_f32 = acc + bias + resid                  (resid may alias out)
  EPI_BIAS_ROWADD_F32 = 3, // out_f32 = acc + bias + rowadd[m % rowadd_mod] (pos_embed / projected PE)
  EPI_LN_BF16_PAD = 4,     // out_bf16[pad66(m)] = LN_N(acc)       N == BN  (neck conv1x1 + LayerNorm2d)
  EPI_LN_F32_NCHW = 5,     // out_f32[b][n][tok] = LN_N(acc)       N == BN  (neck conv3x3 + LayerNorm2d)
  EPI_RESID_LN_F32 = 6,    // out_f32 = LN_N(resid + acc + bias)   N == BN  (decoder keys update, norm4)
  EPI_BIAS_F32 = 7,        // out_f32 = acc + bias
  EPI_UPSCALE1 = 8,        // decoder ConvT(256->64,k2,s2)+LayerNorm2d+GELU, out [K][128][128][64] fp32
  EPI_UPSCALE2_MASK = 9,   // decoder ConvT(64->32,k2,s2)+GELU, dot with hyper_in -> low-res logits
  // LayerNorm folded into the GEMMs either side of it (see stats_* below); separate kernels, so that the plain
  // epilogues above carry none of this
  EPI_FOLD_BF16 = 10,      // out_bf16 = rstd * acc - rstd * mean * colsum + bias          (A = raw bf16 x)
  EPI_FOLD_GELU_BF16 = 11, // out_bf16 = gelu_erf(the same)
  EPI_BIAS_RESID_STATS_F32 = 12,  // EPI_BIAS_RESID_F32 + per-row partial (sum, sum of squares) + bf16 copy of the output
};

struct GemmArgs {
  int M, N, K;
  int ldo;  // output row stride in elements
  void* out;
  const float* bias;   // [N] or nullptr
  const float* resid;  // fp32 [M][ldo]
  const float* rowadd; // fp32 [rowadd_mod][N]
  int rowadd_mod;
  const float* ln_w;
  const float* ln_b;
  float ln_eps;
  int a_conv3x3;       // A is the zero-padded [B][66][66][C] map; K = 9*C (3-D tensor map)
  int conv_c;
  const float* hyper;  // EPI_UPSCALE2_MASK: [boxes][32]
  int win_remap;       // EPI_BIAS_BF16: store token rows in 14x14-window order (windowed encoder blocks)
  const float* pad_bias;  // with win_remap: fp32 [N]; the epilogue warps also fill the rows of the padded window tokens (804 per image,
  int pad_images;         // encoder_ops.cu fill_pad_rows_kernel) with bf16(pad_bias) while the first tile's MMAs run.  Null: not filled.
  int resid_mod;       // EPI_RESID_LN_F32: resid row = m % resid_mod when non-zero (keys shared by all boxes)
  const int* row_img;  // EPI_RESID_LN_F32 with resid_mod: resid row += row_img[m / resid_mod] * resid_mod (keys of the box's image)
  // LayerNorm folded into the GEMMs either side of it (encoder blocks).  The producer of the residual
  // stream also emits what the LayerNorm would have read; the consumer applies the normalisation to its accumulator:
  //   LN(x) W^T + b  ==  rstd * (x (gamma o W)^T) - rstd * mean * colsum(gamma o W) + (b + W beta)
  float* stats_out;       // EPI_BIAS_RESID_STATS_F32: [stats_np][M][2] partial (sum, sum of squares) of each output row, one per
                          // (n tile, column half) = gemm_stats_partials(N) of them
  void* xb_out;           // EPI_BIAS_RESID_STATS_F32: bf16 copy of the output rows [M][ldo] (A of the next GEMM)
  const float* stats_in;  // EPI_FOLD_*: partials of the rows of A over its K columns
  const float* colsum;    // [N] sum over k of the bf16 weight row (with stats_in; `bias` then holds b + W beta)
  int stats_np;           // partials per row in stats_in
  float fold_eps;         // eps of the folded LayerNorm
  int ab_f16;             // 2-byte operands (tf32 == 0) are fp16 instead of bf16 (the decoder's GEMM-to-GEMM intermediates; same 10
                          // explicit mantissa bits as the tf32 operands they replace, half the bytes)
  int resid_f16;          // EPI_RESID_LN_F32: the residual rows are fp16 [..][ldo]
  int rowadd_f16;         // EPI_BIAS_ROWADD_F32: the row-add table is fp16 [rowadd_mod][N]
  int out_f16;            // EPI_BIAS_ROWADD_F32 / EPI_RESID_LN_F32 / EPI_UPSCALE1: store the output as fp16 (round to nearest,
                          // saturating) instead of fp32; same shape, ldo in elements.
                          // The decoder's per-box k|v|q are consumed once, by fp32 FMAs, after a tf32 GEMM produced them: fp16 keeps the
                          // same 10 explicit mantissa bits and halves their HBM traffic.
};

// Partials per row that an EPI_BIAS_RESID_STATS_F32 launch with N output columns writes to stats_out (BN = 256, two column
// halves per tile).
inline int gemm_stats_partials(int N) { return 2 * ((N + 255) / 256); }

}  // namespace inksam
