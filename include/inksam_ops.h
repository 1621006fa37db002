/* libinksam -- kernel-level entry points (C ABI).
 *
 * These expose each sm_100a kernel of the SAM hot path on its own so that tests/ can compare every
 * kernel with the CPU oracle.  They are NOT the drop-in boundary (that is include/inksam.h); the
 * product path reaches the same kernels through inksam_encode / inksam_decode_boxes.
 *
 * Conventions: every pointer is a device pointer owned by the caller; `stream` is a cudaStream_t
 * passed as void*; calls are asynchronous on that stream; the return value is 0 on success, non-zero
 * otherwise with the message available from inksam_last_error().
 */
#ifndef INKSAM_OPS_H_
#define INKSAM_OPS_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Epilogue selectors of the tcgen05 GEMM (mirror inksam::GemmEpi). */
enum {
  INKSAM_EPI_BIAS_BF16 = 0,
  INKSAM_EPI_BIAS_GELU_BF16 = 1,
  INKSAM_EPI_BIAS_RESID_F32 = 2,
  INKSAM_EPI_BIAS_ROWADD_F32 = 3,
  INKSAM_EPI_LN_BF16_PAD = 4,
  INKSAM_EPI_LN_F32_NCHW = 5,
  INKSAM_EPI_RESID_LN_F32 = 6,
  INKSAM_EPI_BIAS_F32 = 7,
  INKSAM_EPI_UPSCALE1 = 8,
  INKSAM_EPI_UPSCALE2_MASK = 9
};

/* out[M,N] = epilogue(A[M,K] * W[N,K]^T): replaces every nn.Linear on the path, e.g.
 * SA/modeling/image_encoder.py:227,238 and SA/modeling/common.py:21-26.
 * tf32 == 0: A and W are bf16; tf32 == 1: A and W are fp32 consumed as tf32.
 * cg: 1 = one CTA per tile (UMMA M=128), 2 = CTA pair (UMMA M=256). bn: tile width (256 or 128). */
int inksam_op_gemm(int cg, int tf32, int bn, int epi, const void* A, int lda, const void* W, int ldw, int M, int N,
                   int K, void* out, int ldo, const float* bias, const float* resid, const float* rowadd,
                   int rowadd_mod, const float* ln_w, const float* ln_b, float ln_eps, const float* hyper,
                   int a_conv3x3, int conv_c, int win_remap, int resid_mod, void* stream);

/* qkv projection of a windowed encoder block as the engine launches it (CTA pairs, bias + bf16 epilogue with win_remap): A bf16
 * [B*4096][lda] in natural token order, W bf16 [N][ldw]; out bf16 [B*25*196][N] in 14x14-window order.  pad_bias != null: the
 * rows of the 804 padded window tokens per image are filled with bf16(pad_bias[N]) by the GEMM's epilogue warps (what
 * inksam_op_fill_pad_rows does as a kernel of its own); null: they are left untouched. */
int inksam_op_gemm_qkv_windowed(const void* A, int lda, const void* W, int ldw, int B, int N, int K, void* out, const float* bias,
                                const float* pad_bias, void* stream);

/* Encoder attention for B images: softmax(q k^T * hd^-0.5 + decomposed rel-pos bias) v, all heads
 * (SA/modeling/image_encoder.py:224-240,325-361).  windowed == 0: qkv is bf16 [B*4096][3*D] in token order and rel is
 * the bf16 table [256][hd] (rows 0..126 rel_pos_h, 128..254 rel_pos_w).  windowed == 1: qkv is bf16 [B*25*196][3*D] in
 * 14x14-window order with the qkv bias in the rows of padded tokens, rel is [64][hd] (rows 0..26 / 32..58).
 * out: bf16 [B*4096][D] in token order. */
int inksam_op_attention(int windowed, const void* qkv, const void* rel, void* out, int B, int heads, int head_dim, void* stream);

/* Debug / profiling aid: buf = device array of [items][16] int64 that CTA 0 of the windowed attention kernel fills with
 * the clock64() of its pipeline events (MMA issue, barrier hand-offs) on the following launches; NULL switches it off. */
int inksam_op_attention_trace(long long* buf);

/* nn.LayerNorm over the last dim (fp32 in, bf16 out); plain != 0 only casts. image_encoder.py:168,180. */
int inksam_op_layernorm(const float* x, const float* w, const float* b, void* out, int M, int D, float eps, int plain, void* stream);

/* LayerNorm folded into the GEMMs either side of it (the encoder blocks' norm1 -> qkv and norm2 -> lin1,
 * image_encoder.py:180-197):   LN(x) W^T + b == rstd * (x (gamma o W)^T) - rstd * mean * colsum(gamma o W) + (b + W beta).
 *   inksam_op_fold_ln_weights: W fp32 [N][K] -> w_out bf16 [N][K] = gamma o W, colsum [N] (of the rounded values),
 *                              bias_out [N] = bias + W beta.
 *   inksam_op_row_stats_bf16:  x fp32 [M][D] -> out bf16 [M][D] (plain cast) and stats [M][2] = (sum, sum of squares).
 *   inksam_op_gemm_fold:       CTA-pair bf16 GEMM, BN 256.  epi 2 (bias + fp32 residual): also writes stats_out
 *                              [inksam_op_gemm_stats_partials(N)][M][2] and the bf16 copy xb_out [M][ldo] of its output
 *                              (both, or both null for the plain epilogue).  epi 0 / 1 (bf16 out, optional GELU): A is the raw bf16 x, W / colsum /
 *                              bias come from inksam_op_fold_ln_weights, stats_in [stats_np][M][2] are partials of the
 *                              rows of A over its K columns. */
int inksam_op_fold_ln_weights(const float* W, const float* gamma, const float* beta, const float* bias, void* w_out,
                              float* colsum, float* bias_out, int N, int K, void* stream);
int inksam_op_row_stats_bf16(const float* x, void* out, float* stats, int M, int D, void* stream);
int inksam_op_gemm_stats_partials(int N);
int inksam_op_gemm_fold(int epi, const void* A, int lda, const void* W, int ldw, int M, int N, int K, void* out, int ldo,
                        const float* bias, const float* resid, float* stats_out, void* xb_out, const float* stats_in,
                        const float* colsum, int stats_np, float eps, int win_remap, void* stream);

/* Sam.preprocess + PatchEmbed im2col: out bf16 [B*4096][768]. hw is a HOST array. modeling/sam.py:164-174. */
int inksam_op_patchify(const void* images, int dtype, int B, const int* hw, long long img_stride, long long row_pitch,
                       long long plane_stride, const float* mean, const float* stdv, void* out, void* stream);

/* qkv bias into the padded-token rows of a window-ordered qkv matrix. image_encoder.py:258-259. */
int inksam_op_fill_pad_rows(void* qkv, const float* bias, int B, int cols, void* stream);

/* Sam.postprocess_masks + threshold, fused. modeling/sam.py:133-162, predictor.py:238-241. */
int inksam_op_postprocess(const float* low, unsigned char* masks, float* logits, int K, int in_h, int in_w, int H, int W,
                          float thr, void* stream);
/* The same with bit-packed masks: mask_bits is uint8 [K][H][(W+7)/8], leftmost pixel in the MSB, rows padded to whole
 * bytes with zero bits -- numpy.packbits(axis=-1) / PIL mode "1" rows, the format InkLayer/runner.py:51 converts to. */
int inksam_op_postprocess_packed(const float* low, unsigned char* mask_bits, int K, int in_h, int in_w, int H, int W, float thr,
                                 void* stream);

/* The two cross attentions of the mask decoder's TwoWayAttentionBlock on the image side (SA/modeling/transformer.py:151-182,
 * Attention with downsample_rate 2: 8 heads x 16, :185-240), after the q / k / v projections, before out_proj.
 *   inksam_op_decoder_t2i: token -> image.  q [boxes][7][128]; kv rows [4096][ld] fp32 per box, k in columns [0,128) and v in
 *     [128,256), box b at kv + (box_img ? box_img[b] : b) * box_stride elements (box_stride 0: all boxes share the keys);
 *     out [boxes][7][128] = softmax(q k^T / 4) v per head.  scratch: inksam_op_decoder_t2i_scratch_bytes(boxes) device bytes.
 *   inksam_op_decoder_i2t: image -> token.  q = columns [qcol, qcol+128) of the rows [4096][ld] at qsrc (+ box offset as
 *     above); kv [boxes][2][7][128] (k then v of the 7 tokens); out [boxes*4096][128]. */
size_t inksam_op_decoder_t2i_scratch_bytes(int boxes);
int inksam_op_decoder_t2i(const float* q, const void* kv, int kv_half, long long box_stride, int ld, const int* box_img, float* out,
                          void* scratch, int boxes, void* stream);
int inksam_op_decoder_i2t(const void* qsrc, int q_half, long long box_stride, int ld, int qcol, const int* box_img, const float* kv,
                          void* out, int out_half, int boxes, void* stream);
/* kv_half / q_half != 0: the k|v|q rows are fp16 (as inksam_op_gemm_rowadd_f16 writes them) instead of fp32; box_stride, ld and
 * qcol count elements either way.  The engine stores the per-box k|v|q of the decoder that way: they are written once by a tf32
 * GEMM and read once by these kernels' fp32 FMAs, and fp16 keeps the 10 explicit mantissa bits of a tf32 operand.
 * inksam_op_gemm_rowadd_f16: the tcgen05 GEMM with the row-add epilogue (inksam_op_gemm, epi 3) storing
 * out_f16[m][n] = fp16(acc + bias[n] + rowadd[m % rowadd_mod][n]), round to nearest, saturating at +-65504. */
int inksam_op_gemm_rowadd_f16(int cg, int tf32, int bn, const void* A, int lda, const void* W, int ldw, int M, int N, int K,
                              void* out_f16, int ldo, const float* bias, const float* rowadd, int rowadd_mod, void* stream);
/* out_half != 0 (inksam_op_decoder_i2t): the attention output is stored as fp16 [boxes*4096][128].
 * inksam_op_gemm_f16: inksam_op_gemm with fp16 operands A [M][lda] and W [N][ldw] (kind::f16 MMAs, fp32 accumulation) for the
 * epilogues the decoder's fp16 intermediates go through -- epi 3 (row-add), 6 (residual + LayerNorm, N == bn == 256),
 * 8 (upscale 1, bn 256), 9 (upscale 2 + mask, bn 128); out_f16 != 0 stores the output as fp16 (epi 3, 6, 8); resid_f16 != 0:
 * the residual rows of epi 6 are fp16; rowadd_f16 != 0: the row-add table of epi 3 is fp16. */
int inksam_op_gemm_f16(int cg, int bn, int epi, const void* A, int lda, const void* W, int ldw, int M, int N, int K, void* out, int ldo,
                       int out_f16, const float* bias, const void* resid, const void* rowadd, int rowadd_mod, const float* ln_w,
                       const float* ln_b, float ln_eps, const float* hyper, int resid_mod, int resid_f16, int rowadd_f16, void* stream);

/* ResizeLongestSide.apply_image on the device: Pillow's 8-bit bilinear resample (torchvision resize of a PIL image,
 * SA/utils/transforms.py:26-31; algorithm of Pillow libImaging/Resample.c), bit-exact with the host path.
 * src: uint8 [H][W][3] with src_row_pitch bytes per row; dst: uint8 [out_h][out_w][3] with dst_row_pitch bytes per row
 * (e.g. a 1024x1024x3 container, pitch 3072); workspace: inksam_resize_workspace_bytes(H, W, out_h, out_w) device bytes
 * owned by the caller, in use until the call's work on `stream` has finished. */
size_t inksam_resize_workspace_bytes(int H, int W, int out_h, int out_w);
int inksam_resize_u8(const uint8_t* src, int H, int W, long long src_row_pitch, uint8_t* dst, int out_h, int out_w,
                     long long dst_row_pitch, void* workspace, void* stream);

/* inksam_resize_u8 for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * Same arguments and result as inksam_resize_u8 minus the workspace: the coefficient tables and the intermediate image
 * live in a file-scope `static __device__` array of inksam_resize_static_max_bytes() = 16 MiB, which every GPU that
 * loads the library holds until the process exits.  A resize whose inksam_resize_workspace_bytes(H, W, out_h, out_w)
 * exceeds that (e.g. H * out_w * 3 > ~16 MB) is rejected with a non-zero return and inksam_last_error() set.  The call
 * allocates and frees nothing.  All calls on a device share the array (it keeps the previous call's contents; nothing
 * in it needs to start at zero): two calls that overlap on the GPU (different streams, no ordering) race. */
size_t inksam_resize_static_max_bytes(void);
int inksam_resize_u8_static(const uint8_t* src, int H, int W, long long src_row_pitch, uint8_t* dst, int out_h, int out_w,
                            long long dst_row_pitch, void* stream);

const char* inksam_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* INKSAM_OPS_H_ */
