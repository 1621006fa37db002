// C-ABI kernel-level entry points (include/inksam_ops.h).
#include "common.h"
#include "inksam_ops.h"
#include "decoder_ops.h"

using namespace inksam;

static int device_sms() {
  static PerDeviceInt sms_dev;
  int& sms = sms_dev();
  if (sms == 0) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, current_device());
  return sms;
}

extern "C" {

const char* inksam_last_error(void) { return get_error(); }

int inksam_op_gemm(int cg, int tf32, int bn, int epi, const void* A, int lda, const void* W, int ldw, int M, int N,
                   int K, void* out, int ldo, const float* bias, const float* resid, const float* rowadd,
                   int rowadd_mod, const float* ln_w, const float* ln_b, float ln_eps, const float* hyper,
                   int a_conv3x3, int conv_c, int win_remap, int resid_mod, void* stream) {
  GemmArgs g{};
  g.M = M; g.N = N; g.K = K; g.ldo = ldo; g.out = out;
  g.bias = bias; g.resid = resid; g.rowadd = rowadd; g.rowadd_mod = rowadd_mod;
  g.ln_w = ln_w; g.ln_b = ln_b; g.ln_eps = ln_eps; g.hyper = hyper;
  g.a_conv3x3 = a_conv3x3; g.conv_c = conv_c; g.win_remap = win_remap; g.resid_mod = resid_mod;
  return launch_gemm(cg, tf32, bn, epi, A, lda, W, ldw, g, device_sms(), static_cast<cudaStream_t>(stream));
}

int inksam_op_gemm_rowadd_f16(int cg, int tf32, int bn, const void* A, int lda, const void* W, int ldw, int M, int N, int K,
                              void* out_f16, int ldo, const float* bias, const float* rowadd, int rowadd_mod, void* stream) {
  GemmArgs g{};
  g.M = M; g.N = N; g.K = K; g.ldo = ldo; g.out = out_f16; g.bias = bias; g.rowadd = rowadd; g.rowadd_mod = rowadd_mod; g.out_f16 = 1;
  return launch_gemm(cg, tf32, bn, EPI_BIAS_ROWADD_F32, A, lda, W, ldw, g, device_sms(), static_cast<cudaStream_t>(stream));
}

int inksam_op_gemm_f16(int cg, int bn, int epi, const void* A, int lda, const void* W, int ldw, int M, int N, int K, void* out, int ldo,
                       int out_f16, const float* bias, const void* resid, const void* rowadd, int rowadd_mod, const float* ln_w,
                       const float* ln_b, float ln_eps, const float* hyper, int resid_mod, int resid_f16, int rowadd_f16, void* stream) {
  GemmArgs g{};
  g.M = M; g.N = N; g.K = K; g.ldo = ldo; g.out = out; g.bias = bias; g.resid = static_cast<const float*>(resid);
  g.rowadd = static_cast<const float*>(rowadd); g.rowadd_f16 = rowadd_f16;
  g.rowadd_mod = rowadd_mod; g.ln_w = ln_w; g.ln_b = ln_b; g.ln_eps = ln_eps; g.hyper = hyper; g.resid_mod = resid_mod;
  g.ab_f16 = 1; g.out_f16 = out_f16; g.resid_f16 = resid_f16;
  return launch_gemm(cg, 0, bn, epi, A, lda, W, ldw, g, device_sms(), static_cast<cudaStream_t>(stream));
}

int inksam_op_gemm_qkv_windowed(const void* A, int lda, const void* W, int ldw, int B, int N, int K, void* out, const float* bias,
                                const float* pad_bias, void* stream) {
  GemmArgs g{};
  g.M = B * 4096; g.N = N; g.K = K; g.ldo = N; g.out = out; g.bias = bias; g.win_remap = 1; g.pad_bias = pad_bias; g.pad_images = B;
  return launch_gemm(2, 0, 256, EPI_BIAS_BF16, A, lda, W, ldw, g, device_sms(), static_cast<cudaStream_t>(stream));
}

int inksam_op_attention(int windowed, const void* qkv, const void* rel, void* out, int B, int heads, int head_dim, void* stream) {
  return launch_attention(windowed, qkv, rel, out, B, heads, head_dim, static_cast<cudaStream_t>(stream));
}

int inksam_op_attention_trace(long long* buf) {
  set_attention_trace(buf);
  return 0;
}

int inksam_op_layernorm(const float* x, const float* w, const float* b, void* out, int M, int D, float eps, int plain, void* stream) {
  return launch_layernorm_bf16(x, w, b, out, M, D, eps, plain, nullptr, static_cast<cudaStream_t>(stream));
}

int inksam_op_row_stats_bf16(const float* x, void* out, float* stats, int M, int D, void* stream) {
  if (!stats) return set_error("inksam_op_row_stats_bf16: null stats");
  return launch_layernorm_bf16(x, nullptr, nullptr, out, M, D, 0.f, 1, stats, static_cast<cudaStream_t>(stream));
}

int inksam_op_fold_ln_weights(const float* W, const float* gamma, const float* beta, const float* bias, void* w_out,
                              float* colsum, float* bias_out, int N, int K, void* stream) {
  if (!W || !gamma || !beta || !bias || !w_out || !colsum || !bias_out) return set_error("inksam_op_fold_ln_weights: null argument");
  return launch_fold_ln_weights(W, gamma, beta, bias, w_out, colsum, bias_out, N, K, static_cast<cudaStream_t>(stream));
}

int inksam_op_gemm_stats_partials(int N) { return gemm_stats_partials(N); }

int inksam_op_gemm_fold(int epi, const void* A, int lda, const void* W, int ldw, int M, int N, int K, void* out, int ldo,
                        const float* bias, const float* resid, float* stats_out, void* xb_out, const float* stats_in,
                        const float* colsum, int stats_np, float eps, int win_remap, void* stream) {
  if (epi != EPI_BIAS_BF16 && epi != EPI_BIAS_GELU_BF16 && epi != EPI_BIAS_RESID_F32)
    return set_error("inksam_op_gemm_fold: epilogue %d has no folded-LayerNorm form", epi);
  if (epi == EPI_BIAS_RESID_F32 ? (stats_in != nullptr) : (stats_out != nullptr || xb_out != nullptr))
    return set_error("inksam_op_gemm_fold: epilogue %d given the other side's arguments", epi);
  if (stats_in != nullptr && (!colsum || !bias || stats_np < 1)) return set_error("inksam_op_gemm_fold: stats_in needs colsum, bias and stats_np");
  if ((stats_out != nullptr) != (xb_out != nullptr)) return set_error("inksam_op_gemm_fold: stats_out and xb_out come together");
  if (epi == EPI_BIAS_RESID_F32 && stats_out != nullptr) epi = EPI_BIAS_RESID_STATS_F32;
  if (epi == EPI_BIAS_BF16 && stats_in != nullptr) epi = EPI_FOLD_BF16;
  if (epi == EPI_BIAS_GELU_BF16 && stats_in != nullptr) epi = EPI_FOLD_GELU_BF16;
  GemmArgs g{};
  g.M = M; g.N = N; g.K = K; g.ldo = ldo; g.out = out; g.bias = bias; g.resid = resid; g.win_remap = win_remap;
  g.stats_out = stats_out; g.xb_out = xb_out; g.stats_in = stats_in; g.colsum = colsum; g.stats_np = stats_np; g.fold_eps = eps;
  return launch_gemm(2, 0, 256, epi, A, lda, W, ldw, g, device_sms(), static_cast<cudaStream_t>(stream));
}

size_t inksam_op_decoder_t2i_scratch_bytes(int boxes) {
  return boxes > 0 ? decoder_t2i_scratch_floats(boxes) * sizeof(float) + (size_t)boxes * sizeof(int) : 0;
}

int inksam_op_decoder_t2i(const float* q, const void* kv, int kv_half, long long box_stride, int ld, const int* box_img, float* out,
                          void* scratch, int boxes, void* stream) {
  if (!q || !kv || !out || !scratch || boxes < 1) return set_error("inksam_op_decoder_t2i: null argument or no boxes");
  if (ld < 256 || ld % 4 != 0 || (reinterpret_cast<uintptr_t>(kv) & 15) != 0 || box_stride % 4 != 0)
    return set_error("inksam_op_decoder_t2i: kv rows must be 16-byte aligned with k at column 0 and v at column 128 (ld=%d)", ld);
  auto st = static_cast<cudaStream_t>(stream);
  float* partial = static_cast<float*>(scratch);
  int* ticket = reinterpret_cast<int*>(partial + decoder_t2i_scratch_floats(boxes));
  CUDA_TRY(cudaMemsetAsync(ticket, 0, (size_t)boxes * sizeof(int), st));
  return launch_decoder_t2i(q, kv, kv_half, box_stride, ld, box_img, out, partial, ticket, boxes, st);
}

int inksam_op_decoder_i2t(const void* qsrc, int q_half, long long box_stride, int ld, int qcol, const int* box_img, const float* kv,
                          void* out, int out_half, int boxes, void* stream) {
  if (!qsrc || !kv || !out || boxes < 1) return set_error("inksam_op_decoder_i2t: null argument or no boxes");
  if (ld % 4 != 0 || qcol % 4 != 0 || qcol + 128 > ld || (reinterpret_cast<uintptr_t>(qsrc) & 15) != 0 || box_stride % 4 != 0)
    return set_error("inksam_op_decoder_i2t: q rows must be 16-byte aligned (ld=%d qcol=%d)", ld, qcol);
  return launch_decoder_i2t(qsrc, q_half, box_stride, ld, qcol, box_img, kv, out, out_half, boxes, static_cast<cudaStream_t>(stream));
}

int inksam_op_patchify(const void* images, int dtype, int B, const int* hw, long long img_stride, long long row_pitch,
                       long long plane_stride, const float* mean, const float* stdv, void* out, void* stream) {
  return launch_patchify(images, dtype, B, hw, img_stride, row_pitch, plane_stride, mean, stdv, out, static_cast<cudaStream_t>(stream));
}

int inksam_op_fill_pad_rows(void* qkv, const float* bias, int B, int cols, void* stream) {
  return launch_fill_pad_rows(qkv, bias, B, cols, static_cast<cudaStream_t>(stream));
}

int inksam_op_postprocess(const float* low, unsigned char* masks, float* logits, int K, int in_h, int in_w, int H, int W,
                          float thr, void* stream) {
  return launch_postprocess(low, masks, logits, K, in_h, in_w, H, W, thr, static_cast<cudaStream_t>(stream));
}

int inksam_op_postprocess_packed(const float* low, unsigned char* mask_bits, int K, int in_h, int in_w, int H, int W, float thr,
                                 void* stream) {
  return launch_postprocess(low, mask_bits, nullptr, K, in_h, in_w, H, W, thr, static_cast<cudaStream_t>(stream), 1);
}

}  // extern "C"
