// C-ABI kernel-level entry points (include/inksam_ops.h).
#include "common.h"
#include "inksam_ops.h"

using namespace inksam;

static int device_sms() {
  static int sms = 0;
  if (sms == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  }
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

int inksam_op_attention(int windowed, const void* qkv, const void* rel, void* out, int B, int heads, int head_dim, void* stream) {
  return launch_attention(windowed, qkv, rel, out, B, heads, head_dim, static_cast<cudaStream_t>(stream));
}

int inksam_op_layernorm(const float* x, const float* w, const float* b, void* out, int M, int D, float eps, int plain, void* stream) {
  return launch_layernorm_bf16(x, w, b, out, M, D, eps, plain, static_cast<cudaStream_t>(stream));
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

}  // extern "C"
