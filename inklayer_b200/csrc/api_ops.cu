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
                   int a_conv3x3, int conv_c, void* stream) {
  GemmArgs g{};
  g.M = M; g.N = N; g.K = K; g.ldo = ldo; g.out = out;
  g.bias = bias; g.resid = resid; g.rowadd = rowadd; g.rowadd_mod = rowadd_mod;
  g.ln_w = ln_w; g.ln_b = ln_b; g.ln_eps = ln_eps; g.hyper = hyper;
  g.a_conv3x3 = a_conv3x3; g.conv_c = conv_c;
  return launch_gemm(cg, tf32, bn, epi, A, lda, W, ldw, g, device_sms(), static_cast<cudaStream_t>(stream));
}

}  // extern "C"
