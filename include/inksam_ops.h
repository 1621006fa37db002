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
                   int a_conv3x3, int conv_c, void* stream);

const char* inksam_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* INKSAM_OPS_H_ */
