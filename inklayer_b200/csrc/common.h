// Shared host-side declarations for libinksam: error plumbing and kernel launchers.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stddef.h>

#include "gemm_args.h"

namespace inksam {

// Records a printf-style message retrievable through inksam_last_error(); returns -1.
int set_error(const char* fmt, ...);
const char* get_error();

#define CUDA_TRY(expr)                                                                               \
  do {                                                                                               \
    cudaError_t _e = (expr);                                                                         \
    if (_e != cudaSuccess)                                                                           \
      return ::inksam::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
  } while (0)

#define INK_TRY(expr)            \
  do {                           \
    if ((expr) != 0) return -1;  \
  } while (0)

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize), cluster occupancy and the SM count are PER DEVICE: launchers keep their
// "already configured" state in one of these, indexed by the calling thread's current device (one process may hold
// handles on several GPUs: Sam.to('cuda:1') after 'cuda:0').
constexpr int kMaxDevices = 64;
inline int current_device() {
  int d = 0;
  cudaGetDevice(&d);
  return (d >= 0 && d < kMaxDevices) ? d : 0;
}
struct PerDeviceFlag {
  bool done[kMaxDevices] = {};
  bool& operator()() { return done[current_device()]; }
};
struct PerDeviceInt {
  int v[kMaxDevices] = {};
  int& operator()() { return v[current_device()]; }
};

int make_tmap_bf16_box(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld_elems,
                       uint32_t box_cols, uint32_t box_rows);

int launch_gemm(int cg, int tf32, int bn, int epi, const void* A, int lda, const void* W, int ldw, GemmArgs g,
                int num_sms, cudaStream_t st);

int launch_attention(int windowed, const void* qkv, const void* rel, void* out, int B, int heads, int head_dim,
                     cudaStream_t st);
void set_attention_trace(long long* buf);
int launch_patchify(const void* images, int dtype, int B, const int* hw, long long img_stride, long long row_pitch,
                    long long plane_stride, const float* mean, const float* stdv, void* out, cudaStream_t st);
int launch_layernorm_bf16(const float* x, const float* w, const float* b, void* out, int M, int D, float eps, int plain,
                          float* stats, cudaStream_t st);
int launch_fold_ln_weights(const float* W, const float* gamma, const float* beta, const float* bias, void* w_out, float* colsum,
                           float* bias_out, int N, int K, cudaStream_t st);
int launch_fill_pad_rows(void* qkv, const float* bias, int B, int cols, cudaStream_t st);
int launch_repack(const float* in, void* out, int out_bf16, const long long* ishape, int ndim, const int* perm,
                  cudaStream_t st);
// packed: masks are 1 bit per pixel, MSB first, rows padded to whole bytes ([K][H][(W+7)/8]) instead of one byte per pixel
int launch_postprocess(const float* low, unsigned char* masks, float* logits, int K, int in_h, int in_w, int H, int W,
                       float thr, cudaStream_t st, int packed = 0);
// boxes of several images (segments with their own geometry) in one launch; in_hw / orig_hw: [nseg][2]
int launch_postprocess_batch(int nseg, const float* const* low, unsigned char* const* masks, float* const* logits, const int* n,
                             const int* in_hw, const int* orig_hw, float thr, cudaStream_t st, int packed = 0);

}  // namespace inksam
