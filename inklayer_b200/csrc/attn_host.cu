// Host side of the encoder attention kernels: tensor maps over the qkv matrix and the rel-pos table, launch.
#include "common.h"
#include "attn_tc.cuh"
#include "attn_global2.cuh"
#include "attn_window2.cuh"
#include "attn_window3.cuh"
#include "attn_window4.cuh"
#include "attn_window5.cuh"

#include <cstdlib>

namespace inksam {

template <int HD, int KPOLY>
static int launch_global(const __nv_bfloat16* qkv, const __nv_bfloat16* rel, const AttnArgs& a, cudaStream_t st) {
  CUtensorMap tm64, tm16, tr64, tr16;
  const uint64_t rows = (uint64_t)a.B * 4096, cols = 3ull * a.D;
  INK_TRY(make_tmap_bf16_box(&tm64, qkv, rows, cols, cols, 64, 128));
  INK_TRY(make_tmap_bf16_box(&tr64, rel, 256, HD, HD, 64, 128));
  if (HD == 80) {
    INK_TRY(make_tmap_bf16_box(&tm16, qkv, rows, cols, cols, 16, 128));
    INK_TRY(make_tmap_bf16_box(&tr16, rel, 256, HD, HD, 16, 128));
  } else {
    tm16 = tm64; tr16 = tr64;
  }
  using L = AttnGlobal2<HD>;
  auto kern = attn_global2_kernel<HD, KPOLY>;
  static PerDeviceFlag attr_done;
  if (!attr_done()) {
    CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::SMEM_BYTES));
    attr_done() = true;
  }
  kern<<<a.B * a.heads * 16, L::THREADS, L::SMEM_BYTES, st>>>(tm64, tm16, tr64, tr16, a);
  CUDA_TRY(cudaGetLastError());
  return 0;
}

template <int HD>
static int launch_window(const __nv_bfloat16* qkv, const __nv_bfloat16* rel, const AttnArgs& a, cudaStream_t st) {
  CUtensorMap tq64, tq16, tk64, tk16, tr64, tr16;
  const uint64_t rows = (uint64_t)a.B * 25 * 196, cols = 3ull * a.D;
  INK_TRY(make_tmap_bf16_box(&tq64, qkv, rows, cols, cols, 64, 128));
  INK_TRY(make_tmap_bf16_box(&tk64, qkv, rows, cols, cols, 64, 208));
  INK_TRY(make_tmap_bf16_box(&tr64, rel, 64, HD, HD, 64, 64));
  if (HD == 80) {
    INK_TRY(make_tmap_bf16_box(&tq16, qkv, rows, cols, cols, 16, 128));
    INK_TRY(make_tmap_bf16_box(&tk16, qkv, rows, cols, cols, 16, 208));
    INK_TRY(make_tmap_bf16_box(&tr16, rel, 64, HD, HD, 16, 64));
  } else {
    tq16 = tq64; tk16 = tk64; tr16 = tr64;
  }
  using L = AttnWindow2<HD>;
  auto kern = attn_window2_kernel<HD>;
  static PerDeviceFlag attr_done;
  static PerDeviceInt num_sms_dev;
  int& num_sms = num_sms_dev();
  if (!attr_done()) {
    CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::SMEM_BYTES));
    CUDA_TRY(cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, current_device()));
    attr_done() = true;
  }
  const int items = a.B * 25 * a.heads * 2;
  const int pairs = items / 2;  // a CTA takes both row tiles of a (window, head) so that K and V are loaded once
  const int grid = pairs < 2 * num_sms ? pairs : 2 * num_sms;  // persistent: 2 CTAs per SM
  kern<<<grid, L::THREADS, L::SMEM_BYTES, st>>>(tq64, tq16, tk64, tk16, tr64, tr16, a, items);
  CUDA_TRY(cudaGetLastError());
  return 0;
}

template <int HD, int KPOLY>
static int launch_window3(const __nv_bfloat16* qkv, const __nv_bfloat16* rel, const AttnArgs& a, cudaStream_t st) {
  CUtensorMap tq64, tq16, tk64, tk16, tr64, tr16;
  const uint64_t rows = (uint64_t)a.B * 25 * 196, cols = 3ull * a.D;
  INK_TRY(make_tmap_bf16_box(&tq64, qkv, rows, cols, cols, 64, 128));
  INK_TRY(make_tmap_bf16_box(&tk64, qkv, rows, cols, cols, 64, 208));
  INK_TRY(make_tmap_bf16_box(&tr64, rel, 64, HD, HD, 64, 64));
  if (HD == 80) {
    INK_TRY(make_tmap_bf16_box(&tq16, qkv, rows, cols, cols, 16, 128));
    INK_TRY(make_tmap_bf16_box(&tk16, qkv, rows, cols, cols, 16, 208));
    INK_TRY(make_tmap_bf16_box(&tr16, rel, 64, HD, HD, 16, 64));
  } else {
    tq16 = tq64; tk16 = tk64; tr16 = tr64;
  }
  using L = AttnWindow3<HD>;
  auto kern = attn_window3_kernel<HD, KPOLY>;
  static PerDeviceFlag attr_done;
  static PerDeviceInt num_sms_dev;
  int& num_sms = num_sms_dev();
  if (!attr_done()) {
    CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::SMEM_BYTES));
    CUDA_TRY(cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, current_device()));
    attr_done() = true;
  }
  const int pairs = a.B * 25 * a.heads;  // (image, window, head): two row tiles each, K and V loaded once
  const int grid = pairs < num_sms ? pairs : num_sms;  // persistent: one CTA per SM (512 TMEM columns, 187 KB of smem)
  kern<<<grid, L::THREADS, L::SMEM_BYTES, st>>>(tq64, tq16, tk64, tk16, tr64, tr16, a, pairs);
  CUDA_TRY(cudaGetLastError());
  return 0;
}

// INKSAM_GLOBAL_KPOLY=0|8|16: keys (of 64 per sub-tile and row) whose exp2 runs on the FMA pipe instead of the MUFU pipe.
template <int HD>
static int launch_global_any(const __nv_bfloat16* q, const __nv_bfloat16* r, const AttnArgs& a, cudaStream_t st) {
  const char* kp = std::getenv("INKSAM_GLOBAL_KPOLY");
  const int k = kp ? std::atoi(kp) : 0;  // B200: with the exponent clamp gone the kernel is not MUFU-bound; 0 is fastest
  if (k == 0) return launch_global<HD, 0>(q, r, a, st);
  if (k == 16) return launch_global<HD, 16>(q, r, a, st);
  if (k == 8) return launch_global<HD, 8>(q, r, a, st);
  return launch_global<HD, 0>(q, r, a, st);
}

template <int HD, int KPOLY, int VER>
static int launch_window4(const __nv_bfloat16* qkv, const __nv_bfloat16* rel, const AttnArgs& a, cudaStream_t st) {
  CUtensorMap tq64, tq16, tk64, tk16, tr64, tr16;
  const uint64_t rows = (uint64_t)a.B * 25 * 196, cols = 3ull * a.D;
  INK_TRY(make_tmap_bf16_box(&tq64, qkv, rows, cols, cols, 64, 128));
  INK_TRY(make_tmap_bf16_box(&tk64, qkv, rows, cols, cols, 64, 208));
  INK_TRY(make_tmap_bf16_box(&tr64, rel, 64, HD, HD, 64, 64));
  if (HD == 80) {
    INK_TRY(make_tmap_bf16_box(&tq16, qkv, rows, cols, cols, 16, 128));
    INK_TRY(make_tmap_bf16_box(&tk16, qkv, rows, cols, cols, 16, 208));
    INK_TRY(make_tmap_bf16_box(&tr16, rel, 64, HD, HD, 16, 64));
  } else {
    tq16 = tq64; tk16 = tk64; tr16 = tr64;
  }
  constexpr int kSmem = VER == 5 ? AttnWindow5<HD>::SMEM_BYTES : AttnWindow4<HD>::SMEM_BYTES;
  constexpr int kThreads = VER == 5 ? AttnWindow5<HD>::THREADS : AttnWindow4<HD>::THREADS;
  auto kern = VER == 5 ? attn_window5_kernel<HD, KPOLY> : attn_window4_kernel<HD, KPOLY>;
  static PerDeviceFlag attr_done;
  static PerDeviceInt num_sms_dev;
  int& num_sms = num_sms_dev();
  if (!attr_done()) {
    CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
    CUDA_TRY(cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, current_device()));
    attr_done() = true;
  }
  const int pairs = a.B * 25 * a.heads;  // (image, window, head): two row tiles each, K and V loaded once
  const int grid = pairs < num_sms ? pairs : num_sms;  // persistent: one CTA per SM (512 TMEM columns, 213 KB of smem)
  kern<<<grid, kThreads, kSmem, st>>>(tq64, tq16, tk64, tk16, tr64, tr16, a, pairs);
  CUDA_TRY(cudaGetLastError());
  return 0;
}

// The windowed kernel has three generations, selectable per launch for A/B runs and cross-checks in the tests:
// INKSAM_WINDOW_ATTN=2 attn_window2.cuh (two CTAs per SM, serial chain per CTA), 3 attn_window3.cuh (one CTA, 8 softmax warps
// on one item, key-split), 4 attn_window4.cuh (one CTA, role-specialised warps, two softmax groups), 5 attn_window5.cuh
// (the warps of 4 with the one-item-at-a-time schedule of 3).  kDefaultWindowAttn is
// the fastest measured on B200 (profiles/r02_window_attention.md).  INKSAM_WINDOW_KPOLY: share of exp2 on the FMA pipe.
constexpr int kDefaultWindowAttn = 5;
template <int HD>
static int launch_window_any(const __nv_bfloat16* q, const __nv_bfloat16* r, const AttnArgs& a, cudaStream_t st) {
  const char* v = std::getenv("INKSAM_WINDOW_ATTN");
  const int ver = (v && v[0] >= '2' && v[0] <= '5') ? v[0] - '0' : kDefaultWindowAttn;
  const char* kp = std::getenv("INKSAM_WINDOW_KPOLY");
  const int k = kp ? std::atoi(kp) : (ver == 5 ? 16 : 0);
  if (ver == 2) return launch_window<HD>(q, r, a, st);
  if (ver == 4) return k == 8 ? launch_window4<HD, 8, 4>(q, r, a, st) : launch_window4<HD, 0, 4>(q, r, a, st);
  if (ver == 5) {
    if (k == 0) return launch_window4<HD, 0, 5>(q, r, a, st);
    if (k == 4) return launch_window4<HD, 4, 5>(q, r, a, st);
    if (k == 12) return launch_window4<HD, 12, 5>(q, r, a, st);
    if (k == 16) return launch_window4<HD, 16, 5>(q, r, a, st);
    return launch_window4<HD, 8, 5>(q, r, a, st);
  }
  if (k == 16) return launch_window3<HD, 16>(q, r, a, st);
  return launch_window3<HD, 0>(q, r, a, st);
}

static long long* g_attn_trace = nullptr;
// Debug facility: device buffer [items][16] of long long that CTA 0 of the windowed kernel (v4) fills with the clock64() of
// its pipeline events (tools/window_trace.py); nullptr switches it off.
void set_attention_trace(long long* buf) { g_attn_trace = buf; }

// qkv: bf16 [rows][3D] (natural order for global blocks, window order for windowed blocks);
// rel: bf16 rel-pos table ([256][hd] global, [64][hd] windowed); out: bf16 [B*4096][D].
int launch_attention(int windowed, const void* qkv, const void* rel, void* out, int B, int heads, int head_dim,
                     cudaStream_t st) {
  AttnArgs a;
  a.B = B; a.heads = heads; a.D = heads * head_dim;
  a.c1 = (1.0f / sqrtf((float)head_dim)) * kLog2e;
  a.out = static_cast<__nv_bfloat16*>(out);
  a.trace = g_attn_trace;
  const auto* q = static_cast<const __nv_bfloat16*>(qkv);
  const auto* r = static_cast<const __nv_bfloat16*>(rel);
  if (head_dim == 80) return windowed ? launch_window_any<80>(q, r, a, st) : launch_global_any<80>(q, r, a, st);
  if (head_dim == 64) return windowed ? launch_window_any<64>(q, r, a, st) : launch_global_any<64>(q, r, a, st);
  return set_error("attention: head_dim %d not supported (64 or 80)", head_dim);
}

}  // namespace inksam
