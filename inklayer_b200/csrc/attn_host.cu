// Host side of the encoder attention kernels: tensor maps over the qkv matrix and the rel-pos table, launch.
#include "common.h"
#include "attn_tc.cuh"
#include "attn_global2.cuh"
#include "attn_window2.cuh"

namespace inksam {

template <int HD>
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
  auto kern = attn_global2_kernel<HD>;
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

// qkv: bf16 [rows][3D] (natural order for global blocks, window order for windowed blocks);
// rel: bf16 rel-pos table ([256][hd] global, [64][hd] windowed); out: bf16 [B*4096][D].
int launch_attention(int windowed, const void* qkv, const void* rel, void* out, int B, int heads, int head_dim,
                     cudaStream_t st) {
  AttnArgs a;
  a.B = B; a.heads = heads; a.D = heads * head_dim;
  a.c1 = (1.0f / sqrtf((float)head_dim)) * kLog2e;
  a.out = static_cast<__nv_bfloat16*>(out);
  const auto* q = static_cast<const __nv_bfloat16*>(qkv);
  const auto* r = static_cast<const __nv_bfloat16*>(rel);
  if (head_dim == 80) return windowed ? launch_window<80>(q, r, a, st) : launch_global<80>(q, r, a, st);
  if (head_dim == 64) return windowed ? launch_window<64>(q, r, a, st) : launch_global<64>(q, r, a, st);
  return set_error("attention: head_dim %d not supported (64 or 80)", head_dim);
}

}  // namespace inksam
