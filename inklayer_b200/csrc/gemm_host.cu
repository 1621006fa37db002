// Host side of the tcgen05 GEMM: tensor-map construction (driver entry point fetched at run time, so the
// library links against cudart only) and the template dispatch table.
#include "common.h"
#include "gemm_tc.cuh"

#include <cudaTypedefs.h>
#include <mutex>

namespace inksam {

static PFN_cuTensorMapEncodeTiled_v12000 g_encode = nullptr;

static int fetch_encode() {
  static std::once_flag once;
  std::call_once(once, [] {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      g_encode = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(fn);
  });
  return g_encode ? 0 : -1;
}

// Row-major [rows][cols] matrix (cols contiguous); box = 128 bytes of columns x box_rows, 128B swizzle.
int make_tmap_2d(CUtensorMap* out, const void* base, int elem_bytes, uint64_t rows, uint64_t cols, uint64_t ld_elems,
                 uint32_t box_rows) {
  if (fetch_encode()) return set_error("cuTensorMapEncodeTiled entry point not available");
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld_elems * (uint64_t)elem_bytes};
  cuuint32_t box[2] = {(cuuint32_t)(128 / elem_bytes), box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUtensorMapDataType dt = elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_TFLOAT32;
  CUresult r = g_encode(out, dt, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return set_error("cuTensorMapEncodeTiled(2d) failed: %d (rows=%llu cols=%llu ld=%llu)", (int)r,
                                          (unsigned long long)rows, (unsigned long long)cols, (unsigned long long)ld_elems);
  return 0;
}

// fp32 row-major [rows][cols] output matrix for the staged epilogue: box = 16 columns (64 bytes) x 32 rows, 64B swizzle.
int make_tmap_f32_out(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld_elems) {
  if (fetch_encode()) return set_error("cuTensorMapEncodeTiled entry point not available");
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld_elems * 4};
  cuuint32_t box[2] = {16, 32};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = g_encode(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(base), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return set_error("cuTensorMapEncodeTiled(f32 out) failed: %d (rows=%llu cols=%llu ld=%llu)", (int)r,
                                          (unsigned long long)rows, (unsigned long long)cols, (unsigned long long)ld_elems);
  return 0;
}

// fp16 row-major [rows][cols] output (GemmArgs::out_f16): the geometry of the bf16 copy below
int make_tmap_f16_out(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld_elems) {
  if (fetch_encode()) return set_error("cuTensorMapEncodeTiled entry point not available");
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld_elems * 2};
  cuuint32_t box[2] = {32, 32};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = g_encode(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return set_error("cuTensorMapEncodeTiled(f16 out) failed: %d (rows=%llu cols=%llu ld=%llu)", (int)r,
                                          (unsigned long long)rows, (unsigned long long)cols, (unsigned long long)ld_elems);
  return 0;
}

// bf16 row-major [rows][cols] copy of the fp32 output (EPI_BIAS_RESID_STATS_F32): box = 32 columns (64 bytes) x 32 rows,
// 64B swizzle -- the staging geometry of an fp32 half.
int make_tmap_bf16_out(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld_elems) {
  if (fetch_encode()) return set_error("cuTensorMapEncodeTiled entry point not available");
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld_elems * 2};
  cuuint32_t box[2] = {32, 32};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = g_encode(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return set_error("cuTensorMapEncodeTiled(bf16 out) failed: %d (rows=%llu cols=%llu ld=%llu)", (int)r,
                                          (unsigned long long)rows, (unsigned long long)cols, (unsigned long long)ld_elems);
  return 0;
}

// bf16 row-major [rows][cols] matrix, arbitrary box; swizzle 128B (box_cols == 64) or 32B (box_cols == 16).
int make_tmap_bf16_box(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld_elems,
                       uint32_t box_cols, uint32_t box_rows) {
  if (fetch_encode()) return set_error("cuTensorMapEncodeTiled entry point not available");
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld_elems * 2};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  const CUtensorMapSwizzle sw = box_cols == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_32B;
  CUresult r = g_encode(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return set_error("cuTensorMapEncodeTiled(box %ux%u) failed: %d (rows=%llu cols=%llu ld=%llu)",
                                          box_cols, box_rows, (int)r, (unsigned long long)rows,
                                          (unsigned long long)cols, (unsigned long long)ld_elems);
  return 0;
}

// Zero-bordered channels-last map [nb*66][66][C] bf16 for the 3x3 conv; box = 64 channels x 64 x 2 rows.
int make_tmap_conv3(CUtensorMap* out, const void* base, int nb, int C) {
  if (fetch_encode()) return set_error("cuTensorMapEncodeTiled entry point not available");
  cuuint64_t dims[3] = {(cuuint64_t)C, 66, (cuuint64_t)nb * 66};
  cuuint64_t strides[2] = {(cuuint64_t)C * 2, (cuuint64_t)C * 2 * 66};
  cuuint32_t box[3] = {64, 64, 2};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = g_encode(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return set_error("cuTensorMapEncodeTiled(conv3) failed: %d", (int)r);
  return 0;
}

template <int CG, int TF32, int BN, int EPI>
static int launch_one(const CUtensorMap& tA, const CUtensorMap& tB, const CUtensorMap& tO, const CUtensorMap& tX, const GemmArgs& g, int num_sms,
                      cudaStream_t st) {
  using C = GemmCfg<CG, TF32, BN, gemm_staged(EPI)>;
  auto kern = gemm_tc_kernel<CG, TF32, BN, EPI>;
  static PerDeviceFlag attr_done;  // per device (and per instantiation)
  if (!attr_done()) {
    CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES));
    attr_done() = true;
  }
  const int tiles = ((g.M + C::BM * CG - 1) / (C::BM * CG)) * ((g.N + BN - 1) / BN);
  cudaLaunchConfig_t cfg{};
  cfg.blockDim = dim3(384);
  cfg.dynamicSmemBytes = C::SMEM_BYTES;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CG;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  // programmatic dependent launch: the prologue may overlap the tail of the preceding kernel of the stream; the kernel
  // waits (griddepcontrol.wait) before it touches global data
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 2;
  // The tile loop is statically strided over the resident clusters, so the grid must not exceed what is
  // co-resident: CTA pairs need both SMs of a TPC, and not every SM of the 148 can be paired on every part.
  static PerDeviceInt max_clusters_dev;
  int& max_clusters = max_clusters_dev();
  if (max_clusters == 0) {
    max_clusters = num_sms / CG;
    if (CG > 1) {
      cfg.gridDim = dim3((num_sms / CG) * CG);
      int n = 0;
      if (cudaOccupancyMaxActiveClusters(&n, kern, &cfg) == cudaSuccess && n > 0 && n < max_clusters) max_clusters = n;
    }
  }
  int clusters = max_clusters;
  if (tiles < clusters) clusters = tiles;
  cfg.gridDim = dim3(clusters * CG);
  CUDA_TRY(cudaLaunchKernelEx(&cfg, kern, tA, tB, tO, tX, g));
  return 0;
}

#define GEMM_CASE(CG_, TF_, BN_, EPI_)                                             \
  if (cg == CG_ && tf32 == TF_ && bn == BN_ && epi == EPI_)                        \
    return launch_one<CG_, TF_, BN_, EPI_>(tA, tB, tO, tX, g, num_sms, st);

// A: [M][K] (ld = lda elements) or the padded conv map when g.a_conv3x3; W: [N][K] row-major.
int launch_gemm(int cg, int tf32, int bn, int epi, const void* A, int lda, const void* W, int ldw, GemmArgs g,
                int num_sms, cudaStream_t st) {
  const int esz = tf32 ? 4 : 2;
  if (g.N % 32 != 0) return set_error("gemm: N=%d must be a multiple of 32", g.N);
  if ((g.K * esz) % 16 != 0) return set_error("gemm: K=%d row bytes must be a multiple of 16", g.K);
  if (epi_is_fold(epi) && (!g.stats_in || !g.colsum || !g.bias || g.stats_np < 1))
    return set_error("gemm: the folded-LayerNorm epilogue needs stats_in, colsum, bias and stats_np >= 1");
  if (epi == EPI_BIAS_RESID_STATS_F32 && (!g.stats_out || !g.xb_out))
    return set_error("gemm: the residual + statistics epilogue needs stats_out and xb_out");
  CUtensorMap tA, tB;
  if (g.a_conv3x3) {
    if (make_tmap_conv3(&tA, A, g.M / 4096, g.conv_c)) return -1;
  } else {
    if (make_tmap_2d(&tA, A, esz, g.M, g.K, lda, 128)) return -1;
  }
  if (make_tmap_2d(&tB, W, esz, g.N, g.K, ldw, bn / cg)) return -1;
  if (g.out_f16 && epi != EPI_BIAS_ROWADD_F32 && epi != EPI_RESID_LN_F32 && epi != EPI_UPSCALE1)
    return set_error("gemm: fp16 output is built for the row-add, residual + LayerNorm and upscale-1 epilogues only (epi=%d)", epi);
  if (g.ab_f16 && tf32) return set_error("gemm: fp16 operands go with tf32 = 0");
  CUtensorMap tO = tA;  // only the staged fp32 epilogues store through a tensor map
  if ((gemm_staged(epi) == 1 || gemm_staged(epi) == 2) && !g.out_f16) {
    if ((reinterpret_cast<uintptr_t>(g.out) & 15) != 0 || (g.ldo * 4) % 16 != 0)
      return set_error("gemm: epilogue %d stores its fp32 output by TMA and needs a 16-byte aligned output (ldo=%d)", epi, g.ldo);
    if (make_tmap_f32_out(&tO, g.out, g.M, g.N, g.ldo)) return -1;
  }
  CUtensorMap tX = tA;  // bf16 copy of the residual output
  if (epi == EPI_BIAS_RESID_STATS_F32) {
    if ((reinterpret_cast<uintptr_t>(g.xb_out) & 15) != 0 || (g.ldo * 2) % 16 != 0)
      return set_error("gemm: the bf16 copy needs a 16-byte aligned xb_out (ldo=%d)", g.ldo);
    if (make_tmap_bf16_out(&tX, g.xb_out, g.M, g.N, g.ldo)) return -1;
  }
  if (gemm_staged(epi) == 3 && !g.win_remap) {
    if ((reinterpret_cast<uintptr_t>(g.out) & 15) != 0 || (g.ldo * 2) % 16 != 0)
      return set_error("gemm: the bf16 output is stored by TMA and needs a 16-byte aligned out (ldo=%d)", g.ldo);
    if (make_tmap_bf16_out(&tX, g.out, g.M, g.N, g.ldo)) return -1;
  }
  if (g.out_f16 && epi != EPI_UPSCALE1) {
    if ((reinterpret_cast<uintptr_t>(g.out) & 15) != 0 || (g.ldo * 2) % 16 != 0)
      return set_error("gemm: the fp16 output needs a 16-byte aligned out (ldo=%d)", g.ldo);
    if (make_tmap_f16_out(&tX, g.out, g.M, g.N, g.ldo)) return -1;
  }

  GEMM_CASE(1, 0, 256, EPI_BIAS_BF16)
  GEMM_CASE(2, 0, 256, EPI_BIAS_BF16)
  GEMM_CASE(1, 0, 256, EPI_BIAS_GELU_BF16)
  GEMM_CASE(2, 0, 256, EPI_BIAS_GELU_BF16)
  GEMM_CASE(1, 0, 256, EPI_BIAS_RESID_F32)
  GEMM_CASE(2, 0, 256, EPI_BIAS_RESID_F32)
  GEMM_CASE(2, 0, 256, EPI_FOLD_BF16)
  GEMM_CASE(2, 0, 256, EPI_FOLD_GELU_BF16)
  GEMM_CASE(2, 0, 256, EPI_BIAS_RESID_STATS_F32)
  GEMM_CASE(1, 0, 256, EPI_BIAS_ROWADD_F32)
  GEMM_CASE(2, 0, 256, EPI_BIAS_ROWADD_F32)
  GEMM_CASE(1, 0, 256, EPI_LN_BF16_PAD)
  GEMM_CASE(2, 0, 256, EPI_LN_BF16_PAD)
  GEMM_CASE(1, 0, 256, EPI_LN_F32_NCHW)
  GEMM_CASE(2, 0, 256, EPI_LN_F32_NCHW)
  GEMM_CASE(1, 0, 256, EPI_BIAS_F32)
  GEMM_CASE(2, 0, 256, EPI_BIAS_F32)
  // tf32 (fp32 activations and weights in memory): the decoder's image-side linears
  GEMM_CASE(1, 1, 256, EPI_BIAS_F32)
  GEMM_CASE(2, 1, 256, EPI_BIAS_F32)
  GEMM_CASE(1, 1, 256, EPI_BIAS_ROWADD_F32)
  GEMM_CASE(2, 1, 256, EPI_BIAS_ROWADD_F32)
  GEMM_CASE(1, 1, 256, EPI_RESID_LN_F32)
  GEMM_CASE(2, 1, 256, EPI_RESID_LN_F32)
  GEMM_CASE(1, 1, 256, EPI_UPSCALE1)
  GEMM_CASE(2, 1, 256, EPI_UPSCALE1)
  GEMM_CASE(1, 1, 128, EPI_BIAS_F32)
  GEMM_CASE(2, 1, 128, EPI_BIAS_F32)
  GEMM_CASE(1, 1, 128, EPI_BIAS_ROWADD_F32)
  GEMM_CASE(2, 1, 128, EPI_BIAS_ROWADD_F32)
  GEMM_CASE(1, 1, 128, EPI_UPSCALE2_MASK)
  GEMM_CASE(2, 1, 128, EPI_UPSCALE2_MASK)
  // 2-byte operands with the decoder's epilogues: fp16 GEMM-to-GEMM intermediates (GemmArgs::ab_f16)
  GEMM_CASE(2, 0, 128, EPI_BIAS_ROWADD_F32)
  GEMM_CASE(2, 0, 256, EPI_RESID_LN_F32)
  GEMM_CASE(2, 0, 256, EPI_UPSCALE1)
  GEMM_CASE(2, 0, 128, EPI_UPSCALE2_MASK)
  return set_error("gemm: no kernel for cg=%d tf32=%d bn=%d epi=%d", cg, tf32, bn, epi);
}

}  // namespace inksam
