// ResizeLongestSide.apply_image on the device: the uint8 bilinear resize that the reference runs on the
// host through torchvision -> PIL.Image.resize (SA/utils/transforms.py:26-31).  Pillow's 8-bit resampler
// (libImaging/Resample.c: precompute_coeffs, normalize_coeffs_8bpc, ImagingResampleHorizontal_8bpc /
// ImagingResampleVertical_8bpc, ImagingResampleInner) is integer arithmetic on double-precision coefficient tables, so
// the result here is BIT-EXACT with the host path:
//   coefficients: per output index, triangle weights over [center - support, center + support), normalised in fp64
//                 (explicit __dmul_rn / __dadd_rn / __ddiv_rn: no FMA contraction, the same roundings as the C code),
//                 then fixed point int(0.5 + w * 2^22);
//   pass:         2^21 + sum(pixel * coeff) in int32, >> 22, clamped to 0..255; horizontal pass first (only the source
//                 rows the vertical pass reads), vertical pass on the uint8 intermediate.
// HBM-bound byte work: reads H*W*3, writes out_h*out_w*3 (+ the intermediate out_w*H*3 twice).
#include "common.h"
#include "inksam_ops.h"

namespace inksam {
namespace {

constexpr int kPrecisionBits = 32 - 8 - 2;

struct AxisGeom {
  int in_size, out_size, ksize;
  double scale, support, ss;
};

AxisGeom axis_geom(int in_size, int out_size) {
  AxisGeom g;
  g.in_size = in_size; g.out_size = out_size;
  g.scale = (double)((float)in_size - 0.0f) / out_size;  // Pillow carries the box as float
  const double filterscale = g.scale < 1.0 ? 1.0 : g.scale;
  g.support = 1.0 * filterscale;
  g.ksize = (int)ceil(g.support) * 2 + 1;
  g.ss = 1.0 / filterscale;
  return g;
}

inline size_t up256(size_t x) { return (x + 255) & ~size_t(255); }

// bounds: int2 (xmin, count) per output index; kk: ksize ints per output index
__global__ void resize_coeffs_kernel(AxisGeom g, int2* __restrict__ bounds, int* __restrict__ kk) {
  const int xx = blockIdx.x * blockDim.x + threadIdx.x;
  if (xx >= g.out_size) return;
  const double center = __dadd_rn(0.0, __dmul_rn((double)xx + 0.5, g.scale));
  int xmin = (int)__dadd_rn(__dadd_rn(center, -g.support), 0.5);
  if (xmin < 0) xmin = 0;
  int xmax = (int)__dadd_rn(__dadd_rn(center, g.support), 0.5);
  if (xmax > g.in_size) xmax = g.in_size;
  xmax -= xmin;
  int* k = kk + (size_t)xx * g.ksize;
  double ww = 0.0;
  for (int x = 0; x < xmax; ++x) {
    double a = __dmul_rn(__dadd_rn(__dadd_rn((double)(x + xmin), -center), 0.5), g.ss);
    if (a < 0.0) a = -a;
    const double w = a < 1.0 ? __dadd_rn(1.0, -a) : 0.0;
    ww = __dadd_rn(ww, w);
  }
  for (int x = 0; x < g.ksize; ++x) {
    double w = 0.0;
    if (x < xmax) {
      double a = __dmul_rn(__dadd_rn(__dadd_rn((double)(x + xmin), -center), 0.5), g.ss);
      if (a < 0.0) a = -a;
      w = a < 1.0 ? __dadd_rn(1.0, -a) : 0.0;
      if (ww != 0.0) w = __ddiv_rn(w, ww);
    }
    k[x] = (int)__dadd_rn(0.5, __dmul_rn(w, (double)(1 << kPrecisionBits)));
  }
  bounds[xx] = make_int2(xmin, xmax);
}

__device__ __forceinline__ unsigned char clip8(int v) {
  v >>= kPrecisionBits;
  return (unsigned char)(v < 0 ? 0 : (v > 255 ? 255 : v));
}

// dst[y][xx][c] = resample of src row (y0 + y) along x.  One thread per output pixel (3 channels).
__global__ void __launch_bounds__(256) resize_horizontal_kernel(const unsigned char* __restrict__ src, long long src_pitch, int y0,
                                                                int rows, int out_w, int ksize, const int2* __restrict__ bounds,
                                                                const int* __restrict__ kk, unsigned char* __restrict__ dst,
                                                                long long dst_pitch) {
  const int xx = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (xx >= out_w || y >= rows) return;
  const int2 b = bounds[xx];
  const int* k = kk + (size_t)xx * ksize;
  const unsigned char* p = src + (size_t)(y0 + y) * src_pitch + (size_t)b.x * 3;
  int s0 = 1 << (kPrecisionBits - 1), s1 = s0, s2 = s0;
  for (int x = 0; x < b.y; ++x) {
    const int c = k[x];
    s0 += p[3 * x] * c; s1 += p[3 * x + 1] * c; s2 += p[3 * x + 2] * c;
  }
  unsigned char* o = dst + (size_t)y * dst_pitch + (size_t)xx * 3;
  o[0] = clip8(s0); o[1] = clip8(s1); o[2] = clip8(s2);
}

// dst[yy][i] = resample of column byte i (i over w*3 bytes of a row) along y.  One thread per output byte.
__global__ void __launch_bounds__(256) resize_vertical_kernel(const unsigned char* __restrict__ src, long long src_pitch,
                                                              int y_shift, int row_bytes, int out_h, int ksize,
                                                              const int2* __restrict__ bounds, const int* __restrict__ kk,
                                                              unsigned char* __restrict__ dst, long long dst_pitch) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x, yy = blockIdx.y;
  if (i >= row_bytes || yy >= out_h) return;
  const int2 b = bounds[yy];
  const int* k = kk + (size_t)yy * ksize;
  const unsigned char* p = src + (size_t)(b.x - y_shift) * src_pitch + i;
  int s = 1 << (kPrecisionBits - 1);
  for (int y = 0; y < b.y; ++y) s += p[(size_t)y * src_pitch] * k[y];
  dst[(size_t)yy * dst_pitch + i] = clip8(s);
}

struct Plan {
  AxisGeom gh, gv;
  bool need_h, need_v;
  size_t off_hb, off_hk, off_vb, off_vk, off_tmp, total;
};

Plan make_plan(int H, int W, int out_h, int out_w) {
  Plan p{};
  p.need_h = out_w != W; p.need_v = out_h != H;
  p.gh = axis_geom(W, out_w); p.gv = axis_geom(H, out_h);
  size_t off = 0;
  p.off_hb = off; off += up256((size_t)out_w * sizeof(int2));
  p.off_hk = off; off += up256((size_t)out_w * p.gh.ksize * sizeof(int));
  p.off_vb = off; off += up256((size_t)out_h * sizeof(int2));
  p.off_vk = off; off += up256((size_t)out_h * p.gv.ksize * sizeof(int));
  p.off_tmp = off; off += (p.need_h && p.need_v) ? up256((size_t)H * out_w * 3) : 0;
  p.total = off;
  return p;
}

// ------------------------------------------------------------------------------------------------ static scratch
// Scratch of inksam_resize_u8_static: coefficient tables + the uint8 intermediate of the two-pass resample, in a file-scope
// device array instead of a caller-owned workspace, for callers that cannot allocate (CUDA graph capture, serving loops).
// Every GPU that loads the library holds these bytes until the process exits.  All calls share the array; nothing in it
// has to start at zero (the coefficient kernels and the horizontal pass write every byte the later kernels read).
constexpr size_t kStaticResizeBytes = (size_t)16 << 20;  // 16 MiB: e.g. a 4096 x 4096 x 3 source down to 1024 x 1024 (12.6 MB)
static __device__ __align__(256) unsigned char g_resize_scratch[kStaticResizeBytes];

// The array's name is a host-side stand-in in host code, not a device address: the address is looked up once per device,
// with that device current, and cached.
int static_resize_scratch(void** out) {
  static void* ptr[kMaxDevices] = {};
  static PerDeviceFlag have;
  int dev = 0;
  CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) return set_error("inksam_resize_u8_static: device %d out of range", dev);
  if (!have.done[dev]) {
    void* p = nullptr;
    CUDA_TRY(cudaGetSymbolAddress(&p, g_resize_scratch));
    if (p == nullptr) return set_error("inksam_resize_u8_static: cudaGetSymbolAddress returned a null address");
    ptr[dev] = p;
    have.done[dev] = true;
  }
  *out = ptr[dev];
  return 0;
}

}  // namespace
}  // namespace inksam

using namespace inksam;

extern "C" {

size_t inksam_resize_workspace_bytes(int H, int W, int out_h, int out_w) {
  if (H < 1 || W < 1 || out_h < 1 || out_w < 1) return 0;
  return make_plan(H, W, out_h, out_w).total;
}

int inksam_resize_u8(const uint8_t* src, int H, int W, long long src_row_pitch, uint8_t* dst, int out_h, int out_w,
                     long long dst_row_pitch, void* workspace, void* stream) {
  if (!src || !dst) return set_error("inksam_resize_u8: null image pointer");
  if (H < 1 || W < 1 || out_h < 1 || out_w < 1) return set_error("inksam_resize_u8: bad sizes %dx%d -> %dx%d", H, W, out_h, out_w);
  if (H > 65535) return set_error("inksam_resize_u8: H=%d exceeds 65535 rows", H);
  if (src_row_pitch < (long long)W * 3 || dst_row_pitch < (long long)out_w * 3)
    return set_error("inksam_resize_u8: row pitch smaller than a row");
  auto st = static_cast<cudaStream_t>(stream);
  const Plan p = make_plan(H, W, out_h, out_w);
  if (!p.need_h && !p.need_v) {
    CUDA_TRY(cudaMemcpy2DAsync(dst, (size_t)dst_row_pitch, src, (size_t)src_row_pitch, (size_t)W * 3, (size_t)H,
                               cudaMemcpyDeviceToDevice, st));
    return 0;
  }
  if (!workspace) return set_error("inksam_resize_u8: null workspace (inksam_resize_workspace_bytes)");
  char* ws = static_cast<char*>(workspace);
  int2* hb = reinterpret_cast<int2*>(ws + p.off_hb); int* hk = reinterpret_cast<int*>(ws + p.off_hk);
  int2* vb = reinterpret_cast<int2*>(ws + p.off_vb); int* vk = reinterpret_cast<int*>(ws + p.off_vk);
  unsigned char* tmp = reinterpret_cast<unsigned char*>(ws + p.off_tmp);
  if (p.need_h) resize_coeffs_kernel<<<(out_w + 127) / 128, 128, 0, st>>>(p.gh, hb, hk);
  if (p.need_v) resize_coeffs_kernel<<<(out_h + 127) / 128, 128, 0, st>>>(p.gv, vb, vk);
  CUDA_TRY(cudaGetLastError());
  if (p.need_h && !p.need_v) {
    resize_horizontal_kernel<<<dim3((out_w + 255) / 256, H), 256, 0, st>>>(src, src_row_pitch, 0, H, out_w, p.gh.ksize, hb, hk, dst,
                                                                         dst_row_pitch);
  } else if (!p.need_h) {
    resize_vertical_kernel<<<dim3((W * 3 + 255) / 256, out_h), 256, 0, st>>>(src, src_row_pitch, 0, W * 3, out_h, p.gv.ksize, vb, vk,
                                                                           dst, dst_row_pitch);
  } else {
    // every source row is resampled horizontally (ImagingResampleInner restricts this to the rows the vertical pass
    // reads, which changes no value), then the vertical pass runs on the uint8 intermediate
    resize_horizontal_kernel<<<dim3((out_w + 255) / 256, H), 256, 0, st>>>(src, src_row_pitch, 0, H, out_w, p.gh.ksize, hb, hk, tmp,
                                                                         (long long)out_w * 3);
    resize_vertical_kernel<<<dim3((out_w * 3 + 255) / 256, out_h), 256, 0, st>>>(tmp, (long long)out_w * 3, 0, out_w * 3, out_h,
                                                                               p.gv.ksize, vb, vk, dst, dst_row_pitch);
  }
  CUDA_TRY(cudaGetLastError());
  return 0;
}

size_t inksam_resize_static_max_bytes(void) { return kStaticResizeBytes; }

int inksam_resize_u8_static(const uint8_t* src, int H, int W, long long src_row_pitch, uint8_t* dst, int out_h, int out_w,
                            long long dst_row_pitch, void* stream) {
  if (H < 1 || W < 1 || out_h < 1 || out_w < 1)
    return set_error("inksam_resize_u8_static: bad sizes %dx%d -> %dx%d", H, W, out_h, out_w);
  const size_t need = make_plan(H, W, out_h, out_w).total;
  if (need > kStaticResizeBytes)
    return set_error("inksam_resize_u8_static: %dx%d -> %dx%d needs %zu bytes of scratch, the static array holds %zu "
                     "(use inksam_resize_u8 with a caller-owned workspace)", H, W, out_h, out_w, need, kStaticResizeBytes);
  void* scratch = nullptr;
  INK_TRY(static_resize_scratch(&scratch));
  return inksam_resize_u8(src, H, W, src_row_pitch, dst, out_h, out_w, dst_row_pitch, scratch, stream);
}

}  // extern "C"
