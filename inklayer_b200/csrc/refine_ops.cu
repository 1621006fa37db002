// First consumers of the SAM masks, as HBM-bound bit / integer kernels:
//
//   inksam_clean_masks   InkLayer/refinement/mask_cleaner.py:11-36 clean_up_mask: threshold, morphological close with a
//                        k x k rectangle (cv2.morphologyEx MORPH_CLOSE, default border = ignored pixels), 8-connected
//                        components with statistics (cv2.connectedComponentsWithStats), keep a component if its area
//                        exceeds 500 pixels or its bounding box is not square within 10 %.
//   inksam_content_iou   InkLayer/refinement/nms_sketch.py:62-78,186-262: masks restricted to the sketch's stroke pixels
//                        (grey < 250), pairwise intersection counts (union = area_i + area_j - intersection).
//
// Masks are bit-packed once (32 pixels per word, rows padded to whole words, padding bits kept 0), so every later pass
// moves 1/8 of the uint8 bytes.  All arithmetic is integer: results are bit-exact against the reference.
#include "common.h"
#include "inksam_refine.h"

#include <climits>

namespace inksam {
namespace {

// ------------------------------------------------------------------------------------------------ bit packing
// One warp per output word: lane b tests pixel x = 32 w + b.  gate (optional, [H][W]): the bit survives only where
// gate < gate_lt (the sketch's stroke pixels, nms_sketch.py:74).
__global__ void __launch_bounds__(256) pack_bits_kernel(const unsigned char* __restrict__ in, int thresh,
                                                        const unsigned char* __restrict__ gate, int gate_lt, int H, int W,
                                                        int WW, long long total_words, unsigned* __restrict__ bits) {
  const long long word = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (word >= total_words) return;
  const int lane = threadIdx.x & 31;
  const int w = (int)(word % WW);
  const long long ky = word / WW;  // mask * H + y
  const int y = (int)(ky % H);
  const int x = w * 32 + lane;
  bool on = false;
  if (x < W) {
    on = in[ky * W + x] > thresh;
    if (on && gate != nullptr) on = gate[(long long)y * W + x] < gate_lt;
  }
  const unsigned b = __ballot_sync(0xffffffffu, on);
  if (lane == 0) bits[word] = b;
}

__device__ __forceinline__ unsigned tail_mask(int w, int W, int WW) {  // valid-pixel bits of word w of a row
  return (w == WW - 1 && (W & 31)) ? ((1u << (W & 31)) - 1u) : 0xffffffffu;
}

// Horizontal pass of a (2r+1)-wide rectangle: dilate (OR) or erode (AND) over x-r..x+r; pixels outside the image do not
// take part (cv2's morphologyDefaultBorderValue), i.e. they read as 0 for the OR and as 1 for the AND.
template <bool ERODE>
__global__ void __launch_bounds__(256) morph_h_kernel(const unsigned* __restrict__ in, unsigned* __restrict__ out, int W, int WW,
                                                      int r, long long total_words) {
  const long long word = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (word >= total_words) return;
  const int w = (int)(word % WW);
  const unsigned* row = in + (word - w);
  auto get = [&](int i) -> unsigned {  // word i of the row as the operator sees it
    if (i < 0 || i >= WW) return ERODE ? 0xffffffffu : 0u;
    const unsigned v = row[i];
    return ERODE ? (v | ~tail_mask(i, W, WW)) : v;
  };
  unsigned acc = get(w);
  for (int s = 1; s <= r; ++s) {
    const int q = s >> 5, rem = s & 31;
    unsigned from_left, from_right;  // the row shifted so that pixel x sees x-s / x+s
    if (rem == 0) {
      from_left = get(w - q);
      from_right = get(w + q);
    } else {
      from_left = (get(w - q) << rem) | (get(w - q - 1) >> (32 - rem));
      from_right = (get(w + q) >> rem) | (get(w + q + 1) << (32 - rem));
    }
    acc = ERODE ? (acc & from_left & from_right) : (acc | from_left | from_right);
  }
  out[word] = acc & tail_mask(w, W, WW);
}

// Vertical pass: rows y-r..y+r clipped to the image.
template <bool ERODE>
__global__ void __launch_bounds__(256) morph_v_kernel(const unsigned* __restrict__ in, unsigned* __restrict__ out, int H, int WW,
                                                      int r, long long total_words) {
  const long long word = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (word >= total_words) return;
  const int w = (int)(word % WW);
  const long long ky = word / WW;
  const int y = (int)(ky % H);
  const unsigned* col = in + (ky - y) * WW + w;  // word w of row 0 of this mask
  const int y0 = y - r < 0 ? 0 : y - r, y1 = y + r > H - 1 ? H - 1 : y + r;
  unsigned acc = ERODE ? 0xffffffffu : 0u;
  for (int yy = y0; yy <= y1; ++yy) {
    const unsigned v = col[(long long)yy * WW];
    acc = ERODE ? (acc & v) : (acc | v);
  }
  out[word] = acc;
}

// ------------------------------------------------------------------------------------------------ connected components
// Run-based union-find on the bit rows (8-connectivity).  A pixel's first label is the start of its horizontal run, so
// rows are flat from the beginning and a union is needed only once per contact between a run and a run of the row
// above; labels are pixel indices inside the mask (y * W + x), -1 for background.
struct CcStats {  // per-root statistics, indexed like the labels (touched at run starts only)
  int* area; int* minx; int* maxx; int* miny; int* maxy;
};

__device__ __forceinline__ bool bit_at(const unsigned* row, int x, int W) {
  return x >= 0 && x < W && ((row[x >> 5] >> (x & 31)) & 1u);
}

// first pixel of the run that contains the set pixel x of `row`
__device__ __forceinline__ int run_start(const unsigned* row, int x) {
  int w = x >> 5;
  const int b = x & 31;
  unsigned zeros_below = ~row[w] & ((b == 0) ? 0u : (0xffffffffu >> (32 - b)));
  while (zeros_below == 0u) {  // the run reaches bit 0 of this word
    if (w == 0 || !(row[w - 1] >> 31)) return w * 32;
    --w;
    zeros_below = ~row[w];
  }
  return w * 32 + (32 - __clz(zeros_below));
}

// last pixel of the run that starts at the set pixel x
__device__ __forceinline__ int run_end(const unsigned* row, int x, int WW) {
  int w = x >> 5;
  const int b = x & 31;
  unsigned zeros_above = ~row[w] & (0xffffffffu << b);
  while (zeros_above == 0u) {  // the run reaches bit 31 of this word (padding bits are 0, so this ends inside the row)
    if (w == WW - 1 || !(row[w + 1] & 1u)) return w * 32 + 31;
    ++w;
    zeros_above = ~row[w];
  }
  return w * 32 + (__ffs(zeros_above) - 1) - 1;
}

__device__ __forceinline__ int cc_find(const int* L, int a) {
  int l;
  while ((l = *reinterpret_cast<const volatile int*>(L + a)) != a) a = l;
  return a;
}

__device__ __forceinline__ void cc_union(int* L, int a, int b) {
  bool done;
  do {
    a = cc_find(L, a);
    b = cc_find(L, b);
    if (a < b) { const int old = atomicMin(L + b, a); done = (old == b); b = old; }
    else if (b < a) { const int old = atomicMin(L + a, b); done = (old == a); a = old; }
    else done = true;
  } while (!done);
}

__global__ void __launch_bounds__(256) cc_init_kernel(const unsigned* __restrict__ bits, int H, int W, int WW, long long total,
                                                      int* __restrict__ labels, CcStats st) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;  // over K * H * W
  if (idx >= total) return;
  const long long HW = (long long)H * W;
  const int k = (int)(idx / HW);
  const int p = (int)(idx % HW);
  const int y = p / W, x = p % W;
  const unsigned* row = bits + ((long long)k * H + y) * WW;
  int lab = -1;
  if ((row[x >> 5] >> (x & 31)) & 1u) {
    const int xs = run_start(row, x);
    lab = y * W + xs;
    if (xs == x) {
      const long long o = (long long)k * HW + p;
      st.area[o] = 0; st.minx[o] = INT_MAX; st.maxx[o] = -1; st.miny[o] = INT_MAX; st.maxy[o] = -1;
    }
  }
  labels[idx] = lab;
}

__global__ void __launch_bounds__(256) cc_merge_kernel(const unsigned* __restrict__ bits, int H, int W, int WW, long long total,
                                                       int* labels) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const long long HW = (long long)H * W;
  const int k = (int)(idx / HW);
  const int p = (int)(idx % HW);
  const int y = p / W, x = p % W;
  if (y == 0) return;
  const unsigned* row = bits + ((long long)k * H + y) * WW;
  if (!((row[x >> 5] >> (x & 31)) & 1u)) return;
  const unsigned* up = row - WW;
  const bool u0 = bit_at(up, x, W), um = bit_at(up, x - 1, W), up1 = bit_at(up, x + 1, W);
  const bool is_start = !bit_at(row, x - 1, W), is_end = !bit_at(row, x + 1, W);
  int* L = labels + (long long)k * HW;
  // one union per contact between this run and a run of the row above: each upper run is taken at its first pixel
  // inside [run start - 1, run end + 1]
  if (u0 && !um) cc_union(L, L[p], L[p - W]);
  if (is_start && um) cc_union(L, L[p], L[p - W - 1]);
  if (is_end && up1 && !u0) cc_union(L, L[p], L[p - W + 1]);
}

// one thread per run start: the whole run's contribution to its component
__global__ void __launch_bounds__(256) cc_stats_kernel(const unsigned* __restrict__ bits, int H, int W, int WW, long long total,
                                                       const int* __restrict__ labels, CcStats st) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const long long HW = (long long)H * W;
  const int k = (int)(idx / HW);
  const int p = (int)(idx % HW);
  const int y = p / W, x = p % W;
  const unsigned* row = bits + ((long long)k * H + y) * WW;
  if (!((row[x >> 5] >> (x & 31)) & 1u) || bit_at(row, x - 1, W)) return;
  const int xe = run_end(row, x, WW);
  const long long base = (long long)k * HW;
  const long long o = base + cc_find(labels + base, p);
  atomicAdd(st.area + o, xe - x + 1);
  atomicMin(st.minx + o, x); atomicMax(st.maxx + o, xe);
  atomicMin(st.miny + o, y); atomicMax(st.maxy + o, y);
}

// final_mask[labels == i] = 255 for the kept components (mask_cleaner.py:23-34)
__global__ void __launch_bounds__(256) cc_select_kernel(const int* __restrict__ labels, int H, int W, long long total, CcStats st,
                                                        int size_threshold, double aspect_threshold,
                                                        unsigned char* __restrict__ out) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const long long HW = (long long)H * W;
  const long long base = (idx / HW) * HW;
  const int lab = labels[idx];
  unsigned char v = 0;
  if (lab >= 0) {
    const long long o = base + cc_find(labels + base, lab);
    const int width = st.maxx[o] - st.minx[o] + 1, height = st.maxy[o] - st.miny[o] + 1;
    const int mx = width > height ? width : height, mn = width > height ? height : width;
    const double aspect = (double)mx / ((double)mn + 1e-5);
    if (st.area[o] > size_threshold || aspect > aspect_threshold) v = 255;
  }
  out[idx] = v;
}

// ------------------------------------------------------------------------------------------------ pairwise popcount
// inter[i][j] = |bits_i AND bits_j| for j >= i (mirrored); one block per pair, words strided over the block.
__global__ void __launch_bounds__(256) pair_popc_kernel(const unsigned* __restrict__ bits, long long words_per_mask, int K,
                                                        int* __restrict__ inter) {
  const int i = blockIdx.y, j = blockIdx.x;
  if (j < i) return;
  const uint4* a = reinterpret_cast<const uint4*>(bits + (long long)i * words_per_mask);
  const uint4* b = reinterpret_cast<const uint4*>(bits + (long long)j * words_per_mask);
  const long long n4 = words_per_mask >> 2;
  int cnt = 0;
  for (long long t = threadIdx.x; t < n4; t += 256) {
    const uint4 x = a[t], y = b[t];
    cnt += __popc(x.x & y.x) + __popc(x.y & y.y) + __popc(x.z & y.z) + __popc(x.w & y.w);
  }
  for (long long t = (n4 << 2) + threadIdx.x; t < words_per_mask; t += 256)
    cnt += __popc(bits[(long long)i * words_per_mask + t] & bits[(long long)j * words_per_mask + t]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
  __shared__ int red[8];
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = cnt;
  __syncthreads();
  if (threadIdx.x == 0) {
    int s = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) s += red[w];
    inter[(long long)i * K + j] = s;
    inter[(long long)j * K + i] = s;
  }
}

// ------------------------------------------------------------------------------------------------ workspace
void* g_ws = nullptr;
size_t g_ws_bytes = 0;
int g_ws_device = -1;

int refine_workspace(size_t need, void** out) {
  int dev = 0;
  CUDA_TRY(cudaGetDevice(&dev));
  if (g_ws != nullptr && (dev != g_ws_device || need > g_ws_bytes)) {
    CUDA_TRY(cudaDeviceSynchronize());
    cudaFree(g_ws);
    g_ws = nullptr; g_ws_bytes = 0;
  }
  if (g_ws == nullptr) {
    CUDA_TRY(cudaMalloc(&g_ws, need));
    g_ws_bytes = need; g_ws_device = dev;
  }
  *out = g_ws;
  return 0;
}

inline unsigned blocks_for(long long n, int per_block) { return (unsigned)((n + per_block - 1) / per_block); }
inline size_t up256(size_t x) { return (x + 255) & ~size_t(255); }

}  // namespace
}  // namespace inksam

using namespace inksam;

extern "C" {

int inksam_morph_kernel_size(int H, int W) {  // mask_cleaner.py:6-9 calculate_kernel_size(factor = 0.025)
  int k = (int)((H < W ? H : W) * 0.025);
  return (k % 2 != 0) ? k : k + 1;
}

int inksam_clean_masks(const uint8_t* masks_in, int K, int H, int W, int thresh, uint8_t* masks_out, long long* launches,
                       void* stream) {
  if (K < 0 || H < 1 || W < 1) return set_error("inksam_clean_masks: bad shape K=%d H=%d W=%d", K, H, W);
  if (K == 0) return 0;
  if (!masks_in || !masks_out) return set_error("inksam_clean_masks: null argument");
  if ((long long)H * W > INT_MAX) return set_error("inksam_clean_masks: %dx%d exceeds 2^31 pixels", H, W);
  auto st = static_cast<cudaStream_t>(stream);
  const int ksz = inksam_morph_kernel_size(H, W);
  const int r = ksz / 2;
  const int WW = (W + 31) / 32;
  const long long HW = (long long)H * W;
  constexpr int kChunk = 32;  // masks per pass: bounds the label / statistics workspace at 24 bytes per pixel
  const int nc = K < kChunk ? K : kChunk;
  const size_t bits_bytes = up256((size_t)nc * H * WW * 4), lab_bytes = up256((size_t)nc * HW * 4);
  void* ws = nullptr;
  INK_TRY(refine_workspace(2 * bits_bytes + 6 * lab_bytes, &ws));
  char* base = static_cast<char*>(ws);
  unsigned* b0 = reinterpret_cast<unsigned*>(base);
  unsigned* b1 = reinterpret_cast<unsigned*>(base + bits_bytes);
  int* labels = reinterpret_cast<int*>(base + 2 * bits_bytes);
  CcStats cs;
  cs.area = labels + lab_bytes / 4; cs.minx = cs.area + lab_bytes / 4; cs.maxx = cs.minx + lab_bytes / 4;
  cs.miny = cs.maxx + lab_bytes / 4; cs.maxy = cs.miny + lab_bytes / 4;
  long long n_launch = 0;
  for (int k0 = 0; k0 < K; k0 += kChunk) {
    const int n = K - k0 < kChunk ? K - k0 : kChunk;
    const long long words = (long long)n * H * WW, pixels = (long long)n * HW;
    pack_bits_kernel<<<blocks_for(words, 8), 256, 0, st>>>(masks_in + (size_t)k0 * HW, thresh, nullptr, 0, H, W, WW, words, b0);
    // close = dilate then erode (cv2.morphologyEx MORPH_CLOSE), each separable
    morph_h_kernel<false><<<blocks_for(words, 256), 256, 0, st>>>(b0, b1, W, WW, r, words);
    morph_v_kernel<false><<<blocks_for(words, 256), 256, 0, st>>>(b1, b0, H, WW, r, words);
    morph_h_kernel<true><<<blocks_for(words, 256), 256, 0, st>>>(b0, b1, W, WW, r, words);
    morph_v_kernel<true><<<blocks_for(words, 256), 256, 0, st>>>(b1, b0, H, WW, r, words);
    cc_init_kernel<<<blocks_for(pixels, 256), 256, 0, st>>>(b0, H, W, WW, pixels, labels, cs);
    cc_merge_kernel<<<blocks_for(pixels, 256), 256, 0, st>>>(b0, H, W, WW, pixels, labels);
    cc_stats_kernel<<<blocks_for(pixels, 256), 256, 0, st>>>(b0, H, W, WW, pixels, labels, cs);
    cc_select_kernel<<<blocks_for(pixels, 256), 256, 0, st>>>(labels, H, W, pixels, cs, 500, 1.1, masks_out + (size_t)k0 * HW);
    n_launch += 9;
  }
  CUDA_TRY(cudaGetLastError());
  if (launches) *launches += n_launch;
  return 0;
}

int inksam_content_iou(const uint8_t* masks, const uint8_t* sketch_gray, int K, int H, int W, int gray_lt, int32_t* inter_out,
                       long long* launches, void* stream) {
  if (K < 0 || H < 1 || W < 1) return set_error("inksam_content_iou: bad shape K=%d H=%d W=%d", K, H, W);
  if (K == 0) return 0;
  if (!masks || !inter_out) return set_error("inksam_content_iou: null argument");
  auto st = static_cast<cudaStream_t>(stream);
  const int WW = (W + 31) / 32;
  const long long wpm = ((long long)H * WW + 3) & ~3LL;  // words per mask, padded to 16 bytes
  void* ws = nullptr;
  INK_TRY(refine_workspace(up256((size_t)K * wpm * 4), &ws));
  unsigned* bits = static_cast<unsigned*>(ws);
  if (wpm != (long long)H * WW) CUDA_TRY(cudaMemsetAsync(bits, 0, (size_t)K * wpm * 4, st));
  for (int k = 0; k < K; ++k) {  // one pack launch per mask only when rows do not fill the padded stride
    if (wpm == (long long)H * WW && k > 0) break;
    const int n = (wpm == (long long)H * WW) ? K : 1;
    const long long words = (long long)n * H * WW;
    pack_bits_kernel<<<blocks_for(words, 8), 256, 0, st>>>(masks + (size_t)k * H * W, 0, sketch_gray, gray_lt, H, W, WW, words,
                                                            bits + (size_t)k * wpm);
    if (launches) *launches += 1;
  }
  pair_popc_kernel<<<dim3(K, K), 256, 0, st>>>(bits, wpm, K, inter_out);
  if (launches) *launches += 1;
  CUDA_TRY(cudaGetLastError());
  return 0;
}

}  // extern "C"
