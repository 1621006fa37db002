// Fused mask post-process (HBM-bound): Sam.postprocess_masks + threshold in one pass
// (SA/modeling/sam.py:133-162, SA/predictor.py:238-241):
//   low-res logits [K][256][256] --bilinear--> 1024x1024 --crop [:h,:w]--> --bilinear--> (H,W) --> (> thr) as uint8.
// Both interpolations are align_corners=False without antialiasing; source indices, clamping and the lambda
// arithmetic follow ATen's upsample_bilinear2d (area_pixel_compute_source_index): the 1024x1024 intermediate is never
// materialised, each output pixel evaluates its 4 taps on the virtual 1024 grid, each tap 4 taps on the 256 grid.
//
// One launch covers every box of a decode pass (boxes of several images with different geometries): a 1-D grid of
// tiles, each tile = 16 output rows of one box.  Two code paths per image segment:
//   fast   : original size == resized input size (the second interpolation is the identity) and 16-byte-aligned rows.
//            A thread produces a 4-row x 16-column patch from 3 low-res rows x 6 columns; the horizontal interpolation
//            of a low-res row is shared by the output rows that use it; one 16-byte store per row.
//   general: per-column tap tables (outer tap + its two inner taps) are built once per tile in shared memory; a warp
//            produces 128 consecutive pixels of one row per pass (mask bytes assembled from ballots, one 4-byte store
//            per lane); the two virtual rows share their low-res rows whenever they fall between the same pair.
// Algorithmic bytes per box: 262,144 read + H*W written; H*ceil(W/8) written in the bit-packed format.
#include "common.h"
#include "decoder_ops.h"

namespace inksam {

struct Tap { int i0, i1; float l0, l1; };

__device__ __forceinline__ Tap src_tap(int dst, float scale, int in_size) {
  float r = scale * ((float)dst + 0.5f) - 0.5f;
  if (r < 0.f) r = 0.f;
  Tap t;
  t.i0 = (int)r;
  if (t.i0 > in_size - 1) t.i0 = in_size - 1;
  t.i1 = t.i0 + (t.i0 < in_size - 1 ? 1 : 0);
  t.l1 = fminf(fmaxf(r - (float)t.i0, 0.f), 1.f);
  t.l0 = 1.f - t.l1;
  return t;
}

struct PostSeg {
  const float* low;      // [nbox][256][256]
  unsigned char* masks;  // [nbox][H][W] bytes 0/1, or (packed) [nbox][H][(W+7)/8] bits, or null
  float* logits;         // [nbox][H][W] or null
  int in_h, in_w, H, W;
  int packed, pitch;     // packed: 1 bit per pixel, leftmost pixel in the MSB, rows padded to whole bytes (numpy packbits /
                         // PIL mode "1" row layout, what InkLayer/runner.py:51 converts the masks to); pitch = bytes per row
  int nbox, tile0;       // tile0: index of the segment's first tile in the grid
  float sy, sx;
};
struct PostBatch {
  int nseg, pad;
  PostSeg s[kMaxDecodeImages];
};

constexpr int kRows = 16;      // output rows per tile
constexpr int kChunk = 1024;   // output columns per pass of the general path


// ---------------------------------------------------------------- fast path: 4 rows x 16 columns per thread
__device__ __forceinline__ void post_fast(const PostSeg& sg, int box, int band, float thr) {
  const int W = sg.W, H = sg.H;
  const int segs = W >> 4;                   // 16-pixel segments per row
  const float* lo = sg.low + (size_t)box * 65536;
  unsigned char* out = sg.masks + (size_t)box * H * (sg.packed ? sg.pitch : W);
  // the tile is 16 rows: four 4-row groups; threads sweep (group, segment)
  for (int it = threadIdx.x; it < (kRows / 4) * segs; it += blockDim.x) {
    const int grp = it / segs, s = it - grp * segs;
    const int y0 = band * kRows + grp * 4;   // multiple of 4
    if (y0 >= H) continue;
    const int m = y0 >> 2;
    // low-res rows m-1, m, m+1 (clamped) and columns 4s-1 .. 4s+4 (clamped)
    const int rm = m > 0 ? m - 1 : 0, rp = m < 255 ? m + 1 : 255;
    const int c0 = 4 * s;
    const int cl = c0 > 0 ? c0 - 1 : 0, cr = c0 + 4 < 256 ? c0 + 4 : 255;
    float h[3][16];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const float* row = lo + (r == 0 ? rm : (r == 1 ? m : rp)) * 256;
      const float4 mid = *reinterpret_cast<const float4*>(row + c0);
      const float c[6] = {row[cl], mid.x, mid.y, mid.z, mid.w, row[cr]};
      // X = 16s + 4j + k: k = 0,1 interpolate columns (j-1, j) with lambda .625 / .875; k = 2,3 columns (j, j+1) with
      // .125 / .375  (r = X/4 - 0.375 is exact in fp32, so these are the values src_tap produces)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        h[r][4 * j + 0] = 0.375f * c[j] + 0.625f * c[j + 1];
        h[r][4 * j + 1] = 0.125f * c[j] + 0.875f * c[j + 1];
        h[r][4 * j + 2] = 0.875f * c[j + 1] + 0.125f * c[j + 2];
        h[r][4 * j + 3] = 0.625f * c[j + 1] + 0.375f * c[j + 2];
      }
      if (s == 0) {  // X = 0, 1: the source coordinate clamps to 0 -> lambda (1, 0) on columns (0, 1)
        h[r][0] = 1.f * c[1] + 0.f * c[2];
        h[r][1] = h[r][0];
      }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int y = y0 + k;
      if (y >= H) break;
      const Tap ty = src_tap(y, 0.25f, 256);
      // source rows (m-1, m) for k = 0, 1 and (m, m+1) for k = 2, 3.  Where ATen clamps (y = 0, 1 -> rows (0, 1) with
      // lambda (1, 0); m = 255 -> rows (255, 255)) the clamped h rows hold the same data or carry a zero weight.
      constexpr int kTop[4] = {0, 0, 1, 1};
      const int a = kTop[k], b = kTop[k] + 1;
      unsigned int w[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        unsigned int word = 0;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int x = 4 * q + e;
          const float v = ty.l0 * h[a][x] + ty.l1 * h[b][x];
          word |= (v > thr ? 1u : 0u) << (8 * e);
        }
        w[q] = word;
      }
      if (sg.packed) {  // 16 pixels -> 2 bytes, pixel 0 in bit 7 of the first byte
        unsigned int bits = 0;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const unsigned int nib = (w[q] & 1u) << 3 | ((w[q] >> 8) & 1u) << 2 | ((w[q] >> 16) & 1u) << 1 | ((w[q] >> 24) & 1u);
          bits = (bits << 4) | nib;
        }
        *reinterpret_cast<unsigned short*>(out + (size_t)y * sg.pitch + 2 * s) = (unsigned short)((bits >> 8) | ((bits & 0xffu) << 8));
      } else {
        *reinterpret_cast<uint4*>(out + (size_t)y * W + 16 * s) = make_uint4(w[0], w[1], w[2], w[3]);
      }
    }
  }
}

// ---------------------------------------------------------------- general path
// A warp produces 128 consecutive pixels of one row per pass: lane l evaluates pixels l, l+32, l+64, l+96 (conflict-free
// table reads, coalesced logit loads), the four ballots are re-sliced so that lane l stores the 4 mask bytes 4l..4l+3.
__device__ __forceinline__ void post_general(const PostSeg& sg, int box, int band, float thr, uint2* tab_i, float4* tab_l) {
  const int W = sg.W, H = sg.H, in_h = sg.in_h, in_w = sg.in_w;
  const float* lo = sg.low + (size_t)box * 65536;
  const size_t obase = (size_t)box * H * W;
  const int y_begin = band * kRows;
  const int rows = H - y_begin < kRows ? H - y_begin : kRows;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  for (int xc = 0; xc < W; xc += kChunk) {
    const int cw = W - xc < kChunk ? W - xc : kChunk;
    __syncthreads();
    for (int i = threadIdx.x; i < cw; i += blockDim.x) {
      const Tap tx = src_tap(xc + i, sg.sx, in_w);
      const Tap ta = src_tap(tx.i0, 0.25f, 256), tb = src_tap(tx.i1, 0.25f, 256);
      tab_i[i] = make_uint2((unsigned)ta.i0 | ((unsigned)ta.i1 << 16), (unsigned)tb.i0 | ((unsigned)tb.i1 << 16));
      tab_l[i] = make_float4(ta.l1, tb.l1, tx.l1, 0.f);
    }
    __syncthreads();
    const int groups = (cw + 3) / 128 + 1;  // 128-pixel groups per row, first one shifted left to a 4-byte boundary
    for (int it = warp; it < rows * groups; it += nwarps) {
      const int r = it / groups, g = it - r * groups;
      const int y = y_begin + r;
      const size_t rowo = obase + (size_t)y * W + xc;
      const int mis = (sg.masks && !sg.packed) ? (int)((reinterpret_cast<uintptr_t>(sg.masks) + rowo) & 3) : 0;
      const int xs = 128 * g - mis;  // chunk-relative column of the group's first pixel (negative in the head group)
      if (xs >= cw) continue;
      const Tap ty = src_tap(y, sg.sy, in_h);
      const Tap ya = src_tap(ty.i0, 0.25f, 256), yb = src_tap(ty.i1, 0.25f, 256);
      const float* ra0 = lo + ya.i0 * 256; const float* ra1 = lo + ya.i1 * 256;
      const float* rb0 = lo + yb.i0 * 256; const float* rb1 = lo + yb.i1 * 256;
      const bool same = (ya.i0 == yb.i0) && (ya.i1 == yb.i1);
      unsigned int bal[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int x = xs + 32 * e + lane;
        const bool ok = x >= 0 && x < cw;
        float v = 0.f;
        if (ok) {
          const uint2 ci = tab_i[x];
          const float4 cl = tab_l[x];
          const int pa = ci.x & 0xffff, qa = ci.x >> 16, pb = ci.y & 0xffff, qb = ci.y >> 16;
          const float la0 = 1.f - cl.x, lb0 = 1.f - cl.y, l0 = 1.f - cl.z;
          // horizontal interpolation of the low-res rows at the two virtual columns
          const float a0A = la0 * ra0[pa] + cl.x * ra0[qa], a1A = la0 * ra1[pa] + cl.x * ra1[qa];
          const float a0B = lb0 * ra0[pb] + cl.y * ra0[qb], a1B = lb0 * ra1[pb] + cl.y * ra1[qb];
          float b0A = a0A, b1A = a1A, b0B = a0B, b1B = a1B;
          if (!same) {
            b0A = la0 * rb0[pa] + cl.x * rb0[qa]; b1A = la0 * rb1[pa] + cl.x * rb1[qa];
            b0B = lb0 * rb0[pb] + cl.y * rb0[qb]; b1B = lb0 * rb1[pb] + cl.y * rb1[qb];
          }
          // virtual 1024-grid values at (Y0, X0), (Y0, X1), (Y1, X0), (Y1, X1)
          const float u00 = ya.l0 * a0A + ya.l1 * a1A, u01 = ya.l0 * a0B + ya.l1 * a1B;
          const float u10 = yb.l0 * b0A + yb.l1 * b1A, u11 = yb.l0 * b0B + yb.l1 * b1B;
          v = ty.l0 * (l0 * u00 + cl.z * u01) + ty.l1 * (l0 * u10 + cl.z * u11);
          if (sg.logits) sg.logits[rowo + x] = v;
        }
        bal[e] = __ballot_sync(0xffffffffu, ok && v > thr);
      }
      if (sg.masks && sg.packed) {
        // ballot e holds pixels xs + 32e .. +31 (bit = lane): reversed and byte-swapped it is the 4 packed bytes of those
        // pixels in memory order; lane e < 4 stores them (pixels beyond the row voted 0: the padding bits are zero)
        if (lane < 4) {
          const int px = xs + 32 * lane;  // xs = 128 g here (mis == 0)
          if (px < cw) {
            const unsigned int word = __byte_perm(__brev(bal[lane]), 0, 0x0123);
            unsigned char* o = sg.masks + ((size_t)box * H + y) * sg.pitch + ((xc + px) >> 3);
            const int nb = (cw - px + 7) >> 3;  // bytes of this row that remain from px on
            if (nb >= 4 && (reinterpret_cast<uintptr_t>(o) & 3) == 0) {
              *reinterpret_cast<unsigned int*>(o) = word;
            } else {
#pragma unroll
              for (int e = 0; e < 4; ++e)
                if (e < nb) o[e] = (unsigned char)((word >> (8 * e)) & 0xffu);
            }
          }
        }
      } else if (sg.masks) {
        // lane l owns pixels xs + 4l .. xs + 4l + 3: bits (4l & 31) .. +3 of ballot word l / 8
        const unsigned int src = (lane < 8) ? bal[0] : (lane < 16) ? bal[1] : (lane < 24) ? bal[2] : bal[3];
        const unsigned int bits = (src >> ((4 * lane) & 31)) & 0xfu;
        const unsigned int word = (bits * 0x00204081u) & 0x01010101u;
        const int x = xs + 4 * lane;
        unsigned char* o = sg.masks + rowo + x;
        if (x >= 0 && x + 4 <= cw) {
          *reinterpret_cast<unsigned int*>(o) = word;
        } else {
#pragma unroll
          for (int e = 0; e < 4; ++e)
            if (x + e >= 0 && x + e < cw) o[e] = (unsigned char)((word >> (8 * e)) & 1u);
        }
      }
    }
  }
}

__global__ void __launch_bounds__(256) postprocess_kernel(const __grid_constant__ PostBatch pb, float thr) {
  // per-column taps of the general path: inner tap indices of X0 / X1 (packed 16-bit pairs) and (lambda X0, lambda X1, outer lambda)
  __shared__ uint2 tab_i[kChunk];
  __shared__ float4 tab_l[kChunk];
  // segment of this tile (tile0 ascending)
  int si = 0;
  while (si + 1 < pb.nseg && (int)blockIdx.x >= pb.s[si + 1].tile0) ++si;
  const PostSeg& sg = pb.s[si];
  const int bands = (sg.H + kRows - 1) / kRows;
  const int t = blockIdx.x - sg.tile0;
  const int box = t / bands, band = t - box * bands;
  const bool fast = sg.in_h == sg.H && sg.in_w == sg.W && (sg.W & 15) == 0 && sg.logits == nullptr && sg.masks != nullptr &&
                    (reinterpret_cast<uintptr_t>(sg.masks) & 15) == 0;  // (packed rows of W % 16 == 0 pixels are 2-byte aligned)
  if (fast) post_fast(sg, box, band, thr);
  else post_general(sg, box, band, thr, tab_i, tab_l);
}

static int add_segment(PostBatch& pb, int& tiles, const float* low, unsigned char* masks, float* logits, int K, int in_h,
                       int in_w, int H, int W, int packed) {
  if (in_h < 1 || in_w < 1 || in_h > 1024 || in_w > 1024 || H < 1 || W < 1)
    return set_error("postprocess: bad sizes input %dx%d original %dx%d", in_h, in_w, H, W);
  PostSeg& s = pb.s[pb.nseg++];
  s.low = low; s.masks = masks; s.logits = logits; s.in_h = in_h; s.in_w = in_w; s.H = H; s.W = W; s.nbox = K;
  s.packed = packed ? 1 : 0; s.pitch = (W + 7) / 8;
  s.sy = (float)in_h / (float)H; s.sx = (float)in_w / (float)W;
  s.tile0 = tiles;
  tiles += K * ((H + kRows - 1) / kRows);
  return 0;
}

// low: [K][256][256] fp32; masks: [K][H][W] uint8 (0/1) or null; logits: [K][H][W] fp32 or null.
int launch_postprocess(const float* low, unsigned char* masks, float* logits, int K, int in_h, int in_w, int H, int W,
                       float thr, cudaStream_t st, int packed) {
  if (K <= 0) return 0;
  PostBatch pb{};
  int tiles = 0;
  INK_TRY(add_segment(pb, tiles, low, masks, logits, K, in_h, in_w, H, W, packed));
  postprocess_kernel<<<tiles, 256, 0, st>>>(pb, thr);
  CUDA_TRY(cudaGetLastError());
  return 0;
}

// Several images' boxes in one launch: segment i has n[i] boxes whose logits start at low[i], output at masks[i] / logits[i].
int launch_postprocess_batch(int nseg, const float* const* low, unsigned char* const* masks, float* const* logits, const int* n,
                             const int* in_hw, const int* orig_hw, float thr, cudaStream_t st, int packed) {
  if (nseg > kMaxDecodeImages) return set_error("postprocess: %d segments > %d", nseg, kMaxDecodeImages);
  PostBatch pb{};
  int tiles = 0;
  for (int i = 0; i < nseg; ++i) {
    if (n[i] <= 0) continue;
    INK_TRY(add_segment(pb, tiles, low[i], masks ? masks[i] : nullptr, logits ? logits[i] : nullptr, n[i], in_hw[2 * i],
                        in_hw[2 * i + 1], orig_hw[2 * i], orig_hw[2 * i + 1], packed));
  }
  if (tiles == 0) return 0;
  postprocess_kernel<<<tiles, 256, 0, st>>>(pb, thr);
  CUDA_TRY(cudaGetLastError());
  return 0;
}

}  // namespace inksam
