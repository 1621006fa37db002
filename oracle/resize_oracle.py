"""TEST INFRASTRUCTURE ONLY (tests/ and __graft_entry__.smoke may import this; the product never does).

CPU restatement of the image resize inside `ResizeLongestSide.apply_image`
(InkLayer/third_party/segment-anything/segment_anything/utils/transforms.py:26-31):

    target_size = get_preprocess_shape(H, W, 1024)                 # transforms.py:93-102
    np.array(resize(to_pil_image(image), target_size))             # torchvision -> PIL.Image.resize(BILINEAR)

The arithmetic lives in a third-party dependency that is NOT vendored under /root/reference: Pillow (the reference pins
no version; this image has Pillow 12.2.0), `src/libImaging/Resample.c`, functions `precompute_coeffs`,
`normalize_coeffs_8bpc`, `ImagingResampleHorizontal_8bpc`, `ImagingResampleVertical_8bpc` and `ImagingResampleInner`.
Its published algorithm for 8-bit images with the bilinear (triangle) filter, restated here:

  * per output index xx of an axis: scale = in / out, filterscale = max(scale, 1), support = 1.0 * filterscale,
    center = (xx + 0.5) * scale, xmin = max(int(center - support + 0.5), 0), xmax = min(int(center + support + 0.5), in),
    weights w[x] = tri((x + xmin - center + 0.5) / filterscale) for x < xmax - xmin, normalised by their sum (double);
  * coefficients become fixed point: int(0.5 + w * 2^22);
  * a pass accumulates 2^21 + sum(pixel * coeff) in int32 and stores clip(acc >> 22, 0, 255) as uint8;
  * horizontal pass first (only over the source rows the vertical pass will read), then the vertical pass on the uint8
    intermediate; a pass whose size does not change is skipped.

PINNED: tests/test_resize_oracle.py checks this restatement bit for bit against Pillow itself (installed in the build
container and on the GPU box) over a sweep of sizes, and against committed fixtures made by tests/golden/make_resize_golden.py.
"""
from __future__ import annotations

from typing import Tuple

import numpy as np

PRECISION_BITS = 32 - 8 - 2


def preprocess_shape(oldh: int, oldw: int, long_side: int = 1024) -> Tuple[int, int]:
    """transforms.py:93-102."""
    scale = long_side * 1.0 / max(oldh, oldw)
    return int(oldh * scale + 0.5), int(oldw * scale + 0.5)


def precompute_coeffs(in_size: int, out_size: int):
    """Resample.c precompute_coeffs + normalize_coeffs_8bpc for the bilinear filter over the whole axis (box = 0..in).
    Returns (bounds int32 [out][2] = (xmin, count), kk int32 [out][ksize])."""
    scale = float(np.float32(in_size) - np.float32(0.0)) / out_size   # the box is float in Pillow
    filterscale = max(scale, 1.0)
    support = 1.0 * filterscale
    ksize = int(np.ceil(support)) * 2 + 1
    ss = 1.0 / filterscale
    bounds = np.zeros((out_size, 2), np.int32)
    kk = np.zeros((out_size, ksize), np.int32)
    for xx in range(out_size):
        center = 0.0 + (xx + 0.5) * scale
        xmin = int(center - support + 0.5)
        if xmin < 0:
            xmin = 0
        xmax = int(center + support + 0.5)
        if xmax > in_size:
            xmax = in_size
        xmax -= xmin
        ww = 0.0
        w = np.zeros(ksize, np.float64)
        for x in range(xmax):
            a = (x + xmin - center + 0.5) * ss
            if a < 0.0:
                a = -a
            w[x] = 1.0 - a if a < 1.0 else 0.0
            ww += w[x]
        for x in range(xmax):
            if ww != 0.0:
                w[x] /= ww
        bounds[xx] = (xmin, xmax)
        kk[xx] = (0.5 + w * (1 << PRECISION_BITS)).astype(np.int64).astype(np.int32)  # all weights are >= 0: truncation
    return bounds, kk


def _pass(img: np.ndarray, bounds: np.ndarray, kk: np.ndarray, axis: int) -> np.ndarray:
    """One resample pass along `axis` (0 = vertical, 1 = horizontal) of a uint8 [H][W][C] image."""
    src = np.moveaxis(img, axis, 0).astype(np.int64)            # [in][other][C]
    out_size, ksize = kk.shape
    acc = np.full((out_size,) + src.shape[1:], 1 << (PRECISION_BITS - 1), np.int64)
    for t in range(ksize):
        idx = np.minimum(bounds[:, 0] + t, src.shape[0] - 1)     # taps beyond `count` carry a zero coefficient
        coeff = np.where(t < bounds[:, 1], kk[:, t], 0).astype(np.int64)
        acc += src[idx] * coeff.reshape((-1,) + (1,) * (src.ndim - 1))
    out = np.clip(acc >> PRECISION_BITS, 0, 255).astype(np.uint8)
    return np.moveaxis(out, 0, axis)


def resize_bilinear_u8(image: np.ndarray, out_h: int, out_w: int) -> np.ndarray:
    """PIL.Image.fromarray(image).resize((out_w, out_h), BILINEAR) for a uint8 [H][W][C] array (ImagingResampleInner)."""
    H, W = image.shape[:2]
    img = np.ascontiguousarray(image)
    need_h, need_v = out_w != W, out_h != H
    if need_v:
        vb, vk = precompute_coeffs(H, out_h)
    if need_h:
        hb, hk = precompute_coeffs(W, out_w)
        if need_v:  # only the source rows the vertical pass reads are resampled horizontally
            first, last = int(vb[0, 0]), int(vb[-1, 0] + vb[-1, 1])
            img = img[first:last]
            vb = vb.copy()
            vb[:, 0] -= first
        img = _pass(img, hb, hk, 1)
    if need_v:
        img = _pass(img, vb, vk, 0)
    return img


def apply_image(image: np.ndarray, long_side: int = 1024) -> np.ndarray:
    """transforms.py:26-31."""
    h, w = preprocess_shape(image.shape[0], image.shape[1], long_side)
    if (h, w) == image.shape[:2]:
        return image
    return resize_bilinear_u8(image, h, w)
