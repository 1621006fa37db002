"""The first consumers of the SAM masks, on the GPU: `clean_up_mask` of
InkLayer/refinement/mask_cleaner.py and the mask arithmetic + greedy loop of `sketch_nms` in
InkLayer/refinement/nms_sketch.py.  Same function names, argument meaning and return values as the reference, but masks
come as arrays / device tensors instead of PNG files re-read from disk for every pair (nms_sketch.py:186-224 reloads the
sketch and two masks per comparison).  The data-parallel parts (threshold, morphological close, connected components,
stroke-restricted pairwise intersections) are CUDA kernels behind include/inksam_refine.h; the box rules and the greedy
loop are host logic on the K x K result.  There is no CPU path for the kernels."""
from __future__ import annotations

import ctypes
from typing import Optional, Sequence, Union

import numpy as np
import torch

from . import _lib

Array = Union[np.ndarray, torch.Tensor]
launch_count = ctypes.c_longlong(0)  # kernels launched through this module (bench.py's gpu_launches)


def _dev(t: Array, device) -> torch.Tensor:
    if isinstance(t, np.ndarray):
        t = torch.from_numpy(np.ascontiguousarray(t))
    if t.dtype == torch.bool:
        t = t.view(torch.uint8)
    if t.dtype != torch.uint8:
        raise RuntimeError(f"masks must be uint8 or bool, got {t.dtype}")
    return t.to(device).contiguous()


def _stream(t: torch.Tensor) -> ctypes.c_void_p:
    return ctypes.c_void_p(torch.cuda.current_stream(t.device).cuda_stream)


def calculate_kernel_size(image_shape, factor: float = 0.025):
    """mask_cleaner.py:6-9."""
    k = int(min(image_shape) * factor)
    k = k if k % 2 != 0 else k + 1
    return (k, k)


def clean_masks(masks: Array, thresh: int = 127, device="cuda") -> torch.Tensor:
    """clean_up_mask for a stack of masks [K,H,W] (uint8; foreground where value > thresh; pass thresh=0 for the bool
    masks predict_torch returns).  Returns a uint8 device tensor [K,H,W] of 0 / 255 (mask_cleaner.py:11-36)."""
    m = _dev(masks, device)
    if m.dim() != 3:
        raise RuntimeError(f"clean_masks expects [K,H,W], got {tuple(m.shape)}")
    K, H, W = m.shape
    out = torch.empty_like(m)
    with torch.cuda.device(m.device):
        _lib.check(_lib.load().inksam_clean_masks(ctypes.c_void_p(m.data_ptr()), K, H, W, int(thresh),
                                                  ctypes.c_void_p(out.data_ptr()), ctypes.byref(launch_count), _stream(m)))
    return out


def clean_up_mask(binary_mask: np.ndarray) -> np.ndarray:
    """Drop-in for mask_cleaner.py:11 on one host mask (uint8 0/255 as read from mask_{i}.png)."""
    return clean_masks(binary_mask[None], 127).cpu().numpy()[0]


def content_iou_matrix(masks: Array, sketch_gray: Optional[Array], device="cuda") -> torch.Tensor:
    """inter[i][j] = number of stroke pixels (sketch grey < 250) shared by masks i and j; the diagonal holds each mask's
    stroke-pixel count, so union = inter[i,i] + inter[j,j] - inter[i,j] (nms_sketch.py:62-78,226-236).  int32 [K,K]."""
    m = _dev(masks, device)
    K, H, W = m.shape
    g = None
    if sketch_gray is not None:
        g = _dev(sketch_gray, m.device)
        if tuple(g.shape) != (H, W):
            raise RuntimeError(f"sketch {tuple(g.shape)} and masks {(H, W)} differ in size")
    inter = torch.empty((K, K), dtype=torch.int32, device=m.device)
    with torch.cuda.device(m.device):
        _lib.check(_lib.load().inksam_content_iou(ctypes.c_void_p(m.data_ptr()), ctypes.c_void_p(0 if g is None else g.data_ptr()),
                                                  K, H, W, 250, ctypes.c_void_p(inter.data_ptr()), ctypes.byref(launch_count),
                                                  _stream(m)))
    return inter


# ------------------------------------------------------------------------------------------------ host box rules
def compute_bbox_iou(box1, box2) -> float:
    """refinement/utils.py:11-33."""
    x1_i, y1_i = max(box1[0], box2[0]), max(box1[1], box2[1])
    x2_i, y2_i = min(box1[2], box2[2]), min(box1[3], box2[3])
    area_1 = (box1[2] - box1[0]) * (box1[3] - box1[1])
    area_2 = (box2[2] - box2[0]) * (box2[3] - box2[1])
    if x2_i < x1_i or y2_i < y1_i:
        return 0.0
    area_i = (x2_i - x1_i) * (y2_i - y1_i)
    return area_i / (area_1 + area_2 - area_i)


def get_dynamic_threshold(h: int, w: int, base_threshold: float = 8.0) -> float:
    """nms_sketch.py:7-21."""
    return base_threshold * (np.sqrt(w ** 2 + h ** 2) / 1000)


def is_contained_bbox(small_box, big_box, epsilon=5.0) -> bool:
    """nms_sketch.py:176-183."""
    return bool(small_box[0] >= big_box[0] - epsilon and small_box[1] >= big_box[1] - epsilon
                and small_box[2] <= big_box[2] + epsilon and small_box[3] <= big_box[3] + epsilon)


def share_corner(box1, box2, epsilon=5.0) -> bool:
    """nms_sketch.py:24-59."""
    c1 = [(box1[0], box1[1]), (box1[0], box1[3]), (box1[2], box1[1]), (box1[2], box1[3])]
    c2 = [(box2[0], box2[1]), (box2[0], box2[3]), (box2[2], box2[1]), (box2[2], box2[3])]
    for p in c1:
        for q in c2:
            if ((p[0] - q[0]) ** 2 + (p[1] - q[1]) ** 2) ** 0.5 <= epsilon:
                return True
    return False


def filter_full_or_empty_bbox(sketch_gray: np.ndarray, bboxes: np.ndarray, size_threshold=0.9, max_contained_boxes=5) -> np.ndarray:
    """nms_sketch.py:128-173 (sketch already loaded as grey uint8)."""
    img_height, img_width = sketch_gray.shape
    img_area = img_height * img_width
    if np.max(bboxes) <= 1.0:
        bboxes = (bboxes * np.array([img_width, img_height, img_width, img_height])).astype(int)
    kept = []
    for i, box in enumerate(bboxes):
        valid_area = (box[2] - box[0]) * (box[3] - box[1]) / img_area < size_threshold
        x_min, y_min, x_max, y_max = box
        x_min = int(max(0, min(x_min, img_width - 1))); y_min = int(max(0, min(y_min, img_height - 1)))
        x_max = int(max(0, min(x_max, img_width - 1))); y_max = int(max(0, min(y_max, img_height - 1)))
        valid_content = np.count_nonzero(sketch_gray[y_min:y_max + 1, x_min:x_max + 1]) > 0
        contained = 0
        for other in bboxes:
            if np.array_equal(box, other):
                continue
            if box[0] <= other[0] and box[1] <= other[1] and box[2] >= other[2] and box[3] >= other[3]:
                contained += 1
        if valid_area and valid_content and contained <= max_contained_boxes:
            kept.append(i)
    return np.array(kept)


def content_iou(inter: np.ndarray, bboxes: np.ndarray, scores: Sequence[float], box1_index: int, box2_index: int, h: int, w: int):
    """nms_sketch.py:186-262 for one pair, reading the intersection counts from the K x K matrix instead of two PNGs."""
    box1, box2 = bboxes[box1_index].astype(float), bboxes[box2_index].astype(float)
    if np.all(box1 <= 1.0) and np.all(box2 <= 1.0):
        box1, box2 = box1 * np.array([w, h, w, h]), box2 * np.array([w, h, w, h])
    area1 = (box1[2] - box1[0]) * (box1[3] - box1[1])
    area2 = (box2[2] - box2[0]) * (box2[3] - box2[1])
    if area1 > area2:
        larger_box, smaller_box, larger_index, smaller_index = box1, box2, box1_index, box2_index
    else:
        larger_box, smaller_box, larger_index, smaller_index = box2, box1, box2_index, box1_index
    intersection_sum = int(inter[box1_index, box2_index])
    union_sum = int(inter[box1_index, box1_index]) + int(inter[box2_index, box2_index]) - intersection_sum
    sketch_iou = intersection_sum / union_sum if union_sum > 0 else 0.0
    eps = get_dynamic_threshold(h, w)
    if not is_contained_bbox(smaller_box, larger_box, eps) or not share_corner(smaller_box, larger_box, eps):
        return 0.0, 0.0, larger_index
    better_idx = larger_index if scores[larger_index] > scores[smaller_index] else smaller_index
    return sketch_iou, compute_bbox_iou(smaller_box, larger_box), better_idx


def sketch_nms(sketch_gray: np.ndarray, bboxes: np.ndarray, scores: np.ndarray, masks: Array, sketch_iou_threshold: float,
               bbox_iou_threshold: float = 0.7, device="cuda") -> np.ndarray:
    """nms_sketch.py:283-360.  sketch_gray: uint8 [H,W]; masks: [K,H,W] (what the reference keeps in masks/mask_{i}.png,
    host or device).  Returns the kept ORIGINAL box indices in the reference's order.

    Reference quirk, reproduced for drop-in parity: the reference calls content_iou with FILTERED indices and content_iou
    opens `mask_{filtered index}.png` (nms_sketch.py:209-221,311-318) although the files were written under ORIGINAL
    indices (runner.py:57-60); after filter_full_or_empty_bbox has dropped a box, filtered box f is therefore compared
    through the mask of original box f, not kept[f].  tests/golden/refine_nms_drop_mid.npz pins this."""
    if len(bboxes) == 0:
        return np.array([])
    kept = filter_full_or_empty_bbox(sketch_gray, bboxes)
    if len(kept) == 0:
        return np.array([])
    h, w = sketch_gray.shape
    inter_all = content_iou_matrix(masks, sketch_gray, device).cpu().numpy()  # one K x K int32 D2H for all pairs
    fb, fs = bboxes[kept], scores[kept]
    inter = inter_all[:len(kept), :len(kept)]  # mask_{f}.png for filtered index f (see the docstring)
    order = np.argsort(-fs)
    original_indices = kept[order]
    keep = np.ones(len(fb), dtype=bool)
    for i in range(len(fb)):
        if not keep[i]:
            continue
        remaining = order[i + 1:]
        if len(remaining) == 0:
            continue
        res = [content_iou(inter, fb, fs, order[i], r, h, w) for r in remaining]
        ious_sketch = np.array([r[0] for r in res]); ious_bbox = np.array([r[1] for r in res])
        larger = np.array([r[2] for r in res])
        for ov in np.where(np.logical_or(ious_sketch > sketch_iou_threshold, ious_bbox > bbox_iou_threshold))[0]:
            compared = remaining[ov]
            if larger[ov] == compared:
                keep[i] = False
                break
            keep[np.where(order == compared)[0][0]] = False
    return original_indices[keep]
