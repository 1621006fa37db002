"""CPU oracle for the mask refinement rows next to the SAM hot path -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

A numpy / scipy.ndimage restatement of the integer arithmetic in
  InkLayer/refinement/mask_cleaner.py:6-36      (calculate_kernel_size, clean_up_mask)
  InkLayer/refinement/nms_sketch.py:62-78       (refine_mask_to_sketch_regions)
  InkLayer/refinement/nms_sketch.py:186-262     (content_iou: the intersection / union sums and the box rules)
  InkLayer/refinement/nms_sketch.py:283-360     (sketch_nms greedy loop), InkLayer/refinement/utils.py:11-33
without OpenCV (the reference calls cv2.threshold / morphologyEx / connectedComponentsWithStats).  Only tests/ may
import it.  Pinned by tests/golden/refine_*.npz, produced by running the reference's own functions in the build container
(tests/golden/make_refine_golden.py); tests/test_refine_oracle.py checks this file against them bit for bit.
"""
from __future__ import annotations

import numpy as np
from scipy import ndimage


def calculate_kernel_size(image_shape, factor: float = 0.025):
    """mask_cleaner.py:6-9."""
    k = int(min(image_shape) * factor)
    k = k if k % 2 != 0 else k + 1
    return (k, k)


def morph_close(binary: np.ndarray, k: int) -> np.ndarray:
    """cv2.morphologyEx(MORPH_CLOSE, rect k x k): dilate then erode; pixels outside the image do not take part
    (cv2's morphologyDefaultBorderValue), i.e. the window is clipped to the image."""
    dil = ndimage.maximum_filter(binary.astype(np.uint8), size=(k, k), mode="constant", cval=0)
    ero = ndimage.minimum_filter(dil, size=(k, k), mode="constant", cval=1)
    return ero.astype(bool)


def clean_up_mask(mask: np.ndarray, thresh: int = 127, size_threshold: int = 500, aspect_ratio_threshold: float = 1.1) -> np.ndarray:
    """mask_cleaner.py:11-36.  mask: uint8 [H,W]; returns uint8 0/255."""
    binary = mask > thresh                                                   # :12 cv2.threshold(127, 255, THRESH_BINARY)
    k = calculate_kernel_size(binary.shape)[0]                               # :13
    closed = morph_close(binary, k)                                          # :14-15
    labels, n = ndimage.label(closed, structure=np.ones((3, 3), dtype=int))  # :17-19, connectivity 8
    out = np.zeros(mask.shape, np.uint8)
    if n == 0:
        return out
    areas = np.bincount(labels.ravel(), minlength=n + 1)
    for i, sl in enumerate(ndimage.find_objects(labels), start=1):           # :25-34
        height, width = sl[0].stop - sl[0].start, sl[1].stop - sl[1].start
        aspect = max(width, height) / (min(width, height) + 1e-5)
        if areas[i] > size_threshold or aspect > aspect_ratio_threshold:
            out[labels == i] = 255
    return out


def refine_mask_to_sketch_regions(sketch_gray: np.ndarray, pred_mask: np.ndarray) -> np.ndarray:
    """nms_sketch.py:62-78 with the sketch already grey and at mask size (the reference resizes with PIL bilinear, the
    identity when the sizes agree, which they do on the pipeline's own files)."""
    return np.logical_and(pred_mask.astype(bool), sketch_gray < 250)


def content_iou_matrix(masks: np.ndarray, sketch_gray) -> np.ndarray:
    """inter[i][j] = |refined_i & refined_j| (nms_sketch.py:226-236); diagonal = refined areas."""
    K = masks.shape[0]
    refined = masks.astype(bool) if sketch_gray is None else np.stack([refine_mask_to_sketch_regions(sketch_gray, m) for m in masks])
    flat = refined.reshape(K, -1).astype(np.int64)
    return (flat @ flat.T).astype(np.int32)


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


def is_contained_bbox(small, big, epsilon):
    """nms_sketch.py:176-183."""
    return small[0] >= big[0] - epsilon and small[1] >= big[1] - epsilon and small[2] <= big[2] + epsilon and small[3] <= big[3] + epsilon


def share_corner(b1, b2, epsilon):
    """nms_sketch.py:24-59."""
    c1 = [(b1[0], b1[1]), (b1[0], b1[3]), (b1[2], b1[1]), (b1[2], b1[3])]
    c2 = [(b2[0], b2[1]), (b2[0], b2[3]), (b2[2], b2[1]), (b2[2], b2[3])]
    return any(((p[0] - q[0]) ** 2 + (p[1] - q[1]) ** 2) ** 0.5 <= epsilon for p in c1 for q in c2)


def pair_scores(inter: np.ndarray, bboxes: np.ndarray, scores: np.ndarray, i: int, j: int, h: int, w: int):
    """content_iou for the pair (i, j) given the intersection matrix: nms_sketch.py:186-262."""
    box1, box2 = bboxes[i].astype(float), bboxes[j].astype(float)
    if np.all(box1 <= 1.0) and np.all(box2 <= 1.0):
        box1, box2 = box1 * np.array([w, h, w, h]), box2 * np.array([w, h, w, h])
    area1 = (box1[2] - box1[0]) * (box1[3] - box1[1])
    area2 = (box2[2] - box2[0]) * (box2[3] - box2[1])
    if area1 > area2:
        larger_box, smaller_box, larger_index, smaller_index = box1, box2, i, j
    else:
        larger_box, smaller_box, larger_index, smaller_index = box2, box1, j, i
    inter_sum = int(inter[i, j])
    union_sum = int(inter[i, i]) + int(inter[j, j]) - inter_sum
    sketch_iou = inter_sum / union_sum if union_sum > 0 else 0.0
    eps = 8.0 * (np.sqrt(w ** 2 + h ** 2) / 1000)                             # get_dynamic_threshold, :7-21
    if not is_contained_bbox(smaller_box, larger_box, eps) or not share_corner(smaller_box, larger_box, eps):
        return 0.0, 0.0, larger_index
    better = larger_index if scores[larger_index] > scores[smaller_index] else smaller_index
    return sketch_iou, compute_bbox_iou(smaller_box, larger_box), better


def sketch_nms_from_matrix(inter: np.ndarray, bboxes: np.ndarray, scores: np.ndarray, kept: np.ndarray, h: int, w: int,
                           sketch_iou_threshold: float, bbox_iou_threshold: float = 0.7) -> np.ndarray:
    """The greedy loop of sketch_nms (nms_sketch.py:283-360) after filter_full_or_empty_bbox has produced `kept`
    (indices into bboxes); `inter` is indexed by ORIGINAL mask index (mask_{i}.png).

    Reference quirk, reproduced: sketch_nms hands content_iou the FILTERED boxes and FILTERED indices
    (nms_sketch.py:295-296,311-318), and content_iou opens `mask_{box_index}.png` with that filtered index
    (nms_sketch.py:209-221) while runner.py:57-60 wrote the files under ORIGINAL indices.  So once the filter has dropped
    a box, filtered box f is compared through the mask file of original box f, not kept[f]."""
    if len(kept) == 0:
        return np.array([])
    fb, fs = bboxes[kept], scores[kept]
    order = np.argsort(-fs)
    original = kept[order]
    n = len(fb)
    keep = np.ones(n, dtype=bool)
    sub = inter[:n, :n]  # mask_{f}.png for filtered index f
    for i in range(n):
        if not keep[i]:
            continue
        remaining = order[i + 1:]
        if len(remaining) == 0:
            continue
        res = [pair_scores(sub, fb, fs, order[i], r, h, w) for r in remaining]
        ious_s = np.array([r[0] for r in res]); ious_b = np.array([r[1] for r in res]); larger = np.array([r[2] for r in res])
        overlapped = np.where(np.logical_or(ious_s > sketch_iou_threshold, ious_b > bbox_iou_threshold))[0]
        for ov in overlapped:
            compared = remaining[ov]
            if larger[ov] == compared:
                keep[i] = False
                break
            keep[np.where(order == compared)[0][0]] = False
    return original[keep]


def filter_full_or_empty_bbox(sketch_gray: np.ndarray, bboxes: np.ndarray, size_threshold: float = 0.9, max_contained_boxes: int = 5):
    """nms_sketch.py:128-173 with the sketch already loaded as grey uint8."""
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
        contained = sum(1 for other in bboxes if not np.array_equal(box, other)
                        and box[0] <= other[0] and box[1] <= other[1] and box[2] >= other[2] and box[3] >= other[3])
        if valid_area and valid_content and contained <= max_contained_boxes:
            kept.append(i)
    return np.array(kept)


def sketch_nms(sketch_gray: np.ndarray, bboxes: np.ndarray, scores: np.ndarray, masks: np.ndarray, sketch_iou_threshold: float,
               bbox_iou_threshold: float = 0.7) -> np.ndarray:
    """nms_sketch.py:283-360 end to end on in-memory arrays (masks: [K,H,W] as saved in masks/mask_{i}.png)."""
    if len(bboxes) == 0:
        return np.array([])
    kept = filter_full_or_empty_bbox(sketch_gray, bboxes)
    if len(kept) == 0:
        return np.array([])
    h, w = sketch_gray.shape
    inter = content_iou_matrix(masks, sketch_gray)
    return sketch_nms_from_matrix(inter, bboxes, scores, kept, h, w, sketch_iou_threshold, bbox_iou_threshold)
