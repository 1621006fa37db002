"""Seeded synthetic inputs: line-drawing "sketches" and GroundingDINO-like boxes.

The detector stage of the reference cannot run here (no weights, missing deps), so the hot path is fed
white canvases with random black strokes and uniformly drawn boxes, all from numpy Generators seeded by
the image index, identical on every machine.
"""
from __future__ import annotations

import numpy as np


def synthetic_sketch(index: int, height: int = 1024, width: int = 1024) -> np.ndarray:
    """uint8 [H,W,3] white canvas with 40-120 black polylines / ellipses, thickness 1-4 px."""
    import cv2

    rng = np.random.default_rng(1000 + index)
    img = np.full((height, width, 3), 255, np.uint8)
    for _ in range(int(rng.integers(40, 121))):
        t = int(rng.integers(1, 5))
        if rng.random() < 0.5:
            n = int(rng.integers(2, 7))
            pts = np.stack([rng.integers(0, width, n), rng.integers(0, height, n)], axis=1).astype(np.int32)
            cv2.polylines(img, [pts.reshape(-1, 1, 2)], bool(rng.random() < 0.3), (0, 0, 0), t, cv2.LINE_AA)
        else:
            c = (int(rng.integers(0, width)), int(rng.integers(0, height)))
            ax = (int(rng.integers(8, width // 4)), int(rng.integers(8, height // 4)))
            cv2.ellipse(img, c, ax, float(rng.uniform(0, 180)), 0, 360, (0, 0, 0), t, cv2.LINE_AA)
    return img


def synthetic_boxes(index: int, count: int, height: int = 1024, width: int = 1024) -> np.ndarray:
    """float32 [count,4] pixel xyxy boxes in the ORIGINAL image frame."""
    rng = np.random.default_rng(3000 + index)
    x0 = rng.uniform(0, 0.7 * width, count)
    y0 = rng.uniform(0, 0.7 * height, count)
    w = rng.uniform(16, 0.3 * width, count)
    h = rng.uniform(16, 0.3 * height, count)
    b = np.stack([x0, y0, np.minimum(x0 + w, width - 1), np.minimum(y0 + h, height - 1)], axis=1)
    return b.astype(np.float32)


def refine_case(seed: int, K: int, H: int, W: int):
    """Seeded inputs for the refinement rows: K noisy blob masks (uint8 0/255, with specks, holes and thin streaks so that
    clean_up_mask's close / size / aspect rules all fire), a grey sketch with strokes, nested boxes and scores."""
    from scipy import ndimage

    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:H, 0:W]
    masks, boxes = [], []
    for k in range(K):
        cx, cy = rng.uniform(0.2, 0.8) * W, rng.uniform(0.2, 0.8) * H
        rx, ry = rng.uniform(0.08, 0.3) * W, rng.uniform(0.08, 0.3) * H
        if k % 3 == 1 and k > 0:  # nested in the previous box, sharing its top-left corner (the NMS merge rule)
            x0, y0, x1, y1 = boxes[-1]
            rx, ry = (x1 - x0) * 0.3, (y1 - y0) * 0.3
            cx, cy = x0 + rx, y0 + ry
        field = ndimage.gaussian_filter(rng.standard_normal((H, W)), 4) * 3
        m = ((xx - cx) / rx) ** 2 + ((yy - cy) / ry) ** 2 + field < 1.0
        m &= rng.random((H, W)) > 0.03                      # pin holes, closed by the morphology
        specks = rng.random((H, W)) > 0.9995                 # isolated pixels far from the blob
        m |= ndimage.binary_dilation(specks, iterations=int(rng.integers(1, 4)))
        y = int(rng.integers(0, H)); x = int(rng.integers(0, max(W - 40, 1)))
        m[y:y + 2, x:x + 40] = True                          # a thin streak: small but far from square
        masks.append(m)
        boxes.append([max(cx - rx, 0), max(cy - ry, 0), min(cx + rx, W - 1), min(cy + ry, H - 1)])
    masks = (np.stack(masks) * 255).astype(np.uint8)
    sketch = np.full((H, W), 255, np.uint8)
    strokes = ndimage.gaussian_filter(rng.standard_normal((H, W)), 3)
    sketch[np.abs(strokes) < 0.01] = 0
    sketch = ndimage.uniform_filter(sketch.astype(np.float32), 2).astype(np.uint8)  # anti-aliased edges around 250
    bboxes = np.array(boxes, dtype=np.float64).round()
    scores = rng.uniform(0.3, 0.9, K)
    return masks, sketch, bboxes, scores
