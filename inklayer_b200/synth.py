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
