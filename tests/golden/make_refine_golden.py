"""Generates tests/golden/refine_*.npz by running the UNMODIFIED reference refinement functions
(/root/reference/InkLayer/refinement/mask_cleaner.py, nms_sketch.py) on seeded synthetic masks.  Runs only in the build
container; the vectors pin oracle/refine_oracle.py (tests/test_refine_oracle.py) and, on the GPU box, the CUDA kernels.

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_refine_golden.py
"""
import contextlib
import io
import os
import sys
import tempfile

import cv2
import numpy as np
from PIL import Image
from scipy import ndimage

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference")

from InkLayer.refinement.mask_cleaner import clean_up_mask  # noqa: E402
from InkLayer.refinement.nms_sketch import content_iou, sketch_nms  # noqa: E402

from inklayer_b200.synth import refine_case  # noqa: E402


def pack(a):
    return np.packbits(np.asarray(a).astype(bool).reshape(-1))


def make_clean(name, H, W, K, seed):
    masks = refine_case(seed, K, H, W)[0]
    out = np.stack([clean_up_mask(m) for m in masks])
    np.savez_compressed(os.path.join(HERE, f"refine_clean_{name}.npz"), shape=np.array(masks.shape), seed=seed,
                        in_bits=pack(masks > 127), out_bits=pack(out > 0), out_values=np.unique(out))
    print(name, masks.shape, "kept fraction", (out > 0).sum() / max((masks > 127).sum(), 1))


def make_nms(name, H, W, K, seed, thr, full_box_at=None):
    masks, sketch, bboxes, scores = refine_case(seed, K, H, W)
    if full_box_at is not None:  # a box covering the whole sketch in the MIDDLE of the list: dropped by
        bboxes[full_box_at] = [0, 0, W - 1, H - 1]  # filter_full_or_empty_bbox, so filtered and original indices diverge
    with tempfile.TemporaryDirectory() as d:
        sp = os.path.join(d, "input.png")
        Image.fromarray(sketch).save(sp)
        md = os.path.join(d, "masks")
        os.makedirs(md)
        for i, m in enumerate(masks):
            cv2.imwrite(os.path.join(md, f"mask_{i}.png"), m)
        with contextlib.redirect_stdout(io.StringIO()), contextlib.redirect_stderr(io.StringIO()):
            keep = sketch_nms(sp, bboxes.copy(), scores.copy(), md, sketch_iou_threshold=thr)
            pairs = [(i, j) for i in range(K) for j in range(i + 1, K)][:: max(1, K // 3)]
            vals = [content_iou(sp, bboxes, scores, i, j, md) for i, j in pairs]
    np.savez_compressed(os.path.join(HERE, f"refine_nms_{name}.npz"), shape=np.array(masks.shape), seed=seed, thr=thr,
                        full_box_at=np.int64(-1 if full_box_at is None else full_box_at),
                        keep=np.asarray(keep, dtype=np.int64), pairs=np.array(pairs), sketch_iou=np.array([v[0] for v in vals]),
                        bbox_iou=np.array([v[1] for v in vals]), better=np.array([v[2] for v in vals]))
    print(name, "keep", keep)


if __name__ == "__main__":
    make_clean("300x200", 300, 200, 3, 11)
    make_clean("750x750", 750, 750, 3, 12)
    make_clean("1024x1000", 1024, 1000, 2, 13)
    make_nms("512_k12", 512, 512, 12, 21, 0.2)
    make_nms("600x750_k9", 600, 750, 9, 22, 0.5)
    make_nms("drop_mid", 512, 640, 12, 23, 0.2, full_box_at=3)
