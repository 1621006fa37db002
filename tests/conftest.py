import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


def load_golden(name):
    return dict(np.load(os.path.join(GOLDEN, name + ".npz")))


def golden_case_inputs(name):
    """Same seeded inputs tests/golden/make_golden.py used (config name, weight seed, image, boxes)."""
    import json

    from inklayer_b200.synth import synthetic_boxes, synthetic_sketch

    cases = {
        "tiny_1024": ("vit_tiny", 0, 0, (1024, 1024), 4),
        "tiny80_600x750": ("vit_tiny80", 1, 1, (600, 750), 4),
        "vit_b_750_bunny": ("vit_b", 0, 2, (750, 750), "bunny"),
        "vit_h_1024": ("vit_h", 0, 3, (1024, 1024), 4),
    }
    cfg_name, seed, idx, (H, W), nb = cases[name]
    img = synthetic_sketch(idx, H, W)
    if nb == "bunny":
        bunny = json.load(open(os.path.join(GOLDEN, "bunny_boxes.json")))["bboxes"][:16]
        boxes = np.array(bunny, np.float32) * np.array([W, H, W, H], np.float32)
    else:
        boxes = synthetic_boxes(idx, nb, H, W)
    return cfg_name, seed, img, boxes


def mask_iou(a, b):
    a = np.asarray(a).astype(bool).reshape(a.shape[0], -1)
    b = np.asarray(b).astype(bool).reshape(b.shape[0], -1)
    inter = (a & b).sum(1).astype(np.float64)
    union = (a | b).sum(1).astype(np.float64)
    return np.where(union > 0, inter / np.maximum(union, 1), 1.0)
