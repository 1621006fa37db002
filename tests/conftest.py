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
    if GOLDEN not in sys.path:
        sys.path.insert(0, GOLDEN)
    from cases import case_inputs

    return case_inputs(name)


def golden_masks(gold):
    """All K reference masks of a golden case as bool [K,1,H,W]."""
    shape = tuple(int(v) for v in gold["mask_shape"])
    return np.unpackbits(gold["mask_bits"])[: int(np.prod(shape))].reshape(shape).astype(bool)


def mask_iou(a, b):
    a = np.asarray(a).astype(bool).reshape(a.shape[0], -1)
    b = np.asarray(b).astype(bool).reshape(b.shape[0], -1)
    inter = (a & b).sum(1).astype(np.float64)
    union = (a | b).sum(1).astype(np.float64)
    return np.where(union > 0, inter / np.maximum(union, 1), 1.0)
