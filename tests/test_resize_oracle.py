"""CPU: the resize oracle (oracle/resize_oracle.py, a restatement of Pillow's 8-bit bilinear resampler) against the
committed Pillow outputs and against Pillow itself, bit for bit; and the host ResizeLongestSide against the oracle."""
import os

import numpy as np
import pytest

from oracle import resize_oracle as R

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "resize_pillow.npz")


def test_oracle_matches_committed_pillow_outputs():
    g = np.load(GOLD)
    n = 0
    while f"in_{n}" in g:
        ref = g[f"out_{n}"]
        got = R.resize_bilinear_u8(g[f"in_{n}"], ref.shape[0], ref.shape[1])
        assert got.dtype == np.uint8 and np.array_equal(got, ref), f"case {n}"
        n += 1
    assert n >= 9


@pytest.mark.parametrize("hw", [(750, 750), (600, 750), (300, 200), (1333, 751), (64, 1000), (37, 53), (1025, 1024), (1, 7),
                                (777, 3001), (1024, 512)])
def test_oracle_apply_image_matches_pillow_here(hw):
    from PIL import Image

    img = np.random.default_rng(hw[0] * 7 + hw[1]).integers(0, 256, hw + (3,), dtype=np.uint8)
    h, w = R.preprocess_shape(*hw)
    ref = np.asarray(Image.fromarray(img).resize((w, h), Image.BILINEAR)) if (h, w) != hw else img
    assert np.array_equal(R.apply_image(img), ref)


def test_host_transform_matches_oracle():
    from inklayer_b200.resize import ResizeLongestSide

    img = np.random.default_rng(3).integers(0, 256, (420, 333, 3), dtype=np.uint8)
    assert np.array_equal(ResizeLongestSide(1024).apply_image(img), R.apply_image(img))
