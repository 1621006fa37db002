"""Golden vectors for the uint8 bilinear resize (row f2): outputs of Pillow itself (the library the reference's
ResizeLongestSide.apply_image calls through torchvision, SA/utils/transforms.py:26-31) on seeded random images.

    python tests/golden/make_resize_golden.py        # writes tests/golden/resize_pillow.npz (Pillow 12.2.0 here)
"""
import os

import numpy as np
import PIL
from PIL import Image

CASES = [  # (H, W, out_h, out_w): up, down, mixed, one axis unchanged, extreme ratios
    (37, 53, 64, 92), (120, 90, 45, 34), (75, 75, 102, 102), (60, 200, 31, 102), (200, 60, 200, 30), (9, 140, 9, 64),
    (333, 17, 100, 5), (1, 7, 15, 102), (130, 131, 64, 64),
]


def main():
    out = {"pillow_version": np.array(PIL.__version__)}
    for n, (H, W, h, w) in enumerate(CASES):
        img = np.random.default_rng(7000 + n).integers(0, 256, (H, W, 3), dtype=np.uint8)
        out[f"in_{n}"] = img
        out[f"out_{n}"] = np.asarray(Image.fromarray(img).resize((w, h), Image.BILINEAR))
    np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "resize_pillow.npz"), **out)


if __name__ == "__main__":
    main()
