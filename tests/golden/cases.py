"""The seeded inputs of the golden cases, shared by tests/golden/make_golden.py (which runs the unmodified reference on
them) and the tests (which run the oracle / the CUDA path on them).  No reference import here."""
import json
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
BUNNY = json.load(open(os.path.join(HERE, "bunny_boxes.json")))["bboxes"][:16]  # first 16 of output/bunny_cook_sketch/bboxes.json
BUNNY_PNG = os.path.join(HERE, "bunny_cook_sketch.png")  # copy of the reference's data/bunny_cook_sketch.png (input data)

# case -> (config, weight seed, sketch index | "bunny_png", (H, W), boxes: count | "bunny", full low-res logits kept for)
CASES = {
    "tiny_1024": ("vit_tiny", 0, 0, (1024, 1024), 4, 4),
    "tiny80_600x750": ("vit_tiny80", 1, 1, (600, 750), 4, 4),
    "vit_b_750_bunny": ("vit_b", 0, 2, (750, 750), "bunny", 16),
    "vit_h_1024": ("vit_h", 0, 3, (1024, 1024), 4, 4),
    # three more ViT-H weight / sketch / box sets (the variants of profiles/r01r_parity_variants.txt, two at other sizes)
    "vit_h_v1_1024": ("vit_h", 1, 11, (1024, 1024), 8, 8),
    "vit_h_v2_600x750": ("vit_h", 2, 12, (600, 750), 8, 8),
    "vit_h_v3_1024x768": ("vit_h", 3, 13, (1024, 768), 8, 8),
    # sketch 0 of bench.py's rank 0 with its 16 boxes: the `parity` field of the bench line
    "vit_h_bench0": ("vit_h", 0, 0, (1024, 1024), 16, 4),
    # BASELINE.json config 1 as stated: the real 750x750 RGBA PNG, first 16 GroundingDINO boxes, ViT-H fp32, entered
    # through InkLayer/segmentor/sam.py::run_SAM (the image below is what the CALLER holds: RGB, before run_SAM's swap)
    "vit_h_750_bunny": ("vit_h", 0, "bunny_png", (750, 750), "bunny", 16),
}


def bunny_rgb() -> np.ndarray:
    """InkLayer/runner.py:49: Image.open(path).convert("RGB")."""
    from PIL import Image

    return np.array(Image.open(BUNNY_PNG).convert("RGB"))


def case_inputs(name):
    """-> config name, weight seed, uint8 RGB image as the caller holds it, float32 [K,4] pixel xyxy boxes."""
    from inklayer_b200.synth import synthetic_boxes, synthetic_sketch

    cfg_name, seed, idx, (H, W), nb, _ = CASES[name]
    if idx == "bunny_png":
        img = bunny_rgb()
        assert img.shape[:2] == (H, W)
    else:
        img = synthetic_sketch(idx, H, W)
    if nb == "bunny":  # InkLayer/utils/processing.py:6-28: normalised xyxy -> pixels
        boxes = np.array(BUNNY, np.float32) * np.array([W, H, W, H], np.float32)
    else:
        boxes = synthetic_boxes(idx, nb, H, W)
    return cfg_name, seed, img, boxes
