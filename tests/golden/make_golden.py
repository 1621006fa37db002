"""Generates tests/golden/*.npz by running the UNMODIFIED reference (vendored segment_anything under
/root/reference) on seeded synthetic weights and inputs.  Runs only in the build container; the vectors it
writes are what pins oracle/sam_oracle.py (tests/test_oracle.py) and, on the GPU box, the CUDA path.

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden.py [case ...]
"""
import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference/InkLayer/third_party/segment-anything")

from segment_anything import SamPredictor  # noqa: E402
from segment_anything.build_sam import _build_sam  # noqa: E402

from inklayer_b200.config import CONFIGS  # noqa: E402
from inklayer_b200.synth import synthetic_boxes, synthetic_sketch  # noqa: E402
from inklayer_b200.weights import synthetic_state_dict  # noqa: E402

# case -> (config, weight seed, sketch index, (H, W), boxes)
BUNNY = json.load(open(os.path.join(HERE, "bunny_boxes.json")))["bboxes"][:16]
CASES = {
    "tiny_1024": ("vit_tiny", 0, 0, (1024, 1024), 4),
    "tiny80_600x750": ("vit_tiny80", 1, 1, (600, 750), 4),
    "vit_b_750_bunny": ("vit_b", 0, 2, (750, 750), "bunny"),
    "vit_h_1024": ("vit_h", 0, 3, (1024, 1024), 4),
}


def case_inputs(name):
    cfg_name, seed, idx, (H, W), nb = CASES[name]
    img = synthetic_sketch(idx, H, W)
    if nb == "bunny":  # InkLayer/utils/processing.py:6-28: normalised xyxy -> pixels
        boxes = np.array(BUNNY, np.float32) * np.array([W, H, W, H], np.float32)
    else:
        boxes = synthetic_boxes(idx, nb, H, W)
    return cfg_name, seed, img, boxes


def main(names):
    torch.manual_seed(0)
    for name in names:
        cfg_name, seed, img, boxes = case_inputs(name)
        cfg = CONFIGS[cfg_name]
        sam = _build_sam(cfg.embed_dim, cfg.depth, cfg.num_heads, list(cfg.global_attn_indexes))
        sam.load_state_dict(synthetic_state_dict(cfg, seed), strict=True)
        pred = SamPredictor(sam)
        pred.set_image(img)
        tb = pred.transform.apply_boxes_torch(torch.from_numpy(boxes), img.shape[:2])
        masks, iou, low = pred.predict_torch(None, None, boxes=tb, multimask_output=False)
        emb = pred.features[0].numpy()
        np.savez_compressed(
            os.path.join(HERE, name + ".npz"),
            boxes=boxes, boxes_input_frame=tb.numpy(),
            emb_sub=emb[:, ::4, ::4].copy(), emb_sum=np.float64(emb.astype(np.float64).sum()),
            emb_abs_sum=np.float64(np.abs(emb.astype(np.float64)).sum()),
            low_sub=low.numpy()[:, :, ::4, ::4].copy(), low_std=np.float64(low.double().std().item()),
            iou=iou.numpy(), mask_bits=np.packbits(masks[:4].numpy().reshape(-1)),
            mask_shape=np.array(masks[:4].shape), mask_area=masks.flatten(1).sum(1).numpy(),
        )
        print(name, "emb std", emb.std(), "low std", low.std().item(), "areas", masks.flatten(1).sum(1)[:4].tolist())


if __name__ == "__main__":
    main(sys.argv[1:] or list(CASES))
