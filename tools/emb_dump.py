"""Dumps the ViT-H embedding and low-res logits of the golden case `vit_h_1024` computed by the package in the current
directory (np.save), for comparing builds against the fp32 oracle: python tools/emb_dump.py out_prefix [--oracle]."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.getcwd())
sys.path.insert(0, os.path.join(os.getcwd(), "tests"))
from conftest import golden_case_inputs  # noqa: E402


def main():
    prefix = sys.argv[1]
    cfg_name, seed, img, boxes = golden_case_inputs("vit_h_1024")
    variant = int(os.environ.get("VARIANT", "0"))
    if variant:  # other weights / sketch / boxes than the golden case
        from inklayer_b200.synth import synthetic_boxes, synthetic_sketch
        seed, img, boxes = seed + variant, synthetic_sketch(10 + variant, 1024, 1024), synthetic_boxes(10 + variant, 8, 1024, 1024)
    from inklayer_b200 import CONFIGS
    from inklayer_b200.weights import synthetic_state_dict
    sd = synthetic_state_dict(CONFIGS[cfg_name], seed)
    if "--oracle" in sys.argv:
        from oracle import sam_oracle as O
        ora = O.OraclePredictor(O.CONFIGS[cfg_name], sd)
        ora.set_image(img)
        tb = O.apply_boxes(torch.from_numpy(boxes), img.shape[:2])
        _, _, low = ora.predict_torch(tb)
        np.save(prefix + "_emb.npy", ora.features.numpy()); np.save(prefix + "_low.npy", low.numpy())
        return
    from inklayer_b200 import SamPredictor, build_sam_from_state_dict
    pred = SamPredictor(build_sam_from_state_dict(sd, CONFIGS[cfg_name]).to("cuda"))
    pred.set_image(img)
    tb = pred.transform.apply_boxes_torch(torch.from_numpy(boxes), img.shape[:2]).to("cuda")
    _, _, low = pred.predict_torch(None, None, boxes=tb, multimask_output=False)
    np.save(prefix + "_emb.npy", pred.features.cpu().numpy()); np.save(prefix + "_low.npy", low.cpu().numpy())


if __name__ == "__main__":
    main()
