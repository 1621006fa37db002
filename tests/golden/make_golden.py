"""Generates tests/golden/*.npz by running the UNMODIFIED reference (vendored segment_anything, and for the config-1
case InkLayer's own `run_SAM`, under /root/reference) on seeded synthetic weights and inputs.  Runs only in the build
container; the vectors it writes are what pins oracle/sam_oracle.py (tests/test_oracle.py) and, on the GPU box, the
CUDA path (tests/test_gpu_e2e.py, tests/test_gpu_config1.py, bench.py's `parity` field).

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden.py [case ...]

Every case stores the FULL low-res logits [K,1,256,256] (fp32) of up to `low_keep` boxes, ALL K binarised masks
(bit-packed), the IoU head output, a ::4 sub-sample of the embedding plus its checksums, and the boxes in both frames.
"""
import os
import sys
import tempfile

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.dont_write_bytecode = True
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)
sys.path.insert(0, "/root/reference")
sys.path.insert(0, "/root/reference/InkLayer/third_party/segment-anything")

from segment_anything import SamPredictor  # noqa: E402
from segment_anything.build_sam import _build_sam  # noqa: E402

from inklayer_b200.config import CONFIGS  # noqa: E402
from inklayer_b200.synth import synthetic_boxes, synthetic_sketch  # noqa: E402
from inklayer_b200.weights import synthetic_state_dict  # noqa: E402

from cases import BUNNY, CASES, case_inputs  # noqa: E402


def reference_run_sam(img_rgb, boxes_norm, sd):
    """The reference's own entry (InkLayer/segmentor/sam.py:16-43) with the synthetic weights saved as its checkpoint
    and the boxes converted by its own process_boxes_ours (InkLayer/utils/processing.py:6-28)."""
    from PIL import Image

    from InkLayer.segmentor.sam import run_SAM
    from InkLayer.utils.processing import process_boxes_ours

    pil = Image.fromarray(img_rgb)
    boxes_filt = process_boxes_ours({"bboxes": boxes_norm}, pil)
    with tempfile.TemporaryDirectory() as d:
        ck = os.path.join(d, "sam_synth.pth")
        torch.save(sd, ck)
        masks = run_SAM(image_pil=pil, boxes_filt=boxes_filt, sam_checkpoint=ck)
    return np.stack(masks), boxes_filt.numpy()


def main(names):
    torch.manual_seed(0)
    for name in names:
        cfg_name, seed, img, boxes = case_inputs(name)
        low_keep = CASES[name][5]
        cfg = CONFIGS[cfg_name]
        sd = synthetic_state_dict(cfg, seed)
        sam = _build_sam(cfg.embed_dim, cfg.depth, cfg.num_heads, list(cfg.global_attn_indexes))
        sam.load_state_dict(sd, strict=True)
        pred = SamPredictor(sam)
        extra = {}
        model_img = img
        if name == "vit_h_750_bunny":
            import cv2
            model_img = cv2.cvtColor(img, cv2.COLOR_BGR2RGB)  # segmentor/sam.py:24-25: the model sees BGR
            run_masks, boxes_filt = reference_run_sam(img, BUNNY, sd)
            extra["run_sam_boxes_filt"] = boxes_filt          # what process_boxes_ours hands to run_SAM (float32 tensor)
            boxes = boxes_filt.astype(np.float32)
        pred.set_image(model_img)
        tb = pred.transform.apply_boxes_torch(torch.from_numpy(boxes), img.shape[:2])
        masks, iou, low = pred.predict_torch(None, None, boxes=tb, multimask_output=False)
        if name == "vit_h_750_bunny":  # run_SAM and the predictor sequence it wraps agree bit for bit
            assert run_masks.dtype == np.bool_ and np.array_equal(run_masks, masks[:, 0].numpy())
        emb = pred.features[0].numpy()
        np.savez_compressed(
            os.path.join(HERE, name + ".npz"),
            boxes=boxes, boxes_input_frame=tb.numpy(),
            emb_sub=emb[:, ::4, ::4].copy(), emb_sum=np.float64(emb.astype(np.float64).sum()),
            emb_abs_sum=np.float64(np.abs(emb.astype(np.float64)).sum()),
            low=low.numpy()[:low_keep].copy(), low_sub=low.numpy()[:, :, ::4, ::4].copy(),
            low_std=np.float64(low.double().std().item()),
            iou=iou.numpy(), mask_bits=np.packbits(masks.numpy().reshape(-1)),
            mask_shape=np.array(masks.shape), mask_area=masks.flatten(1).sum(1).numpy(), **extra,
        )
        print(name, "emb std", emb.std(), "low std", low.std().item(), "areas", masks.flatten(1).sum(1)[:4].tolist(), flush=True)


if __name__ == "__main__":
    main(sys.argv[1:] or list(CASES))
