"""GPU timing of the mask decoder + post-process alone: 8 images x 16 boxes (the bench step's decode) and 1 image x 256
boxes (BASELINE.json config 3), 1024x1024 outputs.  The decoder is the same for every encoder size, so the 2-block ViT is
used to keep start-up short."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from inklayer_b200 import CONFIGS, build_sam_from_state_dict
from inklayer_b200.synth import synthetic_boxes
from inklayer_b200.weights import synthetic_state_dict

cfg = CONFIGS["vit_tiny"]
sam = build_sam_from_state_dict(synthetic_state_dict(cfg, 0), cfg).to("cuda")
g = torch.Generator().manual_seed(0)
for B, K in [(8, 16), (1, 256)]:
    emb = torch.randn(B, 256, 64, 64, generator=g).cuda()
    boxes = [torch.from_numpy(synthetic_boxes(i, K)).cuda() for i in range(B)]
    sizes = [(1024, 1024)] * B
    masks = torch.empty(B * K * 1024 * 1024, dtype=torch.uint8, device="cuda")
    for _ in range(3):
        sam.decode_batch(emb, boxes, sizes, sizes, masks_out=masks)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 10
    e0.record()
    for _ in range(n):
        sam.decode_batch(emb, boxes, sizes, sizes, masks_out=masks)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    flop = B * K * 2.804e9 + B * 0.805e9
    print(f"decode {B} images x {K} boxes: {ms:.3f} ms  ({B*K/ms*1e3:.0f} boxes/s, {flop/ms/1e9:.1f} TFLOP/s algorithmic)", flush=True)
