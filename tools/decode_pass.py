"""One decode pass of the bench step (8 images x 16 boxes, 1024x1024 bit-packed masks) inside a cudaProfilerStart / Stop
range, after 3 warm passes -- for `ncu --profile-from-start off` captures of exactly one pass (tools/gpu_profile_r02.sh;
tools/ncu_summary.py decoder-traffic turns the launch list into profiles/decoder_traffic.json).
  python tools/decode_pass.py [images boxes_per_image]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from inklayer_b200 import CONFIGS, build_sam_from_state_dict  # noqa: E402
from inklayer_b200.synth import synthetic_boxes  # noqa: E402
from inklayer_b200.weights import synthetic_state_dict  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
K = int(sys.argv[2]) if len(sys.argv) > 2 else 16
cfg = CONFIGS["vit_tiny"]  # the decoder does not depend on the encoder size
sam = build_sam_from_state_dict(synthetic_state_dict(cfg, 0), cfg).to("cuda")
emb = torch.randn(B, 256, 64, 64, generator=torch.Generator().manual_seed(0)).cuda()
boxes = [torch.from_numpy(synthetic_boxes(i, K)).cuda() for i in range(B)]
sizes = [(1024, 1024)] * B
masks = torch.empty(B * K * 1024 * 128, dtype=torch.uint8, device="cuda")
for _ in range(3):
    sam.decode_batch(emb, boxes, sizes, sizes, masks_out=masks, packed=True)
torch.cuda.synchronize()
n0 = sam.launch_count
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.profiler.start()
e0.record()
sam.decode_batch(emb, boxes, sizes, sizes, masks_out=masks, packed=True)
e1.record()
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print(f"decode pass: {B} images x {K} boxes, {sam.launch_count - n0} launches, {e0.elapsed_time(e1):.3f} ms")
