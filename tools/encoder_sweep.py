"""BASELINE.json config 5: SAM ViT-B / ViT-L / ViT-H image encoder alone at batch 1 and 32 (tile-shape coverage:
head_dim 64 and 80, D = 768 / 1024 / 1280), synthetic 1024x1024 sketches, bf16 operands.  Prints one JSON line per
(model, batch): ms per batch, images/s, algorithmic TFLOP/s and the fraction of the measured sustained bf16 peak."""
import json, os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from inklayer_b200 import CONFIGS, build_sam_from_state_dict
from inklayer_b200.synth import synthetic_sketch
from inklayer_b200.weights import synthetic_state_dict

GF = {"vit_h": 5641.8, "vit_l": 2837.0, "vit_b": 937.6}  # algorithmic GFLOP per image, SURVEY.md 8(d)
peak = 1385.9
p = os.path.join(ROOT, "MEASURED_PEAKS.json")
if os.path.exists(p):
    peak = json.load(open(p)).get("bf16_tflops_sustained", peak)
models = sys.argv[1:] or ["vit_b", "vit_l", "vit_h"]
for name in models:
    cfg = CONFIGS[name]
    sam = build_sam_from_state_dict(synthetic_state_dict(cfg, 0), cfg).to("cuda")
    for B in (1, 8, 32):
        imgs = torch.from_numpy(np.stack([synthetic_sketch(i % 8) for i in range(B)])).cuda()
        for _ in range(3):
            sam.encode(imgs)
        torch.cuda.synchronize()
        n = 20 if B == 1 else 5
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            sam.encode(imgs)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / n
        tf = GF[name] * B / ms  # GFLOP per ms == TFLOP/s
        print(json.dumps({"model": name, "batch": B, "ms_per_batch": round(ms, 3), "images_per_s": round(B / ms * 1e3, 1),
                          "tflops": round(tf, 1), "frac_of_sustained_peak": round(tf / peak, 3)}), flush=True)
    del sam
    torch.cuda.empty_cache()
