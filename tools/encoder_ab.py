"""Encoder batch time (ViT-H, batch 8 by default, CUDA-graph replay) for the library named by INKSAM_LIB: same-box A/B runs of
two builds.  Entry points that an older build lacks are skipped when the library is loaded."""
import ctypes, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from inklayer_b200 import _lib  # noqa: E402
probe = ctypes.CDLL(_lib.LIB_PATH)
for name in list(_lib.SIGNATURES):
    if not hasattr(probe, name):
        del _lib.SIGNATURES[name]
from inklayer_b200 import CONFIGS, build_sam_from_state_dict  # noqa: E402
from inklayer_b200.weights import synthetic_state_dict  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
cfg = CONFIGS["vit_h"]
sam = build_sam_from_state_dict(synthetic_state_dict(cfg, 0), cfg).to("cuda")
imgs = torch.randint(0, 256, (B, 1024, 1024, 3), dtype=torch.uint8, device="cuda")
sizes = [(1024, 1024)] * B
for _ in range(5):
    sam.encode(imgs, sizes, graph=True)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
n = 30
e0.record()
for _ in range(n):
    sam.encode(imgs, sizes, graph=True)
e1.record(); torch.cuda.synchronize()
print(f"{os.path.basename(_lib.LIB_PATH)}: batch {B}: {e0.elapsed_time(e1) / n:.3f} ms per encode")
