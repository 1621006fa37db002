"""Does a CUDA graph of the encoder launch sequence (about 260 kernels per batch of 8) beat stream launches?"""
import os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from inklayer_b200 import CONFIGS, build_sam_from_state_dict
from inklayer_b200.synth import synthetic_sketch
from inklayer_b200.weights import synthetic_state_dict

cfg = CONFIGS["vit_h"]
sam = build_sam_from_state_dict(synthetic_state_dict(cfg, 0), cfg).to("cuda")
imgs = torch.from_numpy(np.stack([synthetic_sketch(i) for i in range(8)])).cuda()
for _ in range(3):
    f = sam.encode(imgs)
torch.cuda.synchronize()

def timed(fn, n=20):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n

t_stream = timed(lambda: sam.encode(imgs))
g = torch.cuda.CUDAGraph()
s = torch.cuda.Stream()
s.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(s):
    sam.encode(imgs)
torch.cuda.current_stream().wait_stream(s)
with torch.cuda.graph(g):
    fg = sam.encode(imgs)
torch.cuda.synchronize()
g.replay(); torch.cuda.synchronize()
print("graph output equals stream output:", torch.equal(fg, f))
t_graph = timed(g.replay)
t_stream2 = timed(lambda: sam.encode(imgs))
print(f"encode batch 8: stream {t_stream:.3f} ms, graph {t_graph:.3f} ms, stream again {t_stream2:.3f} ms")
