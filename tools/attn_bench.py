"""GPU timing of the encoder attention kernels alone (ViT-H shapes, batch 8) through the op-level C ABI."""
import ctypes, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from inklayer_b200 import _lib

lib = _lib.load()
B, heads, hd = 8, 16, 80
D = heads * hd
dev = "cuda"
g = torch.Generator().manual_seed(0)
def p(t): return ctypes.c_void_p(t.data_ptr())
st = lambda: ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
for windowed, rows, relrows in [(0, B * 4096, 256), (1, B * 25 * 196, 64)]:
    qkv = (torch.randn(rows, 3 * D, generator=g) * 1.0).to(torch.bfloat16).to(dev)
    rel = (torch.randn(relrows, hd, generator=g) * 0.3).to(torch.bfloat16).to(dev)
    out = torch.zeros(B * 4096, D, dtype=torch.bfloat16, device=dev)
    for _ in range(3):
        _lib.check(lib.inksam_op_attention(windowed, p(qkv), p(rel), p(out), B, heads, hd, st()))
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 20
    e0.record()
    for _ in range(n):
        _lib.check(lib.inksam_op_attention(windowed, p(qkv), p(rel), p(out), B, heads, hd, st()))
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    keys = 196 if windowed else 4096
    flop = B * heads * 4096 * keys * hd * 4
    print(f"{'window' if windowed else 'global'}: {ms*1e3:.1f} us  {flop/ms/1e9:.1f} TFLOP/s (algorithmic)", flush=True)
