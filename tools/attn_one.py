"""A few launches of one encoder attention kernel (ViT-H shape, batch 8) for ncu captures:
python tools/attn_one.py window|global [launches]"""
import ctypes, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from inklayer_b200 import _lib

lib = _lib.load()
windowed = 1 if (len(sys.argv) < 2 or sys.argv[1] == "window") else 0
n = int(sys.argv[2]) if len(sys.argv) > 2 else 3
B, heads, hd = 8, 16, 80
D = heads * hd
g = torch.Generator().manual_seed(0)
rows, relrows = (B * 25 * 196, 64) if windowed else (B * 4096, 256)
qkv = torch.randn(rows, 3 * D, generator=g).to(torch.bfloat16).cuda()
rel = (torch.randn(relrows, hd, generator=g) * 0.3).to(torch.bfloat16).cuda()
out = torch.zeros(B * 4096, D, dtype=torch.bfloat16, device="cuda")
p = lambda t: ctypes.c_void_p(t.data_ptr())
for _ in range(n):
    _lib.check(lib.inksam_op_attention(windowed, p(qkv), p(rel), p(out), B, heads, hd, ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)))
torch.cuda.synchronize()
print("ok")
