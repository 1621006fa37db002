"""Timeline of the global attention kernel: CTA 0 records clock64() at its pipeline events (inksam_op_attention_trace).
Needs a build with the hooks compiled in:  make -C inklayer_b200/csrc clean && make -C inklayer_b200/csrc EXTRA=-DINKSAM_ATTN_TRACE  Per 64-key sub-tile t, query tile A: events 0-3 P(t) published by softmax warp 0-3,
4-7 S(t) seen by warp 0-3; issuer of tile A: 9 starts waiting for P(t), 10 P(t) seen, 11 PV(t) and S(t+2) issued."""
import ctypes, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from inklayer_b200 import _lib

lib = _lib.load()
B, heads, hd = 8, 16, 80
D = heads * hd
g = torch.Generator().manual_seed(0)
qkv = torch.randn(B * 4096, 3 * D, generator=g).to(torch.bfloat16).cuda()
rel = (torch.randn(256, hd, generator=g) * 0.3).to(torch.bfloat16).cuda()
out = torch.zeros(B * 4096, D, dtype=torch.bfloat16, device="cuda")
p = lambda t: ctypes.c_void_p(t.data_ptr())
st = lambda: ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
for _ in range(2):
    _lib.check(lib.inksam_op_attention(0, p(qkv), p(rel), p(out), B, heads, hd, st()))
tr = torch.zeros(64 * 16, dtype=torch.int64, device="cuda")
lib.inksam_op_attention_trace(p(tr))
_lib.check(lib.inksam_op_attention(0, p(qkv), p(rel), p(out), B, heads, hd, st()))
torch.cuda.synchronize()
lib.inksam_op_attention_trace(ctypes.c_void_p(0))
t = tr.cpu().view(64, 16).numpy()
t0 = t[0, 4]
names = ["pub0", "pub1", "pub2", "pub3", "S0", "S1", "S2", "S3", "-", "i:wait", "i:P", "i:iss"]
print("  t " + " ".join(f"{n:>7s}" for n in names))
for n in list(range(0, 8)) + list(range(40, 48)):
    print(f"{n:3d} " + " ".join(f"{int(t[n, e] - t0):7d}" if e != 8 else "      -" for e in range(12)))
d = t[8:64] - t[7:63]
print("steady-state period per sub-tile (clk):", float(d[:, 4].mean()))
