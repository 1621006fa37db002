"""Timeline of the windowed attention kernel (v4): CTA 0 records clock64() at its pipeline events (inksam_op_attention_trace);
prints per item the cycle of each event relative to the first, and the steady-state deltas.
Events: 0 T issued  1 S (+bias MMA) issued  2 PV issued | bias warps: 3 T seen  4 T read  5 bias operand stored
        softmax: 6 S seen  7 max pass done  10 P ready | epilogue: 8 O seen  9 S buffer handed back"""
import ctypes, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ["INKSAM_WINDOW_ATTN"] = sys.argv[1] if len(sys.argv) > 1 else "5"
from inklayer_b200 import _lib

lib = _lib.load()
B, heads, hd = 8, 16, 80
D = heads * hd
g = torch.Generator().manual_seed(0)
qkv = torch.randn(B * 25 * 196, 3 * D, generator=g).to(torch.bfloat16).cuda()
rel = (torch.randn(64, hd, generator=g) * 0.3).to(torch.bfloat16).cuda()
out = torch.zeros(B * 4096, D, dtype=torch.bfloat16, device="cuda")
p = lambda t: ctypes.c_void_p(t.data_ptr())
st = lambda: ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
for _ in range(2):
    _lib.check(lib.inksam_op_attention(1, p(qkv), p(rel), p(out), B, heads, hd, st()))
tr = torch.zeros(64 * 16, dtype=torch.int64, device="cuda")
lib.inksam_op_attention_trace(p(tr))
_lib.check(lib.inksam_op_attention(1, p(qkv), p(rel), p(out), B, heads, hd, st()))
torch.cuda.synchronize()
lib.inksam_op_attention_trace(ctypes.c_void_p(0))
t = tr.cpu().view(64, 16).numpy()
n_items = int((t[:, 1] > 0).sum())
t0 = t[0, 0]
names = {0: "T", 1: "S", 2: "PV", 3: "Tseen", 4: "Tread", 5: "Bst", 6: "Sseen", 7: "max", 10: "Prdy", 8: "Oseen", 9: "Ofree"}
order = [0, 3, 4, 5, 1, 6, 7, 10, 2, 8, 9]
print("item " + " ".join(f"{names[e]:>7s}" for e in order))
for n in range(n_items):
    print(f"{n:4d} " + " ".join(f"{int(t[n, e] - t0):7d}" for e in order))
print(f"items {n_items}; span {int(t[n_items - 1, 9] - t0)} clk; per item {(t[n_items - 1, 9] - t0) / n_items:.0f} clk")
