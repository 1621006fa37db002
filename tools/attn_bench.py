"""GPU timing of the encoder attention kernels alone (ViT-H shapes, batch 8 and 1; ViT-B head_dim 64) through the op-level
C ABI.  The windowed kernel is timed in each of its variants (INKSAM_WINDOW_ATTN / INKSAM_WINDOW_KPOLY are read per launch)."""
import ctypes, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from inklayer_b200 import _lib

lib = _lib.load()
dev = "cuda"
g = torch.Generator().manual_seed(0)
def p(t): return ctypes.c_void_p(t.data_ptr())
st = lambda: ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)

def run(windowed, B, heads, hd, n=20):
    D = heads * hd
    rows, relrows = (B * 25 * 196, 64) if windowed else (B * 4096, 256)
    qkv = (torch.randn(rows, 3 * D, generator=g) * 1.0).to(torch.bfloat16).to(dev)
    rel = (torch.randn(relrows, hd, generator=g) * 0.3).to(torch.bfloat16).to(dev)
    out = torch.zeros(B * 4096, D, dtype=torch.bfloat16, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for _ in range(3):
        _lib.check(lib.inksam_op_attention(windowed, p(qkv), p(rel), p(out), B, heads, hd, st()))
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(n):
        flush.zero_()  # qkv of a layer is not L2-resident in the encoder either (252 MB at batch 8)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _lib.check(lib.inksam_op_attention(windowed, p(qkv), p(rel), p(out), B, heads, hd, st()))
        e1.record(); torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    ms = tot / n
    keys = 196 if windowed else 4096
    flop = B * heads * 4096 * keys * hd * 4
    return ms, flop / ms / 1e9, out

for (B, heads, hd) in [(8, 16, 80), (1, 16, 80), (8, 12, 64)]:
    for kp in ("0",):
        os.environ["INKSAM_GLOBAL_KPOLY"] = kp
        ms, tf, _ = run(0, B, heads, hd, n=10)
        print(f"global kpoly {kp:>2} B={B} heads={heads} hd={hd}: {ms*1e3:8.1f} us  {tf:7.1f} TFLOP/s (algorithmic)", flush=True)
    os.environ.pop("INKSAM_GLOBAL_KPOLY", None)
    ref = None
    for name, env in [("v2", {"INKSAM_WINDOW_ATTN": "2"}),
                      
                      ("v4", {"INKSAM_WINDOW_ATTN": "4"}), ("v5 kpoly 0", {"INKSAM_WINDOW_ATTN": "5", "INKSAM_WINDOW_KPOLY": "0"}),
                      ("v5 kpoly 4", {"INKSAM_WINDOW_ATTN": "5", "INKSAM_WINDOW_KPOLY": "4"}),
                      ("v5 kpoly 8", {"INKSAM_WINDOW_ATTN": "5", "INKSAM_WINDOW_KPOLY": "8"}),
                      ("v5 kpoly 12", {"INKSAM_WINDOW_ATTN": "5", "INKSAM_WINDOW_KPOLY": "12"}),
                      ("v5 kpoly 16", {"INKSAM_WINDOW_ATTN": "5", "INKSAM_WINDOW_KPOLY": "16"})]:
        for k in ("INKSAM_WINDOW_ATTN", "INKSAM_WINDOW_KPOLY"):
            os.environ.pop(k, None)
        os.environ.update(env)
        ms, tf, out = run(1, B, heads, hd)
        d = "" if ref is None else f"  max |diff| vs v2 {(out.float() - ref.float()).abs().max().item():.3e}"
        if ref is None:
            ref = out.clone()
        print(f"window {name:<11} B={B} heads={heads} hd={hd}: {ms*1e3:8.1f} us  {tf:7.1f} TFLOP/s{d}", flush=True)
