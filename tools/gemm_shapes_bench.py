"""Times the encoder's four GEMM shapes (ViT-H; batch sizes from the command line, default 8 and 1) through the op-level C ABI,
kernels alone with an L2 flush between launches (unfolded epilogues: bias / bias + GELU / bias + fp32 residual in place)."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from gpu_util import EPI, gemm  # noqa: E402

flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

def t(M, N, K, epi, n=20):
    A = torch.randn(M, K, device="cuda").bfloat16(); W = (torch.randn(N, K, device="cuda") * 0.05).bfloat16()
    bias = torch.randn(N, device="cuda")
    out = torch.zeros(M, N, device="cuda", dtype=torch.float32 if epi == EPI["RESID"] else torch.bfloat16)
    kw = dict(bias=bias, resid=out) if epi == EPI["RESID"] else dict(bias=bias)
    for _ in range(3):
        gemm(epi, A, W, out, cg=2, **kw)
    tot = 0.0
    for _ in range(n):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); gemm(epi, A, W, out, cg=2, **kw); e1.record(); torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / n * 1e3

for B in [int(x) for x in (sys.argv[1:] or ["8", "1"])]:
    M = B * 4096
    for name, N, K, epi in [("qkv", 3840, 1280, EPI["BIAS_BF16"]), ("proj", 1280, 1280, EPI["RESID"]), ("lin1", 5120, 1280, EPI["GELU"]),
                            ("lin2", 1280, 5120, EPI["RESID"])]:
        us = t(M, N, K, epi)
        print(f"B={B} {name:5s} M={M} N={N} K={K}: {us:7.1f} us  {2.0 * M * N * K / us / 1e6:7.1f} TFLOP/s", flush=True)
