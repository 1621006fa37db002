"""Times the four GEMMs of a ViT-H block at batch 8 in their folded-LayerNorm form (CUDA events, kernels alone, working
sets larger than L2).  INKSAM_LIB selects an alternative build for same-box A/B runs."""
import ctypes
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from gpu_util import ptr, stream  # noqa: E402
from inklayer_b200 import _lib  # noqa: E402


def main():
    lib = _lib.load()
    M, D = 8 * 4096, 1280
    g = torch.Generator().manual_seed(0)
    rnd = lambda *s, scale=1.0: (torch.randn(*s, generator=g) * scale).cuda()  # noqa: E731
    x = rnd(M, D)
    xb = torch.empty(M, D, device="cuda", dtype=torch.bfloat16)
    np_ = lib.inksam_op_gemm_stats_partials(D)
    st = torch.zeros(np_, M, 2, device="cuda")
    attn = rnd(M, D).bfloat16()
    hbuf = torch.empty(M, 4 * D, device="cuda", dtype=torch.bfloat16)
    qkv = torch.empty(M, 3 * D, device="cuda", dtype=torch.bfloat16)

    def folded(N):
        W = rnd(N, D, scale=0.03); gamma = 1 + 0.1 * rnd(D); beta = 0.1 * rnd(D); bias = rnd(N, scale=0.1)
        wf = torch.empty(N, D, device="cuda", dtype=torch.bfloat16); cs = torch.empty(N, device="cuda"); bf = torch.empty(N, device="cuda")
        _lib.check(lib.inksam_op_fold_ln_weights(ptr(W), ptr(gamma), ptr(beta), ptr(bias), ptr(wf), ptr(cs), ptr(bf), N, D, stream()))
        return wf, cs, bf

    wq, cq, bq = folded(3 * D)
    w1, c1, b1 = folded(4 * D)
    wp = rnd(D, D, scale=0.03).bfloat16(); bp = rnd(D, scale=0.1)
    w2 = rnd(D, 4 * D, scale=0.02).bfloat16(); b2 = rnd(D, scale=0.1)
    eps = ctypes.c_float(1e-6)

    def call(epi, A, lda, W, ldw, N, K, out, ldo, bias, resid=None, so=None, xo=None, si=None, cs=None, npi=0):
        _lib.check(lib.inksam_op_gemm_fold(epi, ptr(A), lda, ptr(W), ldw, M, N, K, ptr(out), ldo, ptr(bias), ptr(resid), ptr(so), ptr(xo),
                                           ptr(si), ptr(cs), npi, eps, 0, stream()))

    ops = {
        "proj  (resid + stats)": lambda: call(2, attn, D, wp, D, D, D, x, D, bp, resid=x, so=st, xo=xb),
        "qkv   (fold)": lambda: call(0, xb, D, wq, D, 3 * D, D, qkv, 3 * D, bq, si=st, cs=cq, npi=np_),
        "lin1  (fold + gelu)": lambda: call(1, xb, D, w1, D, 4 * D, D, hbuf, 4 * D, b1, si=st, cs=c1, npi=np_),
        "lin2  (resid + stats)": lambda: call(2, hbuf, 4 * D, w2, 4 * D, D, 4 * D, x, D, b2, resid=x, so=st, xo=xb),
        "proj  (plain resid)": lambda: call(2, attn, D, wp, D, D, D, x, D, bp, resid=x),
        "lin2  (plain resid)": lambda: call(2, hbuf, 4 * D, w2, 4 * D, D, 4 * D, x, D, b2, resid=x),
    }
    ops["proj  (resid + stats)"]()  # valid statistics for the consumers
    total = 0.0
    for name, fn in ops.items():
        for _ in range(3):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(20):
            fn()
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / 20
        if "plain" not in name:
            total += us
        print(f"{name:24s} {us:8.1f} us")
        x.normal_()
    print(f"{'block (4 folded GEMMs)':24s} {total:8.1f} us   lib={os.path.basename(_lib.LIB_PATH)}")


if __name__ == "__main__":
    main()
