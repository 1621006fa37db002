"""-m gpu: the launch sequence that exposed an illegal-address fault in one gemm_tc_kernel instantiation during round 1
(DESIGN.md 7b): a single-CTA (CG = 1) launch of an epilogue family right after CTA-pair (CG = 2) launches of the same
family, small and ViT-H-sized shapes alternating, 40 rounds in one process.  Every launch is synchronised and checked, so
a fault or a wrong tile is attributed to its launch."""
import pytest
import torch
import torch.nn.functional as F

from gpu_util import EPI, gemm, rel_err

pytestmark = pytest.mark.gpu


def rnd(*s, scale=1.0, seed=0):
    return (torch.randn(*s, generator=torch.Generator().manual_seed(seed)) * scale).cuda()


def test_cg1_after_cg2_launch_sequence_40_rounds():
    torch.backends.cuda.matmul.allow_tf32 = False
    cases = []
    for (M, N, K) in [(256, 256, 64), (4096, 3840, 1280)]:
        A = rnd(M, K, seed=1).bfloat16(); W = rnd(N, K, scale=0.05, seed=2).bfloat16(); bias = rnd(N, seed=3)
        ref = A.float() @ W.float().t() + bias
        cases.append((A, W, bias, ref, F.gelu(ref), rnd(M, N, seed=4), rnd(64, N, seed=5)))
    for it in range(40):
        for (A, W, bias, ref, ref_gelu, resid, radd) in cases:
            M, N = ref.shape
            for cg in (2, 1):  # the pair launch first, then the single-CTA launch of the same family
                out = torch.zeros(M, N, device="cuda", dtype=torch.bfloat16)
                for name, want in (("BIAS_BF16", ref), ("GELU", ref_gelu)):
                    gemm(EPI[name], A, W, out, cg=cg, bias=bias)
                    torch.cuda.synchronize()
                    assert rel_err(out, want) < 6e-3, f"round {it} cg={cg} {name} {M}x{N}"
                o32 = resid.clone()
                gemm(EPI["RESID"], A, W, o32, cg=cg, bias=bias, resid=o32)
                torch.cuda.synchronize()
                assert rel_err(o32, ref + resid) < 2e-5, f"round {it} cg={cg} RESID {M}x{N}"
                gemm(EPI["ROWADD"], A, W, o32, cg=cg, bias=bias, rowadd=radd, rowadd_mod=64)
                torch.cuda.synchronize()
                assert rel_err(o32, ref + radd[torch.arange(M, device="cuda") % 64]) < 2e-5, f"round {it} cg={cg} ROWADD"
