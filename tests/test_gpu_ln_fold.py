"""-m gpu: LayerNorm folded into the GEMMs either side of it (include/inksam_ops.h, image_encoder.py:180-197), through the
C ABI, against plain torch fp32:   Linear(LayerNorm(x))  ==  rstd * (bf16(x) (gamma o W)^T) - rstd * mean * colsum + (b + W beta).
Tolerances: the folded product carries the bf16 rounding of x and of gamma o W (2^-9 each, averaged over K) like the
unfused path carries that of LN(x) and W, plus the bf16 rounding of the output: max-error bound 8e-3 of the output range,
rel-RMS 4e-3.  Row statistics and the fp32 residual output are fp32 work: 2e-5."""
import ctypes

import pytest
import torch
import torch.nn.functional as F

from gpu_util import EPI, gemm, ptr, rel_err, rel_rms, stream
from inklayer_b200 import _lib

pytestmark = pytest.mark.gpu


def rnd(*s, scale=1.0, seed=0):
    return (torch.randn(*s, generator=torch.Generator().manual_seed(seed)) * scale).cuda()


def fold_weights(W, gamma, beta, bias):
    N, K = W.shape
    wf = torch.empty(N, K, device="cuda", dtype=torch.bfloat16)
    cs = torch.empty(N, device="cuda"); bf = torch.empty(N, device="cuda")
    _lib.check(_lib.load().inksam_op_fold_ln_weights(ptr(W), ptr(gamma), ptr(beta), ptr(bias), ptr(wf), ptr(cs), ptr(bf), N, K, stream()))
    return wf, cs, bf


def gemm_fold(epi, A, W, out, bias=None, resid=None, stats_out=None, xb_out=None, stats_in=None, colsum=None, np_=0, win_remap=0):
    M, K = A.shape
    N = W.shape[0]
    _lib.check(_lib.load().inksam_op_gemm_fold(epi, ptr(A), A.stride(0), ptr(W), W.stride(0), M, N, K, ptr(out), N, ptr(bias), ptr(resid),
                                               ptr(stats_out), ptr(xb_out), ptr(stats_in), ptr(colsum), np_, ctypes.c_float(1e-6),
                                               win_remap, stream()))


def test_fold_ln_weights_vectors():
    N, K = 300, 160
    W = rnd(N, K, scale=0.05, seed=1); gamma = 1 + 0.2 * rnd(K, seed=2); beta = 0.1 * rnd(K, seed=3); bias = rnd(N, seed=4)
    wf, cs, bf = fold_weights(W, gamma, beta, bias)
    want = (W * gamma).bfloat16()
    assert torch.equal(wf, want)                                     # one rounding, of the fp32 product
    assert rel_err(cs, want.float().sum(1)) < 1e-5                   # row sums of the ROUNDED weights
    assert rel_err(bf, bias + W @ beta) < 1e-5


@pytest.mark.parametrize("M,D", [(4096, 1280), (777, 768), (300, 160)])
def test_row_stats_and_cast(M, D):
    x = rnd(M, D, seed=1) + 0.3
    xb = torch.empty(M, D, device="cuda", dtype=torch.bfloat16)
    st = torch.empty(M, 2, device="cuda")
    _lib.check(_lib.load().inksam_op_row_stats_bf16(ptr(x), ptr(xb), ptr(st), M, D, stream()))
    assert torch.equal(xb, x.bfloat16())
    assert rel_err(st[:, 0], x.double().sum(1)) < 2e-5
    assert rel_err(st[:, 1], x.double().pow(2).sum(1)) < 2e-5


@pytest.mark.parametrize("M,N,D,mean,rms_tol", [(4096, 3840, 1280, 0.0, 4e-3), (4096 + 77, 5120, 1280, 0.5, 4e-3), (300, 480, 160, 0.2, 4e-3),
                                                (8192, 2304, 768, 2.0, 1e-2)])
def test_consumer_gemm_matches_layernorm_linear(M, N, D, mean, rms_tol):
    """qkv / lin1 with the LayerNorm folded in == Linear(LayerNorm(x)) in fp32.  The bf16 rounding of the raw x scales with
    |x| rather than |x - mean|, so rows whose mean exceeds their spread lose accuracy (last case: |mean| / std up to 4,
    looser bound); the encoder's token rows sit at |mean| / std ~ 0.02 (tools/ln_fold_probe.py)."""
    torch.backends.cuda.matmul.allow_tf32 = False
    x = rnd(M, D, seed=1) * (0.5 + torch.rand(M, 1, generator=torch.Generator().manual_seed(9)).cuda() * 3) + mean
    W = rnd(N, D, scale=0.05, seed=2); gamma = 1 + 0.2 * rnd(D, seed=3); beta = 0.1 * rnd(D, seed=4); bias = rnd(N, seed=5)
    ref = F.linear(F.layer_norm(x, (D,), gamma, beta, 1e-6), W, bias)
    wf, cs, bf = fold_weights(W, gamma, beta, bias)
    xb = torch.empty(M, D, device="cuda", dtype=torch.bfloat16)
    st = torch.empty(M, 2, device="cuda")
    _lib.check(_lib.load().inksam_op_row_stats_bf16(ptr(x), ptr(xb), ptr(st), M, D, stream()))
    out = torch.zeros(M, N, device="cuda", dtype=torch.bfloat16)
    gemm_fold(EPI["BIAS_BF16"], xb, wf, out, bias=bf, stats_in=st, colsum=cs, np_=1)
    assert rel_err(out, ref) < 4 * rms_tol and rel_rms(out, ref) < rms_tol
    gemm_fold(EPI["GELU"], xb, wf, out, bias=bf, stats_in=st, colsum=cs, np_=1)
    assert rel_err(out, F.gelu(ref)) < 4 * rms_tol and rel_rms(out, F.gelu(ref)) < rms_tol
    if mean > 0.5:
        return
    # the unfused path of the same op (LayerNorm kernel -> bf16 GEMM) is no closer to fp32
    xn = torch.empty(M, D, device="cuda", dtype=torch.bfloat16)
    _lib.check(_lib.load().inksam_op_layernorm(ptr(x), ptr(gamma), ptr(beta), ptr(xn), M, D, ctypes.c_float(1e-6), 0, stream()))
    out2 = torch.zeros(M, N, device="cuda", dtype=torch.bfloat16)
    gemm(EPI["BIAS_BF16"], xn, W.bfloat16(), out2, cg=2, bias=bias)
    gemm_fold(EPI["BIAS_BF16"], xb, wf, out, bias=bf, stats_in=st, colsum=cs, np_=1)
    assert rel_rms(out, ref) < 1.5 * rel_rms(out2, ref) + 1e-4


@pytest.mark.parametrize("offset,outlier", [(0.0, 500.0), (8.0, 0.0), (100.0, 500.0)])
def test_fold_with_large_row_means_and_outlier_channels(offset, outlier):
    """what the fold costs on rows unlike the synthetic weights' (real checkpoints are not available on this box):
    x = offset + N(0, 1) with one channel at `outlier`.  The fold rounds the RAW x to bf16 (8 significant bits), so a common
    offset that LayerNorm would remove costs  ulp(offset) / sqrt(12) / std  of relative error in the normalised row
    (offset 100: ulp 0.5 -> 14 %); an outlier CHANNEL is harmless (it dominates the variance, the other channels keep their
    relative precision).  The test pins that law: measured error within 0.5x .. 2x of the prediction, and inside the normal
    tolerance when there is no offset.  (Remedy if real weights need it, not built: the producer stores bf16(x - c_row) with
    c_row the row mean of the previous sub-layer -- LayerNorm is shift-invariant per row.)"""
    torch.backends.cuda.matmul.allow_tf32 = False
    M, N, D = 2048, 1024, 1280
    x = rnd(M, D, seed=1) + offset
    if outlier:
        x[:, 77] = outlier
    W = rnd(N, D, scale=0.05, seed=2); gamma = 1 + 0.2 * rnd(D, seed=3); beta = 0.1 * rnd(D, seed=4); bias = rnd(N, seed=5)
    ref = F.linear(F.layer_norm(x, (D,), gamma, beta, 1e-6), W, bias)
    wf, cs, bf = fold_weights(W, gamma, beta, bias)
    xb = torch.empty(M, D, device="cuda", dtype=torch.bfloat16)
    st = torch.empty(M, 2, device="cuda")
    _lib.check(_lib.load().inksam_op_row_stats_bf16(ptr(x), ptr(xb), ptr(st), M, D, stream()))
    outb = torch.zeros(M, N, device="cuda", dtype=torch.bfloat16)
    gemm_fold(EPI["BIAS_BF16"], xb, wf, outb, bias=bf, stats_in=st, colsum=cs, np_=1)
    # error of the normalised row that reaches the GEMM: bf16(x) against x, in units of the row's standard deviation
    std = x.std(dim=1, keepdim=True)
    pred = ((xb.float() - x) / std).pow(2).mean().sqrt().item()
    signal = ((x - x.mean(dim=1, keepdim=True)) / std).pow(2).mean().sqrt().item()
    got = rel_rms(outb, ref)
    print(f"offset {offset}, outlier {outlier}: input rounding / std {pred:.2e}, folded GEMM rel-RMS vs fp32 {got:.2e}")
    assert torch.isfinite(outb.float()).all()
    if offset == 0.0:
        assert got < 4e-3                       # an outlier channel alone stays inside the normal tolerance
    else:
        assert 0.3 * pred / signal < got < 2.5 * pred / signal + 4e-3


@pytest.mark.parametrize("M,N,K", [(4096, 1280, 1280), (4096 + 77, 1280, 5120), (300, 160, 640), (2048, 768, 768)])
def test_producer_gemm_emits_stats_and_bf16_copy(M, N, K):
    """proj / lin2: out = resid + acc + bias in place, plus partial (sum, sum of squares) of each output row and bf16(out);
    chained into the consumer, the partials give the same result as statistics of the finished rows."""
    torch.backends.cuda.matmul.allow_tf32 = False
    lib = _lib.load()
    A = rnd(M, K, seed=1).bfloat16(); W = rnd(N, K, scale=0.05, seed=2).bfloat16(); bias = rnd(N, seed=3)
    resid = rnd(M, N, seed=4) + 0.25
    ref = A.float() @ W.float().t() + bias + resid
    np_ = lib.inksam_op_gemm_stats_partials(N)
    assert np_ == 2 * ((N + 255) // 256)
    st = torch.full((np_, M, 2), float("nan"), device="cuda")
    xb = torch.zeros(M, N, device="cuda", dtype=torch.bfloat16)
    out = resid.clone()
    gemm_fold(EPI["RESID"], A, W, out, bias=bias, resid=out, stats_out=st, xb_out=xb)
    assert rel_err(out, ref) < 2e-5
    assert torch.equal(xb, out.bfloat16())
    assert rel_err(st[:, :, 0].sum(0), out.double().sum(1)) < 2e-5
    assert rel_err(st[:, :, 1].sum(0), out.double().pow(2).sum(1)) < 2e-5
    # without the side outputs the epilogue is the plain residual one
    out2 = resid.clone()
    gemm_fold(EPI["RESID"], A, W, out2, bias=bias, resid=out2)
    assert torch.equal(out2, out)
    # chain: consumer on (xb, partials) == Linear(LayerNorm(out))
    N2 = 512
    W2 = rnd(N2, N, scale=0.05, seed=6); gamma = 1 + 0.2 * rnd(N, seed=7); beta = 0.1 * rnd(N, seed=8); b2 = rnd(N2, seed=9)
    wf, cs, bf = fold_weights(W2, gamma, beta, b2)
    y = torch.zeros(M, N2, device="cuda", dtype=torch.bfloat16)
    gemm_fold(EPI["BIAS_BF16"], xb, wf, y, bias=bf, stats_in=st, colsum=cs, np_=np_)
    ref2 = F.linear(F.layer_norm(out, (N,), gamma, beta, 1e-6), W2, b2)
    assert rel_err(y, ref2) < 1.6e-2 and rel_rms(y, ref2) < 4e-3


def test_gemm_fold_rejects_mismatched_arguments():
    lib = _lib.load()
    a = torch.zeros(256, 64, device="cuda", dtype=torch.bfloat16); w = torch.zeros(256, 64, device="cuda", dtype=torch.bfloat16)
    o = torch.zeros(256, 256, device="cuda"); s = torch.zeros(2, 256, 2, device="cuda")
    assert lib.inksam_op_gemm_fold(EPI["ROWADD"], ptr(a), 64, ptr(w), 64, 256, 256, 64, ptr(o), 256, None, None, None, None, None, None, 0,
                                   ctypes.c_float(1e-6), 0, stream()) != 0
    assert lib.inksam_op_gemm_fold(EPI["RESID"], ptr(a), 64, ptr(w), 64, 256, 256, 64, ptr(o), 256, None, ptr(o), None, None, ptr(s), None, 1,
                                   ctypes.c_float(1e-6), 0, stream()) != 0
    assert lib.inksam_op_gemm_fold(EPI["BIAS_BF16"], ptr(a), 64, ptr(w), 64, 256, 256, 64, ptr(o), 256, None, None, None, None, ptr(s), None, 1,
                                   ctypes.c_float(1e-6), 0, stream()) != 0
    assert b"inksam_op_gemm_fold" in lib.inksam_last_error()
