"""-m gpu: every epilogue of the tcgen05 GEMM against plain torch fp32 on the same operands, through the C ABI.
Tolerances: bf16 operands are exact inputs (products accumulate in fp32), so fp32 outputs match to ~1e-5 relative;
bf16 outputs carry one rounding (2^-9); tf32 operands are truncated to 10 mantissa bits (2e-3)."""
import pytest
import torch
import torch.nn.functional as F

from gpu_util import EPI, gemm, rel_err

pytestmark = pytest.mark.gpu


def rnd(*s, scale=1.0, seed=0):
    return (torch.randn(*s, generator=torch.Generator().manual_seed(seed)) * scale).cuda()


@pytest.mark.parametrize("cg", [1, 2])
@pytest.mark.parametrize("M,N,K", [(256, 256, 64), (4096, 3840, 1280), (112, 256, 256), (4096 + 77, 1280, 5120), (300, 160, 160)])
def test_bf16_epilogues(cg, M, N, K):
    torch.backends.cuda.matmul.allow_tf32 = False
    A = rnd(M, K, seed=1).bfloat16(); W = rnd(N, K, scale=0.05, seed=2).bfloat16(); bias = rnd(N, seed=3)
    ref = A.float() @ W.float().t() + bias
    out = torch.zeros(M, N, device="cuda", dtype=torch.bfloat16)
    gemm(EPI["BIAS_BF16"], A, W, out, cg=cg, bias=bias)
    assert rel_err(out, ref) < 6e-3
    gemm(EPI["GELU"], A, W, out, cg=cg, bias=bias)
    assert rel_err(out, F.gelu(ref)) < 6e-3
    resid = rnd(M, N, seed=4)
    o32 = resid.clone()
    gemm(EPI["RESID"], A, W, o32, cg=cg, bias=bias, resid=o32)  # in place, as the encoder uses it
    assert rel_err(o32, ref + resid) < 2e-5
    radd = rnd(64, N, seed=5)
    gemm(EPI["ROWADD"], A, W, o32, cg=cg, bias=bias, rowadd=radd, rowadd_mod=64)
    assert rel_err(o32, ref + radd[torch.arange(M, device="cuda") % 64]) < 2e-5


def test_window_remap_rows():
    """EPI_BIAS_BF16 with win_remap: token (b,y,x) lands at its row of the 14x14 window partition."""
    B, N, K = 2, 96, 64
    M = B * 4096
    A = rnd(M, K, seed=1).bfloat16(); W = rnd(N, K, scale=0.1, seed=2).bfloat16(); bias = rnd(N, seed=3)
    ref = (A.float() @ W.float().t() + bias).reshape(B, 64, 64, N)
    out = torch.full((B * 25 * 196, N), 99.0, device="cuda", dtype=torch.bfloat16)
    gemm(EPI["BIAS_BF16"], A, W, out, bias=bias, win_remap=1)
    padded = F.pad(ref, (0, 0, 0, 6, 0, 6), value=99.0)
    want = padded.reshape(B, 5, 14, 5, 14, N).permute(0, 1, 3, 2, 4, 5).reshape(B * 25 * 196, N)
    real = want != 99.0
    assert rel_err(out.float()[real], want[real]) < 6e-3
    assert (out.float()[~real] == 99.0).all()  # padded-token rows untouched


@pytest.mark.parametrize("cg", [1, 2])
def test_neck_epilogues(cg):
    """conv1x1 + LayerNorm2d -> zero-bordered bf16 map; 3x3 conv (implicit GEMM, 9 taps) + LayerNorm2d -> NCHW fp32
    (SA/modeling/image_encoder.py:88-104)."""
    B, D = 2, 128
    M = B * 4096
    x = rnd(M, D, seed=1).bfloat16(); w0 = rnd(256, D, scale=0.1, seed=2).bfloat16()
    lw = 1 + 0.1 * rnd(256, seed=3); lb = 0.1 * rnd(256, seed=4)
    pad = torch.zeros(B, 66, 66, 256, device="cuda", dtype=torch.bfloat16)
    gemm(EPI["LN_PAD"], x, w0, pad, cg=cg, ln_w=lw, ln_b=lb, eps=1e-6)
    ref1 = F.layer_norm(x.float() @ w0.float().t(), (256,), lw, lb, 1e-6).reshape(B, 64, 64, 256)
    assert rel_err(pad[:, 1:65, 1:65], ref1) < 6e-3
    border = pad.clone(); border[:, 1:65, 1:65] = 0
    assert (border == 0).all()
    w2 = rnd(256, 256, 3, 3, scale=0.03, seed=5).bfloat16()
    wk = w2.permute(0, 2, 3, 1).reshape(256, 9 * 256).contiguous()  # [o][ky][kx][c]
    out = torch.zeros(B, 256, 64, 64, device="cuda")
    gemm(EPI["LN_NCHW"], pad, wk, out, M=M, N=256, K=9 * 256, lda=256, cg=cg, ln_w=lw, ln_b=lb, eps=1e-6, conv3=1, conv_c=256)
    xin = pad[:, 1:65, 1:65].float().permute(0, 3, 1, 2)
    conv = F.conv2d(xin, w2.float(), padding=1)
    ref2 = F.layer_norm(conv.permute(0, 2, 3, 1), (256,), lw, lb, 1e-6).permute(0, 3, 1, 2)
    assert rel_err(out, ref2) < 5e-5


@pytest.mark.parametrize("cg", [1, 2])
def test_tf32_decoder_epilogues(cg):
    torch.backends.cuda.matmul.allow_tf32 = False
    n = 2
    M = n * 4096
    # N = 384 with 128-wide tiles + row-add table (merged k / v / q projections of the keys)
    A = rnd(M, 256, seed=1); W = rnd(384, 256, scale=0.06, seed=2); bias = rnd(384, seed=3); radd = rnd(4096, 384, seed=4)
    out = torch.zeros(M, 384, device="cuda")
    gemm(EPI["ROWADD"], A, W, out, cg=cg, tf32=1, bn=128, bias=bias, rowadd=radd, rowadd_mod=4096)
    ref = A @ W.t() + bias + radd.repeat(n, 1)
    assert rel_err(out, ref) < 2e-3
    gemm(EPI["BIAS_F32"], A, W, out, cg=cg, tf32=1, bn=128, bias=bias)
    assert rel_err(out, A @ W.t() + bias) < 2e-3
    # keys <- LayerNorm(keys + out_proj(a) + b), residual shared by all boxes (resid_mod) and in place
    a1 = rnd(M, 128, seed=5); wo = rnd(256, 128, scale=0.08, seed=6); bo = rnd(256, seed=7)
    keys0 = rnd(4096, 256, seed=8); lw = 1 + 0.1 * rnd(256, seed=9); lb = 0.1 * rnd(256, seed=10)
    keys = torch.zeros(M, 256, device="cuda")
    gemm(EPI["RESID_LN"], a1, wo, keys, cg=cg, tf32=1, bias=bo, resid=keys0, resid_mod=4096, ln_w=lw, ln_b=lb, eps=1e-5)
    ref = F.layer_norm(keys0.repeat(n, 1) + a1 @ wo.t() + bo, (256,), lw, lb, 1e-5)
    assert rel_err(keys, ref) < 2e-3
    k2 = ref.clone()
    gemm(EPI["RESID_LN"], a1, wo, k2, cg=cg, tf32=1, bias=bo, resid=k2, ln_w=lw, ln_b=lb, eps=1e-5)
    assert rel_err(k2, F.layer_norm(ref + a1 @ wo.t() + bo, (256,), lw, lb, 1e-5)) < 2e-3


@pytest.mark.parametrize("cg", [1, 2])
def test_upscale_epilogues(cg):
    """ConvT(256->64,k2,s2)+LayerNorm2d+GELU and ConvT(64->32,k2,s2)+GELU+hypernetwork dot (mask_decoder.py:53-59,136-144)."""
    torch.backends.cuda.matmul.allow_tf32 = False
    n = 2
    keys = rnd(n * 4096, 256, seed=1)
    ct0 = rnd(256, 64, 2, 2, scale=0.06, seed=2); b0 = rnd(64, seed=3)
    lw = 1 + 0.1 * rnd(64, seed=4); lb = 0.1 * rnd(64, seed=5)
    w0 = ct0.permute(2, 3, 1, 0).reshape(256, 256).contiguous()  # rows (dy,dx,o)
    up1 = torch.zeros(n, 128, 128, 64, device="cuda")
    gemm(EPI["UPSCALE1"], keys, w0, up1, cg=cg, tf32=1, bias=b0, ln_w=lw, ln_b=lb, eps=1e-6)
    x = keys.reshape(n, 64, 64, 256).permute(0, 3, 1, 2)
    r = F.conv_transpose2d(x, ct0, b0, stride=2).permute(0, 2, 3, 1)
    ref1 = F.gelu(F.layer_norm(r, (64,), lw, lb, 1e-6))
    assert rel_err(up1, ref1) < 3e-3
    ct3 = rnd(64, 32, 2, 2, scale=0.1, seed=6); b3 = rnd(32, seed=7); hyper = rnd(n, 32, seed=8)
    w3 = ct3.permute(2, 3, 1, 0).reshape(128, 64).contiguous()
    low = torch.zeros(n, 256, 256, device="cuda")
    gemm(EPI["UPSCALE2"], ref1.reshape(n * 16384, 64).contiguous(), w3, low, cg=cg, tf32=1, bn=128, bias=b3, hyper=hyper)
    u = F.gelu(F.conv_transpose2d(ref1.permute(0, 3, 1, 2), ct3, b3, stride=2))  # [n,32,256,256]
    ref2 = torch.einsum("nc,nchw->nhw", hyper, u)
    assert rel_err(low, ref2) < 3e-3


def gemm_f16(epi, A, W, out, bn=256, out_f16=0, bias=None, resid=None, rowadd=None, rowadd_mod=1, ln_w=None, ln_b=None, eps=1e-6,
             hyper=None, resid_mod=0, resid_f16=0, rowadd_f16=0):
    import ctypes
    from gpu_util import ptr, stream
    from inklayer_b200 import _lib
    M, K = A.shape
    N = W.shape[0]
    ldo = N
    _lib.check(_lib.load().inksam_op_gemm_f16(2, bn, epi, ptr(A), A.stride(0), ptr(W), W.stride(0), M, N, K, ptr(out), ldo, out_f16,
                                              ptr(bias), ptr(resid), ptr(rowadd), rowadd_mod, ptr(ln_w), ptr(ln_b), ctypes.c_float(eps),
                                              ptr(hyper), resid_mod, resid_f16, rowadd_f16, stream()))


def test_fp16_operand_decoder_epilogues():
    """The decoder's GEMM-to-GEMM intermediates are stored as fp16 (GemmArgs::ab_f16 / out_f16): the same epilogues as
    test_tf32_decoder_epilogues / test_upscale_epilogues with fp16 A and W (exact products, fp32 accumulation: against torch
    fp32 on the SAME fp16-rounded operands the error is accumulation order only) and fp16 outputs (one rounding, 2^-11)."""
    torch.backends.cuda.matmul.allow_tf32 = False
    n = 2
    M = n * 4096
    # final k|v projection: A = the final keys (fp16), row-add table, fp16 output
    A = rnd(M, 256, seed=1).half(); W = rnd(256, 256, scale=0.06, seed=2).half(); radd = rnd(4096, 256, seed=4)
    o16 = torch.zeros(M, 256, device="cuda", dtype=torch.float16)
    gemm_f16(EPI["ROWADD"], A, W, o16, out_f16=1, rowadd=radd, rowadd_mod=4096)
    ref = A.float() @ W.float().t() + radd.repeat(n, 1)
    assert rel_err(o16, ref) < 6e-4
    # keys <- LayerNorm(keys + out_proj(a) + b): fp16 attention output and weight, fp32 residual; fp32 and fp16 outputs
    a1 = rnd(M, 128, seed=5).half(); wo = rnd(256, 128, scale=0.08, seed=6).half(); bo = rnd(256, seed=7)
    keys0 = rnd(4096, 256, seed=8); lw = 1 + 0.1 * rnd(256, seed=9); lb = 0.1 * rnd(256, seed=10)
    keys = torch.zeros(M, 256, device="cuda")
    gemm_f16(EPI["RESID_LN"], a1, wo, keys, bias=bo, resid=keys0, resid_mod=4096, ln_w=lw, ln_b=lb, eps=1e-5)
    ref = F.layer_norm(keys0.repeat(n, 1) + a1.float() @ wo.float().t() + bo, (256,), lw, lb, 1e-5)
    assert rel_err(keys, ref) < 2e-5
    k16 = torch.zeros(M, 256, device="cuda", dtype=torch.float16)
    gemm_f16(EPI["RESID_LN"], a1, wo, k16, out_f16=1, bias=bo, resid=ref, ln_w=lw, ln_b=lb, eps=1e-5)
    ref2 = F.layer_norm(ref + a1.float() @ wo.float().t() + bo, (256,), lw, lb, 1e-5)
    assert rel_err(k16, ref2) < 6e-4
    assert torch.equal(k16, _same_as_fp32_then_half(a1, wo, bo, ref, lw, lb))
    # fp16 residual rows (the engine's layer-1 keys)
    r16 = ref.half()
    gemm_f16(EPI["RESID_LN"], a1, wo, k16, out_f16=1, bias=bo, resid=r16, resid_f16=1, ln_w=lw, ln_b=lb, eps=1e-5)
    ref2h = F.layer_norm(r16.float() + a1.float() @ wo.float().t() + bo, (256,), lw, lb, 1e-5)
    assert rel_err(k16, ref2h) < 6e-4
    # k|v|q projection with 128-wide tiles on fp16 keys
    Wq = rnd(384, 256, scale=0.06, seed=21).half(); radd3 = rnd(4096, 384, seed=22)
    o3 = torch.zeros(M, 384, device="cuda", dtype=torch.float16)
    gemm_f16(EPI["ROWADD"], A, Wq, o3, bn=128, out_f16=1, rowadd=radd3, rowadd_mod=4096)
    assert rel_err(o3, A.float() @ Wq.float().t() + radd3.repeat(n, 1)) < 6e-4
    r3h = radd3.half()  # fp16 row-add table
    gemm_f16(EPI["ROWADD"], A, Wq, o3, bn=128, out_f16=1, rowadd=r3h, rowadd_mod=4096, rowadd_f16=1)
    assert rel_err(o3, A.float() @ Wq.float().t() + r3h.float().repeat(n, 1)) < 6e-4
    # the two transposed convolutions
    keysh = rnd(M, 256, seed=11).half()
    ct0 = rnd(256, 64, 2, 2, scale=0.06, seed=12); b0 = rnd(64, seed=13)
    lw0 = 1 + 0.1 * rnd(64, seed=14); lb0 = 0.1 * rnd(64, seed=15)
    w0 = ct0.permute(2, 3, 1, 0).reshape(256, 256).contiguous().half()
    up1 = torch.zeros(n, 128, 128, 64, device="cuda", dtype=torch.float16)
    gemm_f16(EPI["UPSCALE1"], keysh, w0, up1, out_f16=1, bias=b0, ln_w=lw0, ln_b=lb0, eps=1e-6)
    ct0h = w0.float().reshape(2, 2, 64, 256).permute(3, 2, 0, 1).contiguous()
    x = keysh.float().reshape(n, 64, 64, 256).permute(0, 3, 1, 2)
    r = F.conv_transpose2d(x, ct0h, b0, stride=2).permute(0, 2, 3, 1)
    ref1 = F.gelu(F.layer_norm(r, (64,), lw0, lb0, 1e-6))
    assert rel_err(up1, ref1) < 8e-4
    ct3 = rnd(64, 32, 2, 2, scale=0.1, seed=16); b3 = rnd(32, seed=17); hyper = rnd(n, 32, seed=18)
    w3 = ct3.permute(2, 3, 1, 0).reshape(128, 64).contiguous().half()
    low = torch.zeros(n, 256, 256, device="cuda")
    gemm_f16(EPI["UPSCALE2"], up1.reshape(n * 16384, 64), w3, low, bn=128, bias=b3, hyper=hyper)
    ct3h = w3.float().reshape(2, 2, 32, 64).permute(3, 2, 0, 1).contiguous()
    u = F.gelu(F.conv_transpose2d(up1.float().permute(0, 3, 1, 2), ct3h, b3, stride=2))
    ref3 = torch.einsum("nc,nchw->nhw", hyper, u)
    assert rel_err(low, ref3) < 3e-5


def _same_as_fp32_then_half(a1, wo, bo, resid, lw, lb):
    """The fp16-output kernel must equal the fp32-output kernel rounded to fp16."""
    k32 = torch.zeros(a1.shape[0], 256, device="cuda")
    gemm_f16(EPI["RESID_LN"], a1, wo, k32, bias=bo, resid=resid, ln_w=lw, ln_b=lb, eps=1e-5)
    return k32.half()


def test_windowed_qkv_gemm_fills_the_padded_token_rows():
    """The qkv GEMM of a windowed block also writes the rows of the padded window tokens (bf16 of the bare bias): identical to
    the plain GEMM followed by inksam_op_fill_pad_rows."""
    from gpu_util import ptr, stream
    from inklayer_b200 import _lib
    lib = _lib.load()
    B, N, K = 2, 3840, 256
    A = rnd(B * 4096, K, seed=1).bfloat16(); W = rnd(N, K, scale=0.05, seed=2).bfloat16()
    bias = rnd(N, seed=3); pad_bias = rnd(N, seed=4)
    want = torch.full((B * 25 * 196, N), 99.0, device="cuda", dtype=torch.bfloat16)
    _lib.check(lib.inksam_op_gemm_qkv_windowed(ptr(A), K, ptr(W), K, B, N, K, ptr(want), ptr(bias), None, stream()))
    assert (want == 99.0).any()  # padded rows untouched without pad_bias
    _lib.check(lib.inksam_op_fill_pad_rows(ptr(want), ptr(pad_bias), B, N, stream()))
    got = torch.full((B * 25 * 196, N), 99.0, device="cuda", dtype=torch.bfloat16)
    _lib.check(lib.inksam_op_gemm_qkv_windowed(ptr(A), K, ptr(W), K, B, N, K, ptr(got), ptr(bias), ptr(pad_bias), stream()))
    assert torch.equal(got, want)
