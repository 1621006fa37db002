"""-m gpu: the HBM-bound helper kernels against the CPU oracle / plain torch fp32, through the C ABI."""
import ctypes

import numpy as np
import pytest
import torch

from gpu_util import ptr, rel_err, stream
from inklayer_b200 import _lib
from oracle import sam_oracle as O

pytestmark = pytest.mark.gpu
CFG = O.CONFIGS["vit_tiny"]


@pytest.mark.parametrize("M,D", [(4096 + 3, 1280), (515, 160), (1000, 128), (64, 768)])
def test_layernorm_matches_torch(M, D):
    lib = _lib.load()
    g = torch.Generator().manual_seed(M + D)
    x = (torch.randn(M, D, generator=g) * 3 + 0.5).cuda()
    w = (1 + 0.1 * torch.randn(D, generator=g)).cuda()
    b = (0.1 * torch.randn(D, generator=g)).cuda()
    out = torch.empty(M, D, dtype=torch.bfloat16, device="cuda")
    _lib.check(lib.inksam_op_layernorm(ptr(x), ptr(w), ptr(b), ptr(out), M, D, ctypes.c_float(1e-6), 0, stream()))
    ref = torch.nn.functional.layer_norm(x, (D,), w, b, 1e-6)
    assert rel_err(out, ref) < 6e-3  # one bf16 rounding
    _lib.check(lib.inksam_op_layernorm(ptr(x), None, None, ptr(out), M, D, ctypes.c_float(0), 1, stream()))
    assert torch.equal(out, x.to(torch.bfloat16))


@pytest.mark.parametrize("h,w,dtype", [(1024, 1024, 0), (819, 1024, 0), (1024, 683, 1)])
def test_patchify_matches_preprocess(h, w, dtype):
    """Sam.preprocess + PatchEmbed unfold (oracle.preprocess / oracle.patch_embed column order)."""
    lib = _lib.load()
    rng = np.random.default_rng(h * 7 + w)
    B = 2
    img = rng.integers(0, 256, (B, h, w, 3), dtype=np.uint8)
    chw = torch.from_numpy(img).permute(0, 3, 1, 2).contiguous()
    x = O.preprocess(CFG, chw)  # [B,3,1024,1024]
    ref = x.reshape(B, 3, 64, 16, 64, 16).permute(0, 2, 4, 1, 3, 5).reshape(B * 4096, 768)
    if dtype == 0:
        dev = torch.from_numpy(img).cuda()
        strides = (h * w * 3, w * 3, 0)
    else:
        dev = chw.float().cuda()
        strides = (3 * h * w * 4, w * 4, h * w * 4)
    hw = (ctypes.c_int * (2 * B))(*([h, w] * B))
    mean = (ctypes.c_float * 3)(*CFG.pixel_mean)
    std = (ctypes.c_float * 3)(*CFG.pixel_std)
    out = torch.empty(B * 4096, 768, dtype=torch.bfloat16, device="cuda")
    _lib.check(lib.inksam_op_patchify(ptr(dev), dtype, B, hw, strides[0], strides[1], strides[2], mean, std, ptr(out), stream()))
    got = out.float().cpu()
    assert (got - ref).abs().max().item() < 0.012  # |x| <= 2.7, bf16 rounding
    pad = ref == 0
    assert torch.equal(got[pad], torch.zeros_like(got[pad]))  # padding is exactly zero (normalise THEN pad)


def test_fill_pad_rows():
    lib = _lib.load()
    B, cols = 2, 480
    q = torch.full((B * 25 * 196, cols), 7.0, dtype=torch.bfloat16, device="cuda")
    bias = torch.randn(cols, generator=torch.Generator().manual_seed(1)).cuda()
    _lib.check(lib.inksam_op_fill_pad_rows(ptr(q), ptr(bias), B, cols, stream()))
    q = q.float().cpu().reshape(B, 5, 5, 14, 14, cols)
    bb = bias.to(torch.bfloat16).float().cpu()
    n_pad = 0
    for wy in range(5):
        for wx in range(5):
            for ly in range(14):
                for lx in range(14):
                    real = wy * 14 + ly < 64 and wx * 14 + lx < 64
                    row = q[:, wy, wx, ly, lx]
                    if real:
                        assert (row == 7.0).all()
                    else:
                        n_pad += 1
                        assert torch.equal(row, bb.expand_as(row))
    assert n_pad == 4900 - 4096


@pytest.mark.parametrize("inp,orig", [((1024, 1024), (1024, 1024)), ((819, 1024), (600, 750)), ((1024, 1024), (750, 750)),
                                      ((1024, 577), (1333, 751)), ((768, 1024), (768, 1024)), ((1024, 683), (1024, 683)),
                                      ((1024, 640), (2048, 1280)), ((3, 1024), (3, 1024)), ((1024, 16), (1024, 16))])
def test_postprocess_matches_interpolate(inp, orig):
    """Fused two-stage bilinear + threshold vs Sam.postprocess_masks (oracle.postprocess_masks): masks + logits (general
    path), masks alone (the 4x16-patch fast path when original == input size and rows are 16-byte aligned) and masks
    written at a misaligned base address."""
    lib = _lib.load()
    K = 3
    low = torch.randn(K, 1, 256, 256, generator=torch.Generator().manual_seed(5)) * 0.1
    ref = O.postprocess_masks(CFG, low, inp, orig)
    H, W = orig
    d_low = low.cuda()
    masks = torch.empty(K, 1, H, W, dtype=torch.uint8, device="cuda")
    logits = torch.empty(K, 1, H, W, dtype=torch.float32, device="cuda")
    _lib.check(lib.inksam_op_postprocess(ptr(d_low), ptr(masks), ptr(logits), K, inp[0], inp[1], H, W,
                                         ctypes.c_float(0.0), stream()))
    assert (logits.cpu() - ref).abs().max().item() < 2e-6
    m_ref = ref > 0.0

    def check_masks(m):
        disagree = (m.cpu().bool() != m_ref)
        assert (ref[disagree].abs() < 2e-6).all()  # only where the logit is within rounding of the threshold
        assert disagree.float().mean().item() < 1e-4
        assert set(m.unique().tolist()) <= {0, 1}

    check_masks(masks)
    only = torch.full((K, 1, H, W), 7, dtype=torch.uint8, device="cuda")
    _lib.check(lib.inksam_op_postprocess(ptr(d_low), ptr(only), None, K, inp[0], inp[1], H, W, ctypes.c_float(0.0), stream()))
    check_masks(only)
    diff = (only != masks).cpu()
    assert (ref[diff].abs() < 2e-6).all()  # the two code paths may contract FMAs differently
    buf = torch.full((K * H * W + 32,), 7, dtype=torch.uint8, device="cuda")
    for shift in (1, 6):
        buf.fill_(7)
        _lib.check(lib.inksam_op_postprocess(ptr(d_low), buf.data_ptr() + shift, None, K, inp[0], inp[1], H, W,
                                             ctypes.c_float(0.0), stream()))
        assert torch.equal(buf[shift:shift + K * H * W].view(K, 1, H, W), masks)
        assert (buf[:shift] == 7).all() and (buf[shift + K * H * W:] == 7).all()  # nothing written outside


@pytest.mark.parametrize("inp,orig", [((1024, 1024), (1024, 1024)), ((819, 1024), (600, 750)), ((1024, 1024), (750, 750)),
                                      ((1024, 577), (1333, 751)), ((1024, 683), (300, 200)), ((1024, 640), (2048, 1280)),
                                      ((1024, 1001), (1024, 1001)), ((3, 1024), (3, 1024)), ((1024, 16), (1024, 16)),
                                      ((1024, 5), (17, 5)), ((1024, 1024), (33, 1031))])
def test_postprocess_packed_is_packbits_of_the_byte_masks(inp, orig):
    """The bit-packed output format (one bit per pixel, MSB first, rows padded to whole bytes -- numpy.packbits / PIL mode
    "1" rows, InkLayer/runner.py:51) holds exactly the bits of the byte masks, for the fast and the general path, odd
    widths, at a misaligned base address, and nothing is written outside [K][H][ceil(W/8)]."""
    import numpy as np

    lib = _lib.load()
    K = 3
    low = (torch.randn(K, 1, 256, 256, generator=torch.Generator().manual_seed(9)) * 0.1).cuda()
    H, W = orig
    pitch = (W + 7) // 8
    masks = torch.empty(K, H, W, dtype=torch.uint8, device="cuda")
    _lib.check(lib.inksam_op_postprocess(ptr(low), ptr(masks), None, K, inp[0], inp[1], H, W, ctypes.c_float(0.0), stream()))
    want = np.packbits(masks.cpu().numpy().astype(bool), axis=-1)
    assert want.shape == (K, H, pitch)
    for shift in (0, 16, 3):
        buf = torch.full((K * H * pitch + 64,), 0xA5, dtype=torch.uint8, device="cuda")
        _lib.check(lib.inksam_op_postprocess_packed(ptr(low), buf.data_ptr() + shift, K, inp[0], inp[1], H, W,
                                                    ctypes.c_float(0.0), stream()))
        got = buf[shift:shift + K * H * pitch].view(K, H, pitch).cpu().numpy()
        np.testing.assert_array_equal(got, want)
        assert (buf[:shift] == 0xA5).all() and (buf[shift + K * H * pitch:] == 0xA5).all()
    assert np.array_equal(np.unpackbits(want, axis=-1, count=W), masks.cpu().numpy())


@pytest.mark.parametrize("hw", [(750, 750), (600, 750), (300, 200), (1333, 751), (2000, 1500), (64, 1000), (37, 53), (1025, 1024),
                                (1, 7), (777, 3001), (1024, 512)])
def test_resize_u8_bit_exact_with_pillow_oracle(hw):
    """inksam_resize_u8 (ResizeLongestSide.apply_image on the device) into a 1024x1024x3 container vs the Pillow
    restatement (oracle/resize_oracle.py): identical bytes, nothing written outside the [h, w] window."""
    from oracle import resize_oracle as R

    lib = _lib.load()
    img = np.random.default_rng(hw[0] * 7 + hw[1]).integers(0, 256, hw + (3,), dtype=np.uint8)
    ref = R.apply_image(img)
    h, w = ref.shape[:2]
    src = torch.from_numpy(img).cuda()
    dst = torch.full((1024, 1024, 3), 9, dtype=torch.uint8, device="cuda")
    need = lib.inksam_resize_workspace_bytes(hw[0], hw[1], h, w)
    ws = torch.empty(max(int(need), 1), dtype=torch.uint8, device="cuda")
    _lib.check(lib.inksam_resize_u8(ptr(src), hw[0], hw[1], hw[1] * 3, ptr(dst), h, w, 1024 * 3, ptr(ws), stream()))
    got = dst.cpu().numpy()
    assert np.array_equal(got[:h, :w], ref)
    assert (got[h:] == 9).all() and (got[:, w:] == 9).all()


def test_sam_resize_longest_side_and_errors():
    from inklayer_b200 import CONFIGS, build_sam_from_state_dict
    from inklayer_b200.weights import synthetic_state_dict
    from oracle import resize_oracle as R

    cfg = CONFIGS["vit_tiny"]
    sam = build_sam_from_state_dict(synthetic_state_dict(cfg, 0), cfg).to("cuda")
    img = np.random.default_rng(11).integers(0, 256, (500, 380, 3), dtype=np.uint8)
    got = sam.resize_longest_side(torch.from_numpy(img).cuda())
    assert np.array_equal(got.cpu().numpy(), R.apply_image(img))
    same = torch.from_numpy(np.zeros((1024, 700, 3), np.uint8)).cuda()
    assert sam.resize_longest_side(same) is same  # already in the input frame
    with pytest.raises(RuntimeError):
        sam.resize_longest_side(torch.zeros(3, 10, 10, dtype=torch.uint8, device="cuda"))
    lib = _lib.load()
    assert lib.inksam_resize_u8(None, 10, 10, 30, None, 5, 5, 15, None, stream()) != 0


def _attn_ref(q, k, v):
    """8 heads x 16, softmax(q k^T / sqrt(16)) v in float64.  q [..., Tq, 128], k / v [..., Tk, 128]."""
    sh = q.shape[:-2]
    qh = q.double().reshape(*sh, -1, 8, 16).transpose(-2, -3)
    kh = k.double().reshape(*sh, -1, 8, 16).transpose(-2, -3)
    vh = v.double().reshape(*sh, -1, 8, 16).transpose(-2, -3)
    p = torch.softmax(qh @ kh.transpose(-1, -2) / 4.0, dim=-1)
    return (p @ vh).transpose(-2, -3).reshape(*sh, -1, 128)


@pytest.mark.parametrize("half", [False, True])
@pytest.mark.parametrize("gain,shared", [(1.0, False), (1.0, True), (12.0, False), (60.0, False), (60.0, True)])
def test_decoder_token_to_image_attention(gain, shared, half):
    """decoder_t2i_kernel against float64 torch: 7 queries x 4096 keys x 8 heads per box.  Large gains make the logits span
    thousands of log2 units with the largest keys late in each key range, so the lazy softmax shift (set by a warp's first key,
    moved when a score exceeds it by 2^24) has to move many times; `shared`: all boxes read the keys of their image."""
    lib = _lib.load()
    boxes, ld = 5, 384
    g = torch.Generator().manual_seed(11)
    q = (torch.randn(boxes, 7, 128, generator=g) * gain).cuda()
    nimg = 2 if shared else boxes
    kv = torch.randn(nimg, 4096, ld, generator=g)
    ramp = torch.linspace(0.2, 2.0, 4096).view(1, 4096, 1)      # later keys are larger
    kv[:, :, :128] *= ramp
    kv = kv.cuda()
    kv_dev = kv.half() if half else kv  # fp16 storage (the engine's per-box k|v|q): the reference sees the same rounded values
    kv = kv_dev.float()
    box_img = torch.tensor([0, 1, 1, 0, 1], dtype=torch.int32, device="cuda") if shared else None
    out = torch.full((boxes, 7, 128), float("nan"), device="cuda")
    scratch = torch.empty(int(lib.inksam_op_decoder_t2i_scratch_bytes(boxes)), dtype=torch.uint8, device="cuda")
    for _ in range(2):  # the ticket counters must be left ready for the next launch
        _lib.check(lib.inksam_op_decoder_t2i(ptr(q), ptr(kv_dev), int(half), 4096 * ld, ld, ptr(box_img), ptr(out), ptr(scratch), boxes,
                                             stream()))
    sel = box_img.long() if shared else torch.arange(boxes, device="cuda")
    ref = _attn_ref(q, kv[sel][:, :, :128], kv[sel][:, :, 128:256])
    assert torch.isfinite(out).all()
    err = (out.double() - ref).abs().max().item() / ref.abs().max().item()
    assert err < 2e-5, err


@pytest.mark.parametrize("half,out_half", [(False, False), (True, False), (False, True), (True, True)])
@pytest.mark.parametrize("gain,shared", [(1.0, False), (40.0, False), (40.0, True)])
def test_decoder_image_to_token_attention(gain, shared, half, out_half):
    """decoder_i2t_kernel against float64 torch: every image token attends to the 7 tokens."""
    lib = _lib.load()
    boxes, ld, qcol = 3, 384, 256
    g = torch.Generator().manual_seed(12)
    nimg = 2 if shared else boxes
    qsrc = (torch.randn(nimg, 4096, ld, generator=g) * gain).cuda()
    q_dev = qsrc.half() if half else qsrc
    qsrc = q_dev.float()
    kvt = torch.randn(boxes, 2, 7, 128, generator=g).cuda()
    box_img = torch.tensor([1, 0, 1], dtype=torch.int32, device="cuda") if shared else None
    out = torch.full((boxes * 4096, 128), float("nan"), device="cuda", dtype=torch.float16 if out_half else torch.float32)
    _lib.check(lib.inksam_op_decoder_i2t(ptr(q_dev), int(half), 4096 * ld, ld, qcol, ptr(box_img), ptr(kvt), ptr(out), int(out_half), boxes,
                                         stream()))
    sel = box_img.long() if shared else torch.arange(boxes, device="cuda")
    ref = _attn_ref(qsrc[sel][:, :, qcol:qcol + 128], kvt[:, 0], kvt[:, 1]).reshape(boxes * 4096, 128)
    assert torch.isfinite(out).all()
    err = (out.double() - ref).abs().max().item() / ref.abs().max().item()
    assert err < (6e-4 if out_half else 2e-5), err  # fp16 output: one rounding (2^-11)


@pytest.mark.parametrize("cg", [1, 2])
@pytest.mark.parametrize("M,N,K,bn", [(4096 * 3, 384, 256, 128), (4096 + 77, 256, 256, 256), (300, 160, 160, 256)])
def test_gemm_rowadd_fp16_output(cg, M, N, K, bn):
    """The decoder's k|v|q projection (tf32 GEMM + row-add table) storing fp16: equal to the fp32-output kernel rounded to fp16
    (round to nearest even), saturating instead of overflowing; rows beyond M untouched."""
    from gpu_util import EPI, gemm
    lib = _lib.load()
    g = torch.Generator().manual_seed(5)
    A = torch.randn(M, K, generator=g).cuda(); W = (torch.randn(N, K, generator=g) * 0.05).cuda()
    radd = torch.randn(64, N, generator=g).cuda(); bias = torch.randn(N, generator=g).cuda()
    radd[3, 7] = 1e6      # beyond the fp16 range: saturates to 65504
    o32 = torch.zeros(M, N, device="cuda")
    gemm(EPI["ROWADD"], A, W, o32, cg=cg, tf32=1, bn=bn, bias=bias, rowadd=radd, rowadd_mod=64)
    o16 = torch.full((M + 5, N), 7.0, device="cuda", dtype=torch.float16)
    _lib.check(lib.inksam_op_gemm_rowadd_f16(cg, 1, bn, ptr(A), K, ptr(W), K, M, N, K, ptr(o16), N, ptr(bias), ptr(radd), 64, stream()))
    want = o32.clamp(-65504.0, 65504.0).half()
    assert torch.equal(o16[:M], want)
    assert (o16[M:] == 7.0).all()
    assert o16[3, 7].item() == 65504.0
