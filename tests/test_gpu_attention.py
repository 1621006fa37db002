"""-m gpu: the tcgen05 attention kernels against the oracle's attention arithmetic (fp32 softmax over bf16-valued
q, k, v with the decomposed rel-pos bias), through the C ABI.  Output is bf16: tolerance 1.5e-2 of the output range
(P is rounded to bf16 before P.V, inputs are identical on both sides)."""
import pytest
import torch
import torch.nn.functional as F

from gpu_util import ptr, rel_err, rel_rms, stream
from inklayer_b200 import _lib
from oracle import sam_oracle as O

pytestmark = pytest.mark.gpu


def ref_attention(q, k, v, rel_h, rel_w, side):
    """q,k,v: [N, side*side, hd] fp32 -> [N, side*side, hd] (oracle.encoder_attention without the linears)."""
    hd = q.shape[-1]
    attn = (q * hd ** -0.5) @ k.transpose(-2, -1)
    rh, rw = O.rel_pos_bias_terms(q, rel_h, rel_w, side)
    S = side
    attn = (attn.view(-1, S, S, S, S) + rh[:, :, :, :, None] + rw[:, :, :, None, :]).view(-1, S * S, S * S)
    return attn.softmax(dim=-1) @ v


def rel_table(rel_h, rel_w, rows):
    t = torch.zeros(rows, rel_h.shape[1])
    t[: rel_h.shape[0]] = rel_h
    t[rows // 2: rows // 2 + rel_w.shape[0]] = rel_w
    return t.to(torch.bfloat16).cuda()


@pytest.mark.parametrize("hd,heads,B", [(64, 2, 1), (80, 2, 1), (80, 3, 2)])
def test_global_attention(hd, heads, B):
    lib = _lib.load()
    g = torch.Generator().manual_seed(hd + heads)
    D = heads * hd
    qkv = (torch.randn(B * 4096, 3 * D, generator=g) * 1.5).to(torch.bfloat16)
    rel_h = (torch.randn(127, hd, generator=g) * 0.3).to(torch.bfloat16).float()
    rel_w = (torch.randn(127, hd, generator=g) * 0.3).to(torch.bfloat16).float()
    out = torch.zeros(B * 4096, D, dtype=torch.bfloat16, device="cuda")
    qd = qkv.cuda()
    _lib.check(lib.inksam_op_attention(0, ptr(qd), ptr(rel_table(rel_h, rel_w, 256)), ptr(out), B, heads, hd, stream()))
    torch.cuda.synchronize()
    t = qkv.float().cuda().reshape(B, 4096, 3, heads, hd).permute(2, 0, 3, 1, 4).reshape(3, B * heads, 4096, hd)
    ref = ref_attention(t[0], t[1], t[2], rel_h.cuda(), rel_w.cuda(), 64)
    ref = ref.reshape(B, heads, 4096, hd).permute(0, 2, 1, 3).reshape(B * 4096, D)
    assert torch.isfinite(out.float()).all()
    assert rel_rms(out, ref) < 1e-2
    assert rel_err(out, ref) < 1.5e-2


@pytest.mark.parametrize("hd,peak_at,gain", [(80, 3000, 50.0), (64, 70, 50.0), (80, 4095, 50.0),
                                             (80, 3000, 170.0), (64, 40, 170.0), (80, 4095, 400.0), (64, 2047, 400.0)])
def test_global_attention_late_peak_rescales(hd, peak_at, gain):
    """The lazy softmax shift comes from the first 32 keys.  A later key ~40 log2 units above it (gain 50) must trigger
    the exact rescale of the running sum and of the O accumulator in TMEM; a key 137 / 320+ log2 units above it (gain
    170 / 400) is beyond the 2^120 exponent clamp of the hot loop and must take the cold path that raises the shift to
    the sub-tile's true maximum BEFORE its probabilities are used (the clamp must never bite silently).  A
    second, lower peak 1000 keys earlier makes the shift move twice."""
    lib = _lib.load()
    heads, B = 2, 1
    g = torch.Generator().manual_seed(7 + hd + peak_at)
    D = heads * hd
    u = torch.ones(hd) / hd ** 0.5
    t = torch.randn(3, B, heads, 4096, hd, generator=g) * 1.0
    t[0] += 5.0 * u                       # every query has a component along u
    t[1, :, :, peak_at] += gain * u       # one key far along u: score about 5 * gain / sqrt(hd) nats above the rest
    if peak_at >= 1000:
        t[1, :, :, peak_at - 1000] += 0.5 * gain * u
    qkv = t.permute(1, 3, 0, 2, 4).reshape(B * 4096, 3 * D).to(torch.bfloat16)
    rel_h = (torch.randn(127, hd, generator=g) * 0.1).to(torch.bfloat16).float()
    rel_w = (torch.randn(127, hd, generator=g) * 0.1).to(torch.bfloat16).float()
    out = torch.zeros(B * 4096, D, dtype=torch.bfloat16, device="cuda")
    _lib.check(lib.inksam_op_attention(0, ptr(qkv.cuda()), ptr(rel_table(rel_h, rel_w, 256)), ptr(out), B, heads, hd, stream()))
    torch.cuda.synchronize()
    tt = qkv.float().cuda().reshape(B, 4096, 3, heads, hd).permute(2, 0, 3, 1, 4).reshape(3, B * heads, 4096, hd)
    ref = ref_attention(tt[0], tt[1], tt[2], rel_h.cuda(), rel_w.cuda(), 64)
    ref = ref.reshape(B, heads, 4096, hd).permute(0, 2, 1, 3).reshape(B * 4096, D)
    assert torch.isfinite(out.float()).all()
    assert rel_rms(out, ref) < 1e-2
    assert rel_err(out, ref) < 1.5e-2


@pytest.fixture(params=["v5", "v5-kpoly8", "v4", "v3", "v2"])
def window_kernel(request, monkeypatch):
    """The windowed kernel variants (read per launch by attn_host.cu): the current one with its default / no / double
    share of polynomial exp2, and the previous kernel as a cross-check."""
    monkeypatch.delenv("INKSAM_WINDOW_ATTN", raising=False)
    monkeypatch.delenv("INKSAM_WINDOW_KPOLY", raising=False)
    monkeypatch.setenv("INKSAM_WINDOW_ATTN", request.param[1])
    if "-kpoly" in request.param:
        monkeypatch.setenv("INKSAM_WINDOW_KPOLY", request.param.split("-kpoly")[1])
    return request.param


@pytest.mark.parametrize("hd,heads,B", [(64, 2, 1), (80, 2, 1), (80, 3, 2), (80, 16, 1), (64, 12, 3)])
def test_window_attention(hd, heads, B, window_kernel):
    """Windowed blocks: 14x14 windows over the zero-padded 70x70 grid; padded tokens carry the qkv bias and are NOT
    masked (SA/modeling/image_encoder.py:170-177,243-289)."""
    lib = _lib.load()
    g = torch.Generator().manual_seed(100 + hd + heads)
    D = heads * hd
    nat = (torch.randn(B, 64, 64, 3 * D, generator=g) * 1.5).to(torch.bfloat16).float()
    bias = (torch.randn(3 * D, generator=g)).to(torch.bfloat16).float()
    rel_h = (torch.randn(27, hd, generator=g) * 0.3).to(torch.bfloat16).float()
    rel_w = (torch.randn(27, hd, generator=g) * 0.3).to(torch.bfloat16).float()
    padded = bias.expand(B, 70, 70, 3 * D).clone()
    padded[:, :64, :64] = nat
    win = padded.reshape(B, 5, 14, 5, 14, 3 * D).permute(0, 1, 3, 2, 4, 5).reshape(B * 25, 196, 3 * D)
    qd = win.reshape(B * 25 * 196, 3 * D).to(torch.bfloat16).cuda()
    out = torch.zeros(B * 4096, D, dtype=torch.bfloat16, device="cuda")
    _lib.check(lib.inksam_op_attention(1, ptr(qd), ptr(rel_table(rel_h, rel_w, 64)), ptr(out), B, heads, hd, stream()))
    torch.cuda.synchronize()
    t = win.cuda().reshape(B * 25, 196, 3, heads, hd).permute(2, 0, 3, 1, 4).reshape(3, B * 25 * heads, 196, hd)
    ref = ref_attention(t[0], t[1], t[2], rel_h.cuda(), rel_w.cuda(), 14)
    ref = ref.reshape(B * 25, heads, 14, 14, hd).permute(0, 2, 3, 1, 4).reshape(B * 25, 14, 14, D)
    ref = O.window_unpartition(ref, 14, 70, 64).reshape(B * 4096, D)
    assert torch.isfinite(out.float()).all()
    assert rel_rms(out, ref) < 1e-2
    assert rel_err(out, ref) < 1.5e-2


@pytest.mark.parametrize("hd,peak_key,gain", [(80, 150, 170.0), (64, 195, 400.0), (80, 40, 50.0), (64, 97, 170.0), (80, 98, 400.0)])
@pytest.mark.parametrize("version", ["4", "5"])
def test_window_attention_late_peak(hd, peak_key, gain, version, monkeypatch):
    """Windowed blocks with two keys of every window far above the others (beyond the first 32 keys and, for gain >= 170,
    more than 120 log2 units above them, 7 log2 units apart): the softmax must stay exact -- the reference subtracts the
    true row maximum.  (A kernel that clamps the exponent gives both keys the same weight and fails this.)"""
    lib = _lib.load()
    monkeypatch.setenv("INKSAM_WINDOW_ATTN", version)
    heads, B = 2, 1
    g = torch.Generator().manual_seed(300 + hd + peak_key)
    D = heads * hd
    u = torch.ones(hd) / hd ** 0.5
    t = torch.randn(3, B * 25, heads, 196, hd, generator=g)
    t[0] += 5.0 * u
    t[1, :, :, peak_key] += gain * u
    t[1, :, :, (peak_key + 61) % 196] += 0.95 * gain * u   # a second peak in the other key half, a few log2 units lower
    win = t.permute(1, 3, 0, 2, 4).reshape(B * 25, 196, 3 * D).to(torch.bfloat16).float()
    rel_h = (torch.randn(27, hd, generator=g) * 0.3).to(torch.bfloat16).float()
    rel_w = (torch.randn(27, hd, generator=g) * 0.3).to(torch.bfloat16).float()
    qd = win.reshape(B * 25 * 196, 3 * D).to(torch.bfloat16).cuda()
    out = torch.zeros(B * 4096, D, dtype=torch.bfloat16, device="cuda")
    _lib.check(lib.inksam_op_attention(1, ptr(qd), ptr(rel_table(rel_h, rel_w, 64)), ptr(out), B, heads, hd, stream()))
    torch.cuda.synchronize()
    tt = win.cuda().reshape(B * 25, 196, 3, heads, hd).permute(2, 0, 3, 1, 4).reshape(3, B * 25 * heads, 196, hd)
    ref = ref_attention(tt[0], tt[1], tt[2], rel_h.cuda(), rel_w.cuda(), 14)
    ref = ref.reshape(B * 25, heads, 14, 14, hd).permute(0, 2, 3, 1, 4).reshape(B * 25, 14, 14, D)
    ref = O.window_unpartition(ref, 14, 70, 64).reshape(B * 4096, D)
    assert torch.isfinite(out.float()).all()
    assert rel_rms(out, ref) < 1e-2
    assert rel_err(out, ref) < 1.5e-2
