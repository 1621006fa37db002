"""CPU numerics probe for the next optimisation: folding LayerNorm into the GEMMs around it.

Compares, on the oracle's fp32 encoder with bf16 rounding emulated on the LN -> Linear pairs only,
  (a) today's scheme:   Linear(bf16(LN(x)), bf16(W)) + b
  (b) the folded scheme: rstd * (bf16(x) @ bf16(gamma * W)^T) - rstd * mean * colsum(bf16(gamma * W)) + (b + W beta)
against the pure fp32 result.  Test infrastructure: imports oracle/, never used by the product."""
import os
import sys

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from inklayer_b200.config import CONFIGS  # noqa: E402
from inklayer_b200.synth import synthetic_sketch  # noqa: E402
from inklayer_b200.weights import synthetic_state_dict  # noqa: E402
from oracle import sam_oracle as O  # noqa: E402


def bf(x):
    return x.to(torch.bfloat16).float()


def ln_linear(mode, x, g, b, W, bias):
    if mode == "fp32":
        return F.linear(O._ln(x, g, b, 1e-6), W, bias)
    if mode == "today":
        return F.linear(bf(O._ln(x, g, b, 1e-6)), bf(W), bias)
    mean = x.mean(-1, keepdim=True)
    var = (x * x).mean(-1, keepdim=True) - mean * mean          # E[x^2] - mean^2 as the epilogue would compute it
    rstd = torch.rsqrt(var + 1e-6)
    Wp = bf(W * g[None, :])
    acc = F.linear(bf(x), Wp)
    return rstd * acc - rstd * mean * Wp.sum(1)[None] + (bias + W @ b)


def encoder(cfg, sd, x, mode):
    t = O.patch_embed(cfg, sd, x)
    if "image_encoder.pos_embed" in sd:
        t = t + sd["image_encoder.pos_embed"]
    ratios = []
    for i in range(cfg.depth):
        pre = f"image_encoder.blocks.{i}."
        ratios.append(float((t.mean(-1).abs() / t.std(-1)).mean()))
        h = t
        is_global = i in cfg.global_attn_indexes
        if is_global:
            qkv = ln_linear(mode, h, sd[pre + "norm1.weight"], sd[pre + "norm1.bias"], sd[pre + "attn.qkv.weight"], sd[pre + "attn.qkv.bias"])
            a = attention_from_qkv(cfg, sd, pre + "attn.", qkv)
        else:
            qkv = ln_linear(mode, h, sd[pre + "norm1.weight"], sd[pre + "norm1.bias"], sd[pre + "attn.qkv.weight"], sd[pre + "attn.qkv.bias"])
            # the reference zero-pads AFTER norm1: padded tokens carry qkv = bias
            H = qkv.shape[1]
            pad = (cfg.window_size - H % cfg.window_size) % cfg.window_size
            full = sd[pre + "attn.qkv.bias"].expand(qkv.shape[0], H + pad, H + pad, -1).clone()
            full[:, :H, :H] = qkv
            n = (H + pad) // cfg.window_size
            win = full.view(-1, n, cfg.window_size, n, cfg.window_size, full.shape[-1]).permute(0, 1, 3, 2, 4, 5)
            win = win.reshape(-1, cfg.window_size, cfg.window_size, full.shape[-1])
            a = attention_from_qkv(cfg, sd, pre + "attn.", win)
            a = O.window_unpartition(a, cfg.window_size, H + pad, H)
        t = t + a
        hmid = ln_linear(mode, t, sd[pre + "norm2.weight"], sd[pre + "norm2.bias"], sd[pre + "mlp.lin1.weight"], sd[pre + "mlp.lin1.bias"])
        t = t + F.linear(F.gelu(hmid), sd[pre + "mlp.lin2.weight"], sd[pre + "mlp.lin2.bias"])
    return O.neck(cfg, sd, t), ratios


def attention_from_qkv(cfg, sd, pre, qkv):
    N, S, _, D3 = qkv.shape
    D = D3 // 3
    nh, hd = cfg.num_heads, cfg.head_dim
    q3 = qkv.reshape(N, S * S, 3, nh, hd).permute(2, 0, 3, 1, 4).reshape(3, N * nh, S * S, hd)
    q, k, v = q3[0], q3[1], q3[2]
    attn = (q * hd ** -0.5) @ k.transpose(-2, -1)
    rel_h, rel_w = O.rel_pos_bias_terms(q, sd[pre + "rel_pos_h"], sd[pre + "rel_pos_w"], S)
    attn = (attn.view(-1, S, S, S, S) + rel_h[:, :, :, :, None] + rel_w[:, :, :, None, :]).view(-1, S * S, S * S)
    out = (attn.softmax(-1) @ v).view(N, nh, S, S, hd).permute(0, 2, 3, 1, 4).reshape(N, S, S, D)
    return F.linear(out, sd[pre + "proj.weight"], sd[pre + "proj.bias"])


@torch.no_grad()
def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "vit_b"
    cfg_o, cfg = O.CONFIGS[name], CONFIGS[name]
    sd = synthetic_state_dict(cfg, 0)
    img = torch.from_numpy(synthetic_sketch(0)).permute(2, 0, 1).float()[None]
    x = O.preprocess(cfg_o, img)
    ref, ratios = encoder(cfg_o, sd, x, "fp32")
    print(f"{name}: |token mean| / token std of the residual stream per block: " + " ".join(f"{r:.2f}" for r in ratios))
    for mode in ("today", "folded"):
        out, _ = encoder(cfg_o, sd, x, mode)
        err = ((out - ref).double().pow(2).mean().sqrt() / ref.double().pow(2).mean().sqrt()).item()
        print(f"{name} {mode:7s}: embedding rel-RMS vs fp32 = {err:.3e}")


if __name__ == "__main__":
    main()
