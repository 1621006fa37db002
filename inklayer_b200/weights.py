"""Synthetic SAM weights with the reference's state_dict layout.

There are no checkpoints on this box, so benches and tests use seeded random weights.
Every tensor is drawn on the CPU from its own generator (seed = crc32(key) ^ seed), so the GPU box, this
container and the reference model loaded through `load_state_dict` all see bit-identical values.

Unlike the reference's default init (build_sam.py with checkpoint=None), nothing is degenerate here:
pos_embed and rel_pos_* are non-zero and the LayerNorm affines are not (1, 0), so parity exercises
positional embedding, the decomposed rel-pos bias and the LN affine (reference zero-inits them:
segment_anything/modeling/image_encoder.py:65-70,221-222).
"""
from __future__ import annotations

import math
import zlib
from typing import Dict, List, Tuple

import torch

from .config import SamConfig


def state_dict_shapes(cfg: SamConfig) -> List[Tuple[str, Tuple[int, ...], str]]:
    """(key, shape, kind) for every tensor of the reference state_dict.  kind: w<fan_in> | b<fan_in> |
    emb | ln_w | ln_b | pos."""
    D, g, hd, p = cfg.embed_dim, cfg.grid, cfg.head_dim, cfg.patch_size
    E = cfg.out_chans
    out: List[Tuple[str, Tuple[int, ...], str]] = []

    def lin(pre: str, o: int, i: int) -> None:
        out.append((pre + ".weight", (o, i), f"w{i}"))
        out.append((pre + ".bias", (o,), f"b{i}"))

    def ln(pre: str, n: int) -> None:
        out.append((pre + ".weight", (n,), "ln_w"))
        out.append((pre + ".bias", (n,), "ln_b"))

    ie = "image_encoder."
    out.append((ie + "pos_embed", (1, g, g, D), "pos"))
    out.append((ie + "patch_embed.proj.weight", (D, 3, p, p), f"w{3 * p * p}"))
    out.append((ie + "patch_embed.proj.bias", (D,), f"b{3 * p * p}"))
    for i in range(cfg.depth):
        b = f"{ie}blocks.{i}."
        side = g if i in cfg.global_attn_indexes else cfg.window_size
        ln(b + "norm1", D)
        out.append((b + "attn.rel_pos_h", (2 * side - 1, hd), "pos"))
        out.append((b + "attn.rel_pos_w", (2 * side - 1, hd), "pos"))
        lin(b + "attn.qkv", 3 * D, D)
        lin(b + "attn.proj", D, D)
        ln(b + "norm2", D)
        lin(b + "mlp.lin1", cfg.mlp_ratio * D, D)
        lin(b + "mlp.lin2", D, cfg.mlp_ratio * D)
    out.append((ie + "neck.0.weight", (E, D, 1, 1), f"w{D}"))
    ln(ie + "neck.1", E)
    out.append((ie + "neck.2.weight", (E, E, 3, 3), f"w{9 * E}"))
    ln(ie + "neck.3", E)

    pe = "prompt_encoder."
    out.append((pe + "pe_layer.positional_encoding_gaussian_matrix", (2, E // 2), "emb"))
    for i in range(4):
        out.append((pe + f"point_embeddings.{i}.weight", (1, E), "emb"))
    out.append((pe + "not_a_point_embed.weight", (1, E), "emb"))
    out.append((pe + "mask_downscaling.0.weight", (4, 1, 2, 2), "w4"))
    out.append((pe + "mask_downscaling.0.bias", (4,), "b4"))
    ln(pe + "mask_downscaling.1", 4)
    out.append((pe + "mask_downscaling.3.weight", (16, 4, 2, 2), "w16"))
    out.append((pe + "mask_downscaling.3.bias", (16,), "b16"))
    ln(pe + "mask_downscaling.4", 16)
    out.append((pe + "mask_downscaling.6.weight", (E, 16, 1, 1), "w16"))
    out.append((pe + "mask_downscaling.6.bias", (E,), "b16"))
    out.append((pe + "no_mask_embed.weight", (1, E), "emb"))

    md = "mask_decoder."
    T = md + "transformer."

    def attn(pre: str, internal: int) -> None:
        for n in ("q_proj", "k_proj", "v_proj"):
            lin(pre + n, internal, E)
        lin(pre + "out_proj", E, internal)

    for l in range(cfg.dec_depth):
        L = f"{T}layers.{l}."
        attn(L + "self_attn.", E)
        ln(L + "norm1", E)
        attn(L + "cross_attn_token_to_image.", E // 2)
        ln(L + "norm2", E)
        lin(L + "mlp.lin1", cfg.dec_mlp_dim, E)
        lin(L + "mlp.lin2", E, cfg.dec_mlp_dim)
        ln(L + "norm3", E)
        ln(L + "norm4", E)
        attn(L + "cross_attn_image_to_token.", E // 2)
    attn(T + "final_attn_token_to_image.", E // 2)
    ln(T + "norm_final_attn", E)
    out.append((md + "iou_token.weight", (1, E), "emb"))
    out.append((md + "mask_tokens.weight", (cfg.num_mask_tokens, E), "emb"))
    out.append((md + "output_upscaling.0.weight", (E, E // 4, 2, 2), f"w{E}"))
    out.append((md + "output_upscaling.0.bias", (E // 4,), f"b{E}"))
    ln(md + "output_upscaling.1", E // 4)
    out.append((md + "output_upscaling.3.weight", (E // 4, E // 8, 2, 2), f"w{E // 4}"))
    out.append((md + "output_upscaling.3.bias", (E // 8,), f"b{E // 4}"))
    for i in range(cfg.num_mask_tokens):
        h = f"{md}output_hypernetworks_mlps.{i}.layers."
        lin(h + "0", E, E)
        lin(h + "1", E, E)
        lin(h + "2", E // 8, E)
    h = md + "iou_prediction_head.layers."
    lin(h + "0", E, E)
    lin(h + "1", E, E)
    lin(h + "2", cfg.num_mask_tokens, E)
    return out


def synthetic_state_dict(cfg: SamConfig, seed: int = 0) -> Dict[str, torch.Tensor]:
    """Seeded fp32 CPU state_dict, loadable by the reference `Sam.load_state_dict` (strict)."""
    sd: Dict[str, torch.Tensor] = {}
    for key, shape, kind in state_dict_shapes(cfg):
        g = torch.Generator(device="cpu").manual_seed((zlib.crc32(key.encode()) ^ (seed * 0x9E3779B1)) & 0x7FFFFFFF)
        if kind[0] in "wb" and kind[1:].isdigit():
            bound = 1.0 / math.sqrt(int(kind[1:]))
            t = (torch.rand(shape, generator=g) * 2 - 1) * bound
        elif kind == "emb":
            t = torch.randn(shape, generator=g)
        elif kind == "pos":
            t = torch.randn(shape, generator=g) * 0.02
        elif kind == "ln_w":
            t = 1.0 + 0.1 * torch.randn(shape, generator=g)
        elif kind == "ln_b":
            t = 0.1 * torch.randn(shape, generator=g)
        else:
            raise ValueError(kind)
        sd[key] = t.contiguous()
    return sd
