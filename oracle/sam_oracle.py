"""CPU oracle for the SAM segmentation hot path -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

A functional fp32 restatement (plain torch CPU ops, no nn.Module) of the arithmetic that
`segment_anything.SamPredictor.set_image / predict_torch` performs in the reference.  Only `tests/`,
`__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference` leg may import this module;
the product path (`inklayer_b200`) never does and fails loudly when its CUDA library is missing.

Parity pinning: the reference has no tests or golden vectors for this path.  The oracle is
pinned against outputs of the reference implementation itself, run in the build container from
/root/reference by `tests/golden/make_golden.py`; the committed vectors live in `tests/golden/*.npz` and
`tests/test_oracle.py` checks this file against them.

All `file:line` citations are relative to
/root/reference/InkLayer/third_party/segment-anything/segment_anything/ ("SA/").
Weights are a flat state_dict with the reference's key names (SA/build_sam.py:103-106).
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.nn.functional as F

Tensor = torch.Tensor


@dataclass(frozen=True)
class SamConfig:
    """Hyper-parameters of one SAM variant (SA/build_sam.py:14-45,55-101)."""

    embed_dim: int = 1280
    depth: int = 32
    num_heads: int = 16
    global_attn_indexes: Tuple[int, ...] = (7, 15, 23, 31)
    window_size: int = 14
    img_size: int = 1024
    patch_size: int = 16
    out_chans: int = 256
    mlp_ratio: int = 4
    # decoder (SA/build_sam.py:84-96)
    dec_depth: int = 2
    dec_heads: int = 8
    dec_mlp_dim: int = 2048
    num_mask_tokens: int = 4
    pixel_mean: Tuple[float, float, float] = (123.675, 116.28, 103.53)
    pixel_std: Tuple[float, float, float] = (58.395, 57.12, 57.375)

    @property
    def grid(self) -> int:
        return self.img_size // self.patch_size

    @property
    def head_dim(self) -> int:
        return self.embed_dim // self.num_heads


CONFIGS: Dict[str, SamConfig] = {
    "vit_h": SamConfig(),
    "vit_l": SamConfig(embed_dim=1024, depth=24, num_heads=16, global_attn_indexes=(5, 11, 17, 23)),
    "vit_b": SamConfig(embed_dim=768, depth=12, num_heads=12, global_attn_indexes=(2, 5, 8, 11)),
    # not a reference variant: a 2-block encoder (one windowed, one global) for second-scale CPU tests
    "vit_tiny": SamConfig(embed_dim=128, depth=2, num_heads=2, global_attn_indexes=(1,)),
    # head_dim 80 like ViT-H, 3 blocks
    "vit_tiny80": SamConfig(embed_dim=160, depth=3, num_heads=2, global_attn_indexes=(2,)),
}


# --------------------------------------------------------------------------------------------- helpers
def _ln(x: Tensor, w: Tensor, b: Tensor, eps: float) -> Tensor:
    return F.layer_norm(x, (x.shape[-1],), w, b, eps)


def _ln2d(x: Tensor, w: Tensor, b: Tensor, eps: float = 1e-6) -> Tensor:
    """LayerNorm over the channel axis of an NCHW map, biased variance (SA/modeling/common.py:31-43)."""
    mu = x.mean(1, keepdim=True)
    var = (x - mu).pow(2).mean(1, keepdim=True)
    y = (x - mu) / torch.sqrt(var + eps)
    return w[None, :, None, None] * y + b[None, :, None, None]


# --------------------------------------------------------------------------------------------- host side
def preprocess_shape(h: int, w: int, long_side: int) -> Tuple[int, int]:
    """SA/utils/transforms.py:93-102: scale so the long side equals `long_side`, round half up."""
    scale = long_side * 1.0 / max(h, w)
    return int(h * scale + 0.5), int(w * scale + 0.5)


def apply_boxes(boxes: Tensor, original_size: Tuple[int, int], long_side: int = 1024) -> Tensor:
    """SA/utils/transforms.py:67-91: xyxy boxes from the original frame to the resized-input frame."""
    oh, ow = original_size
    nh, nw = preprocess_shape(oh, ow, long_side)
    out = boxes.clone().to(torch.float).reshape(-1, 2, 2)
    out[..., 0] = out[..., 0] * (nw / ow)
    out[..., 1] = out[..., 1] * (nh / oh)
    return out.reshape(-1, 4)


def resize_image(image: np.ndarray, long_side: int = 1024) -> np.ndarray:
    """SA/utils/transforms.py:26-31: Pillow bilinear resize of the uint8 HWC image (identity at 1024^2)."""
    from PIL import Image

    h, w = preprocess_shape(image.shape[0], image.shape[1], long_side)
    if (h, w) == image.shape[:2]:
        return image
    return np.array(Image.fromarray(image).resize((w, h), Image.BILINEAR))


# --------------------------------------------------------------------------------------------- encoder
def preprocess(cfg: SamConfig, x: Tensor) -> Tensor:
    """SA/modeling/sam.py:164-174: per-channel normalise, THEN zero-pad right/bottom to img_size.
    x: [B,3,h,w] float (0..255)."""
    mean = torch.tensor(cfg.pixel_mean, dtype=torch.float32).view(1, 3, 1, 1)
    std = torch.tensor(cfg.pixel_std, dtype=torch.float32).view(1, 3, 1, 1)
    x = (x.float() - mean) / std
    h, w = x.shape[-2:]
    return F.pad(x, (0, cfg.img_size - w, 0, cfg.img_size - h))


def patch_embed(cfg: SamConfig, sd: Dict[str, Tensor], x: Tensor) -> Tensor:
    """SA/modeling/image_encoder.py:391-395 + :107-109: non-overlapping conv as a matmul over unfolded
    patches, NHWC out, plus pos_embed.  x: [B,3,1024,1024] -> [B,G,G,D]."""
    p, g = cfg.patch_size, cfg.grid
    B = x.shape[0]
    w = sd["image_encoder.patch_embed.proj.weight"]  # [D,3,p,p]
    b = sd["image_encoder.patch_embed.proj.bias"]
    cols = x.reshape(B, 3, g, p, g, p).permute(0, 2, 4, 1, 3, 5).reshape(B, g, g, 3 * p * p)
    y = cols @ w.reshape(w.shape[0], -1).t() + b
    return y + sd["image_encoder.pos_embed"]


def rel_pos_bias_terms(q: Tensor, rel_pos_h: Tensor, rel_pos_w: Tensor, side: int) -> Tuple[Tensor, Tensor]:
    """SA/modeling/image_encoder.py:292-361 for q_size == k_size == (side, side) and a table that already
    has 2*side-1 rows (true for every SAM variant at 1024^2, so no interpolation branch).
    q: [N, side*side, hd] UNSCALED queries.  Returns rel_h [N,side,side,side(kh)], rel_w [N,side,side,side(kw)]."""
    assert rel_pos_h.shape[0] == 2 * side - 1 and rel_pos_w.shape[0] == 2 * side - 1
    idx = torch.arange(side)[:, None] - torch.arange(side)[None, :] + (side - 1)  # [q, k]
    Rh = rel_pos_h[idx]  # [side(q), side(k), hd]
    Rw = rel_pos_w[idx]
    r_q = q.reshape(q.shape[0], side, side, q.shape[-1])
    rel_h = torch.einsum("nhwc,hkc->nhwk", r_q, Rh)
    rel_w = torch.einsum("nhwc,wkc->nhwk", r_q, Rw)
    return rel_h, rel_w


def encoder_attention(cfg: SamConfig, sd: Dict[str, Tensor], pre: str, x: Tensor) -> Tensor:
    """SA/modeling/image_encoder.py:224-240.  x: [N, S, S, D] (a window or the whole grid)."""
    N, S, _, D = x.shape
    nh, hd = cfg.num_heads, cfg.head_dim
    qkv = F.linear(x, sd[pre + "qkv.weight"], sd[pre + "qkv.bias"])  # [N,S,S,3D]
    qkv = qkv.reshape(N, S * S, 3, nh, hd).permute(2, 0, 3, 1, 4).reshape(3, N * nh, S * S, hd)
    q, k, v = qkv[0], qkv[1], qkv[2]
    attn = (q * hd ** -0.5) @ k.transpose(-2, -1)  # [N*nh, S*S, S*S]
    rel_h, rel_w = rel_pos_bias_terms(q, sd[pre + "rel_pos_h"], sd[pre + "rel_pos_w"], S)
    attn = (attn.view(-1, S, S, S, S) + rel_h[:, :, :, :, None] + rel_w[:, :, :, None, :]).view(-1, S * S, S * S)
    attn = attn.softmax(dim=-1)
    out = (attn @ v).view(N, nh, S, S, hd).permute(0, 2, 3, 1, 4).reshape(N, S, S, D)
    return F.linear(out, sd[pre + "proj.weight"], sd[pre + "proj.bias"])


def window_partition(x: Tensor, ws: int) -> Tuple[Tensor, int]:
    """SA/modeling/image_encoder.py:243-264: zero-pad the grid up to a multiple of ws, cut into windows."""
    B, H, W, C = x.shape
    pad = (ws - H % ws) % ws
    x = F.pad(x, (0, 0, 0, pad, 0, pad))
    Hp = H + pad
    n = Hp // ws
    x = x.view(B, n, ws, n, ws, C).permute(0, 1, 3, 2, 4, 5).reshape(B * n * n, ws, ws, C)
    return x, Hp


def window_unpartition(win: Tensor, ws: int, Hp: int, H: int) -> Tensor:
    """SA/modeling/image_encoder.py:267-289: inverse of window_partition, dropping the padded rows/cols."""
    n = Hp // ws
    B = win.shape[0] // (n * n)
    x = win.view(B, n, n, ws, ws, -1).permute(0, 1, 3, 2, 4, 5).reshape(B, Hp, Hp, -1)
    return x[:, :H, :H, :]


def encoder_block(cfg: SamConfig, sd: Dict[str, Tensor], i: int, x: Tensor) -> Tensor:
    """SA/modeling/image_encoder.py:166-182 (+ MLPBlock, SA/modeling/common.py:13-26)."""
    pre = f"image_encoder.blocks.{i}."
    is_global = i in cfg.global_attn_indexes
    h = _ln(x, sd[pre + "norm1.weight"], sd[pre + "norm1.bias"], 1e-6)
    if is_global:
        h = encoder_attention(cfg, sd, pre + "attn.", h)
    else:
        H = h.shape[1]
        win, Hp = window_partition(h, cfg.window_size)
        win = encoder_attention(cfg, sd, pre + "attn.", win)
        h = window_unpartition(win, cfg.window_size, Hp, H)
    x = x + h
    h = _ln(x, sd[pre + "norm2.weight"], sd[pre + "norm2.bias"], 1e-6)
    h = F.linear(h, sd[pre + "mlp.lin1.weight"], sd[pre + "mlp.lin1.bias"])
    h = F.gelu(h)  # exact erf form (nn.GELU default)
    h = F.linear(h, sd[pre + "mlp.lin2.weight"], sd[pre + "mlp.lin2.bias"])
    return x + h


def neck(cfg: SamConfig, sd: Dict[str, Tensor], x: Tensor) -> Tensor:
    """SA/modeling/image_encoder.py:88-104,114.  x: [B,G,G,D] -> [B,256,G,G]."""
    x = x.permute(0, 3, 1, 2)
    x = F.conv2d(x, sd["image_encoder.neck.0.weight"])
    x = _ln2d(x, sd["image_encoder.neck.1.weight"], sd["image_encoder.neck.1.bias"])
    x = F.conv2d(x, sd["image_encoder.neck.2.weight"], padding=1)
    return _ln2d(x, sd["image_encoder.neck.3.weight"], sd["image_encoder.neck.3.bias"])


@torch.no_grad()
def image_encoder(cfg: SamConfig, sd: Dict[str, Tensor], x: Tensor, taps: Optional[dict] = None) -> Tensor:
    """SA/modeling/image_encoder.py:106-116.  x: preprocessed [B,3,1024,1024] -> [B,256,64,64].
    `taps`, when given, receives intermediate tensors by name for per-kernel tests."""
    t = patch_embed(cfg, sd, x)
    if taps is not None:
        taps["patch_embed"] = t.clone()
    for i in range(cfg.depth):
        t = encoder_block(cfg, sd, i, t)
        if taps is not None:
            taps[f"block{i}"] = t.clone()
    return neck(cfg, sd, t)


# --------------------------------------------------------------------------------------------- prompt encoder
def _pe_encoding(sd: Dict[str, Tensor], coords01: Tensor) -> Tensor:
    """SA/modeling/prompt_encoder.py:185-192: random-Fourier features of points in [0,1]^2."""
    g = sd["prompt_encoder.pe_layer.positional_encoding_gaussian_matrix"]  # [2,128]
    c = (2 * coords01 - 1) @ g
    c = 2 * np.pi * c
    return torch.cat([torch.sin(c), torch.cos(c)], dim=-1)


def dense_pe(cfg: SamConfig, sd: Dict[str, Tensor]) -> Tensor:
    """SA/modeling/prompt_encoder.py:62-71,194-205: PE of the pixel centres of the G x G grid -> [1,256,G,G]."""
    g = cfg.grid
    c = (torch.arange(g, dtype=torch.float32) + 0.5) / g
    yy, xx = torch.meshgrid(c, c, indexing="ij")
    pe = _pe_encoding(sd, torch.stack([xx, yy], dim=-1))  # [G,G,256]
    return pe.permute(2, 0, 1).unsqueeze(0)


def embed_boxes(cfg: SamConfig, sd: Dict[str, Tensor], boxes: Tensor) -> Tensor:
    """SA/modeling/prompt_encoder.py:93-100,207-214: boxes [K,4] (input frame, xyxy) -> sparse [K,2,256]."""
    corners = (boxes.float() + 0.5).reshape(-1, 2, 2) / cfg.img_size
    emb = _pe_encoding(sd, corners)
    emb[:, 0, :] += sd["prompt_encoder.point_embeddings.2.weight"][0]
    emb[:, 1, :] += sd["prompt_encoder.point_embeddings.3.weight"][0]
    return emb


# --------------------------------------------------------------------------------------------- mask decoder
def _dec_attention(sd: Dict[str, Tensor], pre: str, heads: int, q: Tensor, k: Tensor, v: Tensor) -> Tensor:
    """SA/modeling/transformer.py:218-240."""
    q = F.linear(q, sd[pre + "q_proj.weight"], sd[pre + "q_proj.bias"])
    k = F.linear(k, sd[pre + "k_proj.weight"], sd[pre + "k_proj.bias"])
    v = F.linear(v, sd[pre + "v_proj.weight"], sd[pre + "v_proj.bias"])
    b, nq, c = q.shape
    hd = c // heads

    def split(t: Tensor) -> Tensor:
        return t.reshape(b, t.shape[1], heads, hd).transpose(1, 2)

    q, k, v = split(q), split(k), split(v)
    attn = torch.softmax((q @ k.transpose(-2, -1)) / math.sqrt(hd), dim=-1)
    out = (attn @ v).transpose(1, 2).reshape(b, nq, c)
    return F.linear(out, sd[pre + "out_proj.weight"], sd[pre + "out_proj.bias"])


def two_way_transformer(cfg: SamConfig, sd: Dict[str, Tensor], src: Tensor, pos: Tensor, tokens: Tensor,
                        taps: Optional[dict] = None) -> Tuple[Tensor, Tensor]:
    """SA/modeling/transformer.py:62-106,151-182.  src,pos: [K,256,G,G]; tokens: [K,7,256]."""
    T = "mask_decoder.transformer."
    keys = src.flatten(2).permute(0, 2, 1)
    key_pe = pos.flatten(2).permute(0, 2, 1)
    queries, query_pe = tokens, tokens
    H = cfg.dec_heads
    for l in range(cfg.dec_depth):
        L = f"{T}layers.{l}."
        if l == 0:  # skip_first_layer_pe: the self-attention output REPLACES the queries
            queries = _dec_attention(sd, L + "self_attn.", H, queries, queries, queries)
        else:
            q = queries + query_pe
            queries = queries + _dec_attention(sd, L + "self_attn.", H, q, q, queries)
        queries = _ln(queries, sd[L + "norm1.weight"], sd[L + "norm1.bias"], 1e-5)
        q, k = queries + query_pe, keys + key_pe
        queries = queries + _dec_attention(sd, L + "cross_attn_token_to_image.", H, q, k, keys)
        queries = _ln(queries, sd[L + "norm2.weight"], sd[L + "norm2.bias"], 1e-5)
        m = F.relu(F.linear(queries, sd[L + "mlp.lin1.weight"], sd[L + "mlp.lin1.bias"]))
        queries = queries + F.linear(m, sd[L + "mlp.lin2.weight"], sd[L + "mlp.lin2.bias"])
        queries = _ln(queries, sd[L + "norm3.weight"], sd[L + "norm3.bias"], 1e-5)
        q, k = queries + query_pe, keys + key_pe
        keys = keys + _dec_attention(sd, L + "cross_attn_image_to_token.", H, k, q, queries)
        keys = _ln(keys, sd[L + "norm4.weight"], sd[L + "norm4.bias"], 1e-5)
        if taps is not None:
            taps[f"dec_queries{l}"] = queries.clone()
            taps[f"dec_keys{l}"] = keys.clone()
    q, k = queries + query_pe, keys + key_pe
    queries = queries + _dec_attention(sd, T + "final_attn_token_to_image.", H, q, k, keys)
    queries = _ln(queries, sd[T + "norm_final_attn.weight"], sd[T + "norm_final_attn.bias"], 1e-5)
    return queries, keys


def _mlp_relu(sd: Dict[str, Tensor], pre: str, n: int, x: Tensor) -> Tensor:
    """SA/modeling/mask_decoder.py:154-176."""
    for i in range(n):
        x = F.linear(x, sd[f"{pre}layers.{i}.weight"], sd[f"{pre}layers.{i}.bias"])
        if i < n - 1:
            x = F.relu(x)
    return x


@torch.no_grad()
def mask_decoder(cfg: SamConfig, sd: Dict[str, Tensor], embedding: Tensor, boxes: Tensor,
                 taps: Optional[dict] = None) -> Tuple[Tensor, Tensor]:
    """SA/predictor.py:222-235 + SA/modeling/mask_decoder.py:71-149 with multimask_output=False.
    embedding: [1,256,G,G]; boxes: [K,4] input frame.  Returns low_res [K,1,4G,4G], iou [K,1]."""
    M = "mask_decoder."
    K = boxes.shape[0]
    g = cfg.grid
    sparse = embed_boxes(cfg, sd, boxes)
    out_tokens = torch.cat([sd[M + "iou_token.weight"], sd[M + "mask_tokens.weight"]], dim=0)
    tokens = torch.cat([out_tokens.unsqueeze(0).expand(K, -1, -1), sparse], dim=1)  # [K,7,256]
    src = embedding.expand(K, -1, -1, -1) + sd["prompt_encoder.no_mask_embed.weight"].reshape(1, -1, 1, 1)
    pos = dense_pe(cfg, sd).expand(K, -1, -1, -1)
    hs, keys = two_way_transformer(cfg, sd, src, pos, tokens, taps)
    iou_tok, mask_toks = hs[:, 0, :], hs[:, 1:1 + cfg.num_mask_tokens, :]
    up = keys.transpose(1, 2).reshape(K, -1, g, g)
    up = F.conv_transpose2d(up, sd[M + "output_upscaling.0.weight"], sd[M + "output_upscaling.0.bias"], stride=2)
    up = F.gelu(_ln2d(up, sd[M + "output_upscaling.1.weight"], sd[M + "output_upscaling.1.bias"]))
    up = F.gelu(F.conv_transpose2d(up, sd[M + "output_upscaling.3.weight"], sd[M + "output_upscaling.3.bias"], stride=2))
    hyper = torch.stack([_mlp_relu(sd, f"{M}output_hypernetworks_mlps.{i}.", 3, mask_toks[:, i, :])
                         for i in range(cfg.num_mask_tokens)], dim=1)  # [K,4,32]
    masks = (hyper @ up.reshape(K, up.shape[1], -1)).reshape(K, -1, 4 * g, 4 * g)
    iou = _mlp_relu(sd, M + "iou_prediction_head.", 3, iou_tok)
    if taps is not None:
        taps["hs"] = hs.clone()
        taps["hyper"] = hyper.clone()
    return masks[:, 0:1], iou[:, 0:1]


def postprocess_masks(cfg: SamConfig, low_res: Tensor, input_size: Tuple[int, int],
                      original_size: Tuple[int, int]) -> Tensor:
    """SA/modeling/sam.py:133-162: bilinear up to img_size, crop to the un-padded input, bilinear to original."""
    m = F.interpolate(low_res, (cfg.img_size, cfg.img_size), mode="bilinear", align_corners=False)
    m = m[..., : input_size[0], : input_size[1]]
    return F.interpolate(m, original_size, mode="bilinear", align_corners=False)


# --------------------------------------------------------------------------------------------- predictor
class OraclePredictor:
    """The SamPredictor state machine (SA/predictor.py:17-269) over the functional oracle, box prompts only."""

    def __init__(self, cfg: SamConfig, sd: Dict[str, Tensor]):
        self.cfg, self.sd = cfg, {k: v.float() for k, v in sd.items()}
        self.is_image_set = False

    @torch.no_grad()
    def set_image(self, image: np.ndarray, image_format: str = "RGB") -> None:
        assert image_format in ["RGB", "BGR"]
        if image_format != "RGB":  # model.image_format is "RGB" (SA/modeling/sam.py:20)
            image = image[..., ::-1]
        inp = resize_image(np.ascontiguousarray(image), self.cfg.img_size)
        t = torch.as_tensor(inp).permute(2, 0, 1).contiguous()[None]
        self.set_torch_image(t, image.shape[:2])

    @torch.no_grad()
    def set_torch_image(self, t: Tensor, original_size: Tuple[int, int]) -> None:
        assert t.dim() == 4 and t.shape[1] == 3 and max(t.shape[2:]) == self.cfg.img_size
        self.original_size = tuple(original_size)
        self.input_size = tuple(t.shape[-2:])
        self.features = image_encoder(self.cfg, self.sd, preprocess(self.cfg, t))
        self.is_image_set = True

    @torch.no_grad()
    def predict_torch(self, boxes: Tensor, return_logits: bool = False):
        if not self.is_image_set:
            raise RuntimeError("An image must be set with .set_image(...) before mask prediction.")
        low_res, iou = mask_decoder(self.cfg, self.sd, self.features, boxes)
        masks = postprocess_masks(self.cfg, low_res, self.input_size, self.original_size)
        if not return_logits:
            masks = masks > 0.0
        return masks, iou, low_res
