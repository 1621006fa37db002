"""Model hyper-parameters of the SAM variants (mirrors /root/reference .../segment_anything/build_sam.py:14-45,55-101)."""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, Tuple


@dataclass(frozen=True)
class SamConfig:
    embed_dim: int = 1280
    depth: int = 32
    num_heads: int = 16
    global_attn_indexes: Tuple[int, ...] = (7, 15, 23, 31)
    window_size: int = 14
    img_size: int = 1024
    patch_size: int = 16
    out_chans: int = 256
    mlp_ratio: int = 4
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
    # test-only shapes (not reference variants): same kernels, seconds on the CPU oracle
    "vit_tiny": SamConfig(embed_dim=128, depth=2, num_heads=2, global_attn_indexes=(1,)),
    "vit_tiny80": SamConfig(embed_dim=160, depth=3, num_heads=2, global_attn_indexes=(2,)),
}


def config_from_state_dict(sd) -> SamConfig:
    """Recover the encoder shape from a reference state_dict (keys as in build_sam.py:103-106)."""
    d = sd["image_encoder.pos_embed"].shape[-1]
    depth = 1 + max(int(k.split(".")[2]) for k in sd if k.startswith("image_encoder.blocks."))
    hd = sd["image_encoder.blocks.0.attn.rel_pos_h"].shape[1]
    side = sd["image_encoder.pos_embed"].shape[1]
    glob = tuple(i for i in range(depth)
                 if sd[f"image_encoder.blocks.{i}.attn.rel_pos_h"].shape[0] == 2 * side - 1)
    return SamConfig(embed_dim=d, depth=depth, num_heads=d // hd, global_attn_indexes=glob)
