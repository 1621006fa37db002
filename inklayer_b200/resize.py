"""Host-side geometry of the SamPredictor surface: the `ResizeLongestSide` helper that `run_SAM` reaches through
`predictor.transform` (InkLayer/segmentor/sam.py:27-29).  Behaviour follows segment_anything/utils/transforms.py:
the long side is scaled to `target_length`, sizes round half up (:93-102), coordinates scale by new/old per axis
(:33-46, :67-91), and the uint8 image is resized on the host with Pillow's bilinear filter (:26-31), which is the
identity for 1024x1024 sketches."""
from __future__ import annotations

from typing import Sequence, Tuple

import numpy as np
import torch


def _scaled_size(h: int, w: int, target: int) -> Tuple[int, int]:
    s = float(target) / max(h, w)
    return int(h * s + 0.5), int(w * s + 0.5)


class ResizeLongestSide:
    def __init__(self, target_length: int) -> None:
        self.target_length = int(target_length)

    @staticmethod
    def get_preprocess_shape(oldh: int, oldw: int, long_side_length: int) -> Tuple[int, int]:
        return _scaled_size(oldh, oldw, long_side_length)

    def _factors(self, original_size: Sequence[int]) -> Tuple[float, float]:
        h, w = int(original_size[0]), int(original_size[1])
        nh, nw = _scaled_size(h, w, self.target_length)
        return nw / w, nh / h  # (x factor, y factor)

    def apply_image(self, image: np.ndarray) -> np.ndarray:
        nh, nw = _scaled_size(image.shape[0], image.shape[1], self.target_length)
        image = np.ascontiguousarray(image)
        if (nh, nw) == image.shape[:2]:
            return image
        from PIL import Image

        return np.asarray(Image.fromarray(image).resize((nw, nh), Image.BILINEAR))

    def apply_coords(self, coords: np.ndarray, original_size: Sequence[int]) -> np.ndarray:
        fx, fy = self._factors(original_size)
        out = np.array(coords, dtype=float, copy=True)
        out[..., 0] *= fx
        out[..., 1] *= fy
        return out

    def apply_boxes(self, boxes: np.ndarray, original_size: Sequence[int]) -> np.ndarray:
        return self.apply_coords(np.asarray(boxes).reshape(-1, 2, 2), original_size).reshape(-1, 4)

    def apply_coords_torch(self, coords: torch.Tensor, original_size: Sequence[int]) -> torch.Tensor:
        fx, fy = self._factors(original_size)
        out = coords.detach().clone().to(torch.float)
        out[..., 0] = out[..., 0] * fx
        out[..., 1] = out[..., 1] * fy
        return out

    def apply_boxes_torch(self, boxes: torch.Tensor, original_size: Sequence[int]) -> torch.Tensor:
        return self.apply_coords_torch(boxes.reshape(-1, 2, 2), original_size).reshape(-1, 4)
