"""The caller side of the hot path: InkLayer's `run_SAM` adapter and a batched sketch segmenter.

`run_SAM` keeps the reference's signature and return types (InkLayer/segmentor/sam.py:16-43, including its
BGR<->RGB swap and its empty-box return) but builds the model once instead of once per image, and brings all masks
back with one pinned-memory copy instead of one `.cpu()` sync per mask.  `SketchSegmenter.segment_batch` is the
throughput form used by `main.py --dir`-style drivers: B sketches are encoded in one batch, every box of every sketch
is decoded, masks return as numpy bool arrays."""
from __future__ import annotations

from typing import List, Optional, Sequence

import numpy as np
import torch

from .predictor import SamPredictor
from .resize import ResizeLongestSide
from .sam import Sam, build_sam


class SketchSegmenter:
    def __init__(self, sam: Sam, max_batch: int = 8) -> None:
        if sam.device.type != "cuda":
            raise RuntimeError("SketchSegmenter needs a Sam on a CUDA device")
        self.sam = sam
        self.transform = ResizeLongestSide(sam.image_encoder.img_size)
        self.max_batch = max_batch
        S = sam.image_encoder.img_size
        self._host_img = torch.empty((max_batch, S, S, 3), dtype=torch.uint8).pin_memory()
        self._dev_img = torch.empty((max_batch, S, S, 3), dtype=torch.uint8, device=sam.device)
        self._host_masks: Optional[torch.Tensor] = None
        self.h2d_bytes = 0
        self.d2h_bytes = 0

    def _mask_staging(self, nbytes: int) -> torch.Tensor:
        if self._host_masks is None or self._host_masks.numel() < nbytes:
            self._host_masks = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
        return self._host_masks

    @torch.no_grad()
    def segment_batch(self, images: Sequence[np.ndarray], boxes: Sequence[np.ndarray]) -> List[np.ndarray]:
        """images: uint8 [H,W,3] each (RGB as the model should see it); boxes: float [K_i,4] pixel xyxy in the ORIGINAL
        frame of image i.  Returns one bool array [K_i,H,W] per image."""
        assert len(images) == len(boxes)
        out: List[np.ndarray] = []
        for b0 in range(0, len(images), self.max_batch):
            out.extend(self._segment_chunk(images[b0:b0 + self.max_batch], boxes[b0:b0 + self.max_batch]))
        return out

    def _segment_chunk(self, images, boxes) -> List[np.ndarray]:
        dev, B = self.sam.device, len(images)
        sizes, in_sizes = [], []
        for i, im in enumerate(images):
            r = self.transform.apply_image(im)  # host resize, identity at 1024x1024
            h, w = r.shape[:2]
            self._host_img[i, :h, :w].copy_(torch.from_numpy(r))
            sizes.append(tuple(im.shape[:2]))
            in_sizes.append((h, w))
        self._dev_img[:B].copy_(self._host_img[:B], non_blocking=True)
        self.h2d_bytes += B * self._host_img[0].numel()
        feats = self.sam.encode(self._dev_img[:B], in_sizes)
        total = sum(int(len(bx)) * s[0] * s[1] for bx, s in zip(boxes, sizes))
        stage = self._mask_staging(max(total, 1))
        views, off = [], 0
        for i in range(B):
            bx = torch.as_tensor(np.asarray(boxes[i], dtype=np.float32).reshape(-1, 4))
            tb = self.transform.apply_boxes_torch(bx, sizes[i]).to(dev, non_blocking=True)
            self.h2d_bytes += tb.numel() * 4
            masks, _, _ = self.sam.decode_boxes(feats[i], tb, in_sizes[i], sizes[i])
            n = masks.numel()
            stage[off:off + n].copy_(masks.view(torch.uint8).reshape(-1), non_blocking=True)
            views.append((off, n, (masks.shape[0],) + sizes[i]))
            off += n
        self.d2h_bytes += off
        torch.cuda.current_stream(dev).synchronize()
        host = stage.numpy()
        return [host[o:o + n].view(np.bool_).reshape(shape).copy() for o, n, shape in views]


_default: Optional[SketchSegmenter] = None


def run_SAM(image_pil, boxes_filt, sam_checkpoint: Optional[str] = None, segmenter: Optional[SketchSegmenter] = None):
    """Drop-in for InkLayer/segmentor/sam.py:16-43.  image_pil: PIL RGB image; boxes_filt: float tensor/array [K,4]
    pixel xyxy.  Returns a list of K bool arrays [H,W] (or the reference's 2-tuple of empty arrays for K == 0)."""
    global _default
    if len(boxes_filt) == 0:
        return np.array([]), np.array([])
    if segmenter is None:
        if _default is None:
            _default = SketchSegmenter(build_sam(checkpoint=sam_checkpoint).to("cuda"), max_batch=1)
        segmenter = _default
    image = np.array(image_pil)[..., ::-1]  # the reference applies COLOR_BGR2RGB to an RGB array (sam.py:24-25)
    bx = boxes_filt.detach().cpu().numpy() if isinstance(boxes_filt, torch.Tensor) else np.asarray(boxes_filt)
    masks = segmenter.segment_batch([np.ascontiguousarray(image)], [bx])[0]
    return [m for m in masks]
