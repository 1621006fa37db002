"""`SamPredictor`, the stateful surface InkLayer's segmentor calls (InkLayer/segmentor/sam.py:23-36), backed by the
B200 kernels.  Method names, argument meaning, cached state and error behaviour follow
segment_anything/predictor.py:17-269; prompts other than boxes are not part of InkLayer's path and raise
NotImplementedError."""
from __future__ import annotations

from typing import Optional, Tuple

import numpy as np
import torch

from .resize import ResizeLongestSide
from .sam import Sam


class SamPredictor:
    def __init__(self, sam_model: Sam) -> None:
        self.model = sam_model
        self.transform = ResizeLongestSide(sam_model.image_encoder.img_size)
        self._stage_u8 = None    # device staging buffers in the input frame: a stable pointer for the encoder's CUDA graph
        self._stage_f32 = None
        self.reset_image()

    @property
    def device(self) -> torch.device:
        return self.model.device

    def reset_image(self) -> None:
        self.is_image_set = False
        self.features = None
        self.original_size = None
        self.input_size = None

    # ------------------------------------------------------------------ encode
    def set_image(self, image: np.ndarray, image_format: str = "RGB") -> None:
        """predictor.py:34-60: uint8 HWC image -> device -> ResizeLongestSide -> encoder; caches the embedding."""
        assert image_format in ["RGB", "BGR"], f"image_format must be in ['RGB', 'BGR'], is {image_format}."
        if image_format != self.model.image_format:
            image = image[..., ::-1]
        # H2D of the original pixels, then the resize on the device (bit-exact with the reference's host-side Pillow
        # resize, oracle/resize_oracle.py); the encoder kernel reads uint8 HWC directly: no permute / float copy
        src = torch.as_tensor(np.ascontiguousarray(image), device=self.device)
        resized = self.model.resize_longest_side(src)
        self._encode(resized[None], tuple(image.shape[:2]), tuple(resized.shape[:2]))

    @torch.no_grad()
    def set_torch_image(self, transformed_image: torch.Tensor, original_image_size: Tuple[int, ...]) -> None:
        """predictor.py:62-90: 1x3xHxW tensor already resized with ResizeLongestSide (uint8 or float, 0..255)."""
        assert (
            len(transformed_image.shape) == 4
            and transformed_image.shape[1] == 3
            and max(*transformed_image.shape[2:]) == self.model.image_encoder.img_size
        ), f"set_torch_image input must be BCHW with long side {self.model.image_encoder.img_size}."
        t = transformed_image.to(device=self.device, dtype=torch.float32)
        self._encode(t, tuple(original_image_size), tuple(transformed_image.shape[-2:]))

    def _encode(self, t: torch.Tensor, original_size, input_size) -> None:
        """One image through the encoder.  The image is copied into a per-predictor staging buffer of the model's input
        frame so that the ~200 kernel launches of the batch-1 encode replay from a CUDA graph (captured per input size the
        second time it is seen): `set_image` on a stream of sketches is launch-bound otherwise.  The embedding is copied out
        of the graph's buffer: `features` stays valid until reset_image / the next set_image, as in the reference."""
        self.reset_image()
        S = self.model.image_encoder.img_size
        h, w = int(input_size[0]), int(input_size[1])
        if t.dtype == torch.uint8:
            if self._stage_u8 is None:
                self._stage_u8 = torch.zeros((1, S, S, 3), dtype=torch.uint8, device=self.device)
            self._stage_u8[0, :h, :w].copy_(t[0])
            feats = self.model.encode(self._stage_u8, [(h, w)], graph=True)
        else:
            if self._stage_f32 is None:
                self._stage_f32 = torch.zeros((1, 3, S, S), dtype=torch.float32, device=self.device)
            self._stage_f32[0, :, :h, :w].copy_(t[0])
            feats = self.model.encode(self._stage_f32, [(h, w)], graph=True)
        self.original_size = original_size
        self.input_size = input_size
        self.features = feats.clone()
        self.is_image_set = True

    def get_image_embedding(self) -> torch.Tensor:
        if not self.is_image_set:
            raise RuntimeError("An image must be set with .set_image(...) to generate an embedding.")
        assert self.features is not None, "Features must exist if an image has been set."
        return self.features

    # ------------------------------------------------------------------ decode
    @torch.no_grad()
    def predict_torch(self, point_coords: Optional[torch.Tensor], point_labels: Optional[torch.Tensor],
                      boxes: Optional[torch.Tensor] = None, mask_input: Optional[torch.Tensor] = None,
                      multimask_output: bool = True, return_logits: bool = False):
        """predictor.py:168-243 for the call InkLayer makes: boxes [K,4] in the input frame, multimask_output=False.
        Returns masks [K,1,H,W] (bool, or float logits), iou [K,1], low-res logits [K,1,256,256] on the device."""
        if not self.is_image_set:
            raise RuntimeError("An image must be set with .set_image(...) before mask prediction.")
        if point_coords is not None or point_labels is not None or mask_input is not None:
            raise NotImplementedError("inklayer_b200 implements the box-prompt path of SamPredictor only")
        if boxes is None:
            raise NotImplementedError("predict_torch needs boxes (the only prompt InkLayer uses)")
        if multimask_output:
            raise NotImplementedError("multimask_output=True is not built; InkLayer calls with multimask_output=False")
        boxes = boxes.reshape(-1, 4)
        return self.model.decode_boxes(self.features, boxes, self.input_size, self.original_size, return_logits)

    def predict(self, point_coords: Optional[np.ndarray] = None, point_labels: Optional[np.ndarray] = None,
                box: Optional[np.ndarray] = None, mask_input: Optional[np.ndarray] = None,
                multimask_output: bool = True, return_logits: bool = False):
        """predictor.py:92-166, numpy in / numpy out, one box in the ORIGINAL image frame."""
        if not self.is_image_set:
            raise RuntimeError("An image must be set with .set_image(...) before mask prediction.")
        if point_coords is not None or mask_input is not None:
            raise NotImplementedError("inklayer_b200 implements the box-prompt path of SamPredictor only")
        box_t = None
        if box is not None:
            b = self.transform.apply_boxes(np.asarray(box), self.original_size)
            box_t = torch.as_tensor(b, dtype=torch.float, device=self.device)
        masks, iou, low = self.predict_torch(None, None, box_t, None, multimask_output, return_logits=return_logits)
        return masks[0].cpu().numpy(), iou[0].cpu().numpy(), low[0].cpu().numpy()
