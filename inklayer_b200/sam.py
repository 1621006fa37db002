"""`Sam` model object and the `build_sam*` factories of the drop-in surface (segment_anything/build_sam.py:14-107,
segment_anything/modeling/sam.py:18-47).  The object owns a libinksam handle instead of nn.Modules; PyTorch is used
for tensor ownership, streams and host<->device copies only.  Every compute call goes through the C ABI in
include/inksam.h and raises if the CUDA library or a B200 is missing -- there is no CPU path."""
from __future__ import annotations

import ctypes
from types import SimpleNamespace
from typing import Dict, Optional, Sequence, Tuple

import torch

from . import _lib
from .config import CONFIGS, SamConfig, config_from_state_dict
from .weights import state_dict_shapes, synthetic_state_dict


def _stream(device: torch.device) -> ctypes.c_void_p:
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _ptr(t: Optional[torch.Tensor]) -> ctypes.c_void_p:
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


class Sam:
    """Stands where `segment_anything.modeling.Sam` stands for SamPredictor: `.device`, `.image_format`,
    `.mask_threshold`, `.image_encoder.img_size`, `.to()`, `.eval()`, `.load_state_dict()`."""

    mask_threshold: float = 0.0  # modeling/sam.py:19
    image_format: str = "RGB"    # modeling/sam.py:20

    def __init__(self, cfg: SamConfig) -> None:
        self.cfg = cfg
        self.image_encoder = SimpleNamespace(img_size=cfg.img_size)
        self.device = torch.device("cpu")
        self._sd: Optional[Dict[str, torch.Tensor]] = None
        self._handle: Optional[ctypes.c_void_p] = None
        self._graphs: dict = {}        # encode() CUDA graphs: key -> (graph, output, input, kernel launches per replay)
        self._graph_seen: dict = {}
        self._graph_launches = 0       # kernels launched by graph replays (the library counts stream launches only)
        self._graph_gen = 0            # inksam_workspace_generation the cached graphs were captured under
        self._profiling = False

    # ------------------------------------------------------------------ weights
    def load_state_dict(self, state_dict: Dict[str, torch.Tensor], strict: bool = True) -> None:
        """Accepts the reference's flat state_dict (build_sam.py:103-106): same keys, fp32."""
        want = {k: tuple(s) for k, s, _ in state_dict_shapes(self.cfg)}
        missing = [k for k in want if k not in state_dict]
        unexpected = [k for k in state_dict if k not in want]
        if strict and (missing or unexpected):
            raise RuntimeError(f"Error(s) in loading state_dict for Sam: missing {missing[:5]}, unexpected {unexpected[:5]}")
        for k, shape in want.items():
            if k in state_dict and tuple(state_dict[k].shape) != shape:
                raise RuntimeError(f"size mismatch for {k}: checkpoint {tuple(state_dict[k].shape)}, model {shape}")
        self._sd = {k: v.detach() for k, v in state_dict.items() if k in want}
        if self._handle is not None:
            self._upload()

    def state_dict(self) -> Dict[str, torch.Tensor]:
        if self._sd is None:
            raise RuntimeError("no weights loaded")
        return dict(self._sd)

    def eval(self) -> "Sam":
        return self

    def to(self, device) -> "Sam":
        device = torch.device(device)
        if device.type != "cuda":
            raise RuntimeError("inklayer_b200.Sam runs on a CUDA device only (sm_100a kernels, no CPU fallback)")
        if device.index is None:
            device = torch.device("cuda", torch.cuda.current_device())
        if self._handle is not None and device == self.device:
            return self
        self._destroy()
        lib = _lib.load()
        c = _lib.InksamConfig()
        c.embed_dim, c.depth, c.num_heads = self.cfg.embed_dim, self.cfg.depth, self.cfg.num_heads
        c.num_global = len(self.cfg.global_attn_indexes)
        for i, g in enumerate(self.cfg.global_attn_indexes):
            c.global_attn_indexes[i] = g
        c.window_size, c.img_size, c.patch_size, c.out_chans = (self.cfg.window_size, self.cfg.img_size,
                                                                 self.cfg.patch_size, self.cfg.out_chans)
        for i in range(3):
            c.pixel_mean[i], c.pixel_std[i] = self.cfg.pixel_mean[i], self.cfg.pixel_std[i]
        h = ctypes.c_void_p()
        _lib.check(lib.inksam_create(ctypes.byref(c), device.index, ctypes.byref(h)))
        self._handle, self.device = h, device
        if self._sd is not None:
            self._upload()
        return self

    def cuda(self, index: Optional[int] = None) -> "Sam":
        return self.to(torch.device("cuda", torch.cuda.current_device() if index is None else index))

    def _upload(self) -> None:
        self._graphs.clear()  # captured graphs hold pointers into the previous weight set
        self._graph_seen.clear()
        lib = _lib.load()
        with torch.cuda.device(self.device):
            st = _stream(self.device)
            for key, t in self._sd.items():
                d = t.to(device=self.device, dtype=torch.float32).contiguous()
                shape = (ctypes.c_int64 * d.dim())(*d.shape)
                _lib.check(lib.inksam_load_tensor(self._handle, key.encode(), _ptr(d), shape, d.dim(), st))
            _lib.check(lib.inksam_finalize_weights(self._handle, st))
            torch.cuda.current_stream(self.device).synchronize()

    def _destroy(self) -> None:
        self._graphs.clear()
        self._graph_seen.clear()
        if self._handle is not None:
            _lib.load().inksam_destroy(self._handle)
            self._handle = None

    def __del__(self) -> None:
        try:
            self._destroy()
        except Exception:
            pass

    def _require(self) -> ctypes.c_void_p:
        if self._handle is None or self._sd is None:
            raise RuntimeError("Sam is not on a CUDA device or has no weights: call build_sam(...).to('cuda')")
        return self._handle

    @property
    def launch_count(self) -> int:
        return int(_lib.load().inksam_launch_count(self._require())) + self._graph_launches

    def reserve(self, max_batch: int, max_boxes: int = 128) -> int:
        """Sizes the library's workspace once for encodes of up to `max_batch` images and decodes of up to `max_boxes`
        boxes per pass (inksam_reserve): nothing inside that envelope allocates afterwards, so CUDA graphs captured by
        encode(graph=True) stay valid.  Synchronises the device.  Returns the workspace size in bytes."""
        lib = _lib.load()
        with torch.cuda.device(self.device):
            _lib.check(lib.inksam_reserve(self._require(), int(max_batch), int(max_boxes)))
        self._drop_stale_graphs()
        return int(lib.inksam_workspace_bytes(self._require(), int(max_batch), int(max_boxes)))

    def _drop_stale_graphs(self) -> None:
        """Graphs bake in pointers into the library's workspace: when it was reallocated since they were captured
        (inksam_workspace_generation moved) they would replay into retired memory, so they are dropped and re-captured."""
        gen = int(_lib.load().inksam_workspace_generation(self._require()))
        if gen != self._graph_gen:
            self._graphs.clear()
            self._graph_seen.clear()
            self._graph_gen = gen

    def profile(self, on: bool) -> None:
        """Start (and reset) / stop per-category CUDA-event timing of the library's launches."""
        _lib.check(_lib.load().inksam_profile_enable(self._require(), 1 if on else 0))
        self._profiling = bool(on)  # per-launch events are not part of a captured graph: encode() uses stream launches

    def profile_read(self):
        """{category: (ms, work, launches)} since profile(True); waits for the recorded events."""
        ms, work, cnt = (ctypes.c_double * 6)(), (ctypes.c_double * 6)(), (ctypes.c_longlong * 6)()
        _lib.check(_lib.load().inksam_profile_read(self._require(), ms, work, cnt))
        names = {0: "encoder_gemm", 1: "encoder_attention", 3: "decoder_gemm", 5: "postprocess"}
        return {n: (ms[c], work[c], int(cnt[c])) for c, n in names.items()}

    # ------------------------------------------------------------------ compute (C ABI)
    @torch.no_grad()
    def encode(self, images: torch.Tensor, sizes: Optional[Sequence[Tuple[int, int]]] = None, graph: bool = False) -> torch.Tensor:
        """Sam.preprocess + image encoder (modeling/sam.py:164-174, image_encoder.py:106-116) for a batch.
        images: uint8 [B,h,w,3] (HWC) or float32 [B,3,h,w] (values 0..255) on this device, long side <= 1024;
        `sizes` optionally gives each image's valid (h, w) inside the container.  Returns float32 [B,256,64,64].

        graph=True: the ~260 kernel launches of the encoder are replayed from a CUDA graph captured for this input
        buffer (same data pointer, batch and sizes; captured the second time the combination is seen).  The result is
        then a buffer owned by the graph, valid until the next encode of the same input buffer on this stream."""
        self._require()
        if images.device != self.device:
            raise RuntimeError(f"images are on {images.device}, model on {self.device}")
        images = images.contiguous()
        B = images.shape[0]
        if images.dtype == torch.uint8 and images.dim() == 4 and images.shape[3] == 3:
            dtype, hh, ww = 0, images.shape[1], images.shape[2]
            row_pitch, plane = ww * 3, 0
            img_stride = hh * ww * 3
        elif images.dtype == torch.float32 and images.dim() == 4 and images.shape[1] == 3:
            dtype, hh, ww = 1, images.shape[2], images.shape[3]
            row_pitch, plane = ww * 4, hh * ww * 4
            img_stride = 3 * hh * ww * 4
        else:
            raise RuntimeError(f"encode expects uint8 [B,h,w,3] or float32 [B,3,h,w], got {images.dtype} {tuple(images.shape)}")
        if max(hh, ww) > self.cfg.img_size:
            raise RuntimeError(f"image {hh}x{ww} exceeds img_size {self.cfg.img_size}")
        hw = (ctypes.c_int * (2 * B))()
        for i in range(B):
            sh, sw = (hh, ww) if sizes is None else sizes[i]
            hw[2 * i], hw[2 * i + 1] = int(sh), int(sw)

        def run(out: torch.Tensor) -> None:
            with torch.cuda.device(self.device):
                _lib.check(_lib.load().inksam_encode(self._require(), _ptr(images), dtype, B, hw, img_stride, row_pitch, plane,
                                                     _ptr(out), _stream(self.device)))

        if graph and not self._profiling:
            self._drop_stale_graphs()
            key = (images.data_ptr(), dtype, B, hh, ww, tuple(hw))
            hit = self._graphs.get(key)
            if hit is not None:
                hit[0].replay()
                self._graph_launches += hit[3]
                return hit[1]
            self._graph_seen[key] = self._graph_seen.get(key, 0) + 1
            if self._graph_seen[key] >= 2 and len(self._graphs) < 16:
                out = torch.empty((B, self.cfg.out_chans, 64, 64), dtype=torch.float32, device=self.device)
                n0 = int(_lib.load().inksam_launch_count(self._require()))
                run(out)  # workspaces are grown outside the capture
                per_replay = int(_lib.load().inksam_launch_count(self._require())) - n0
                self._drop_stale_graphs()  # that run may have grown the workspace under older graphs
                g = torch.cuda.CUDAGraph()
                # thread_local: SketchSegmenter's hand-off thread may be waiting on events / copying while this captures
                with torch.cuda.graph(g, capture_error_mode="thread_local"):
                    run(out)
                self._graphs[key] = (g, out, images, per_replay)  # the input buffer stays alive with its graph
                self._graph_launches -= per_replay  # the capture pass counted launches that did not execute
                g.replay()
                self._graph_launches += per_replay
                return out
        out = torch.empty((B, self.cfg.out_chans, 64, 64), dtype=torch.float32, device=self.device)
        run(out)
        return out

    @torch.no_grad()
    def resize_longest_side(self, image: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """ResizeLongestSide.apply_image on the device (utils/transforms.py:26-31), bit-exact with the Pillow path the
        reference runs on the host.  image: uint8 [H,W,3] on this device (rows may be strided).  Returns the uint8
        [h,w,3] image; with `out` (a uint8 [>=h,>=w,3] device container, e.g. one slot of a 1024x1024 batch) the result
        is written into out[:h,:w] and that view is returned."""
        from .resize import _scaled_size

        self._require()
        if image.device != self.device or image.dtype != torch.uint8 or image.dim() != 3 or image.shape[2] != 3:
            raise RuntimeError(f"resize_longest_side expects uint8 [H,W,3] on {self.device}, got {image.dtype} {tuple(image.shape)}")
        if image.stride(2) != 1 or image.stride(1) != 3:
            image = image.contiguous()
        H, W = int(image.shape[0]), int(image.shape[1])
        h, w = _scaled_size(H, W, self.cfg.img_size)
        if out is None and (h, w) == (H, W):
            return image
        if out is None:
            out = torch.empty((h, w, 3), dtype=torch.uint8, device=self.device)
        elif (out.device != self.device or out.dtype != torch.uint8 or out.dim() != 3 or out.shape[0] < h or out.shape[1] < w
              or out.shape[2] != 3 or out.stride(2) != 1 or out.stride(1) != 3):
            raise RuntimeError("resize_longest_side: `out` must be a uint8 [>=h,>=w,3] device tensor with packed pixels")
        lib = _lib.load()
        need = int(lib.inksam_resize_workspace_bytes(H, W, h, w))
        ws = torch.empty(max(need, 1), dtype=torch.uint8, device=self.device)  # stream-ordered: safe to drop after the call
        with torch.cuda.device(self.device):
            _lib.check(lib.inksam_resize_u8(_ptr(image), H, W, int(image.stride(0)), _ptr(out), h, w, int(out.stride(0)),
                                            _ptr(ws), _stream(self.device)))
        return out[:h, :w]

    @torch.no_grad()
    def decode_boxes(self, embedding: torch.Tensor, boxes: torch.Tensor, input_size: Tuple[int, int],
                     original_size: Tuple[int, int], return_logits: bool = False):
        """prompt encoder + mask decoder + postprocess_masks + threshold for K boxes of one image
        (predictor.py:222-241).  embedding: float32 [1,256,64,64] or [256,64,64]; boxes: [K,4] xyxy, input frame.
        Returns (masks bool|float [K,1,H,W], iou [K,1], low_res [K,1,256,256])."""
        h = self._require()
        K = int(boxes.shape[0])
        H, W = int(original_size[0]), int(original_size[1])
        dev = self.device
        emb = embedding.reshape(-1, 64, 64).to(device=dev, dtype=torch.float32).contiguous()
        bx = boxes.to(device=dev, dtype=torch.float32).contiguous()
        iou = torch.empty((K, 1), dtype=torch.float32, device=dev)
        low = torch.empty((K, 1, 256, 256), dtype=torch.float32, device=dev)
        masks = None if return_logits else torch.empty((K, 1, H, W), dtype=torch.uint8, device=dev)
        logits = torch.empty((K, 1, H, W), dtype=torch.float32, device=dev) if return_logits else None
        if K > 0:
            with torch.cuda.device(dev):
                _lib.check(_lib.load().inksam_decode_boxes(
                    h, _ptr(emb), _ptr(bx), K, int(input_size[0]), int(input_size[1]), H, W, _ptr(masks), _ptr(iou),
                    _ptr(low), _ptr(logits), ctypes.c_float(self.mask_threshold), _stream(dev)))
        return (logits if return_logits else masks.view(torch.bool)), iou, low


    @torch.no_grad()
    def decode_batch(self, embeddings: torch.Tensor, boxes: Sequence[torch.Tensor], input_sizes: Sequence[Tuple[int, int]],
                     original_sizes: Sequence[Tuple[int, int]], masks_out: Optional[torch.Tensor] = None, packed: bool = False):
        """The boxes of B images in one pass of the decoder (the `main.py --dir` form of predict_torch).
        embeddings: float32 [B,256,64,64]; boxes[i]: [K_i,4] xyxy in image i's input frame (device or host).
        Returns (masks, iou [sum K], low_res [sum K,256,256], offsets) where `masks` is ONE flat uint8 device tensor
        holding image after image K_i maps of original_sizes[i] (written into `masks_out` when given) and
        offsets[i] is image i's first element.  packed=True: the maps are bit-packed, H x ceil(W/8) bytes each, leftmost
        pixel in the MSB (numpy.packbits / PIL mode "1" rows, the form InkLayer/runner.py:51 converts masks to)."""
        h = self._require()
        dev, B = self.device, len(boxes)
        if embeddings.shape[0] != B or len(input_sizes) != B or len(original_sizes) != B:
            raise RuntimeError("decode_batch: embeddings, boxes and sizes disagree on the number of images")
        emb = embeddings.to(device=dev, dtype=torch.float32).contiguous()
        counts = (ctypes.c_int * B)(*[int(b.shape[0]) for b in boxes])
        in_hw = (ctypes.c_int * (2 * B))(*[int(v) for s in input_sizes for v in s])
        or_hw = (ctypes.c_int * (2 * B))(*[int(v) for s in original_sizes for v in s])
        total = sum(counts)
        offsets, n = [], 0
        for i in range(B):
            offsets.append(n)
            H, W = int(original_sizes[i][0]), int(original_sizes[i][1])
            n += counts[i] * H * ((W + 7) // 8 if packed else W)
        if all(b.device == dev for b in boxes):
            bx = torch.cat([b.reshape(-1, 4).to(torch.float32) for b in boxes]).contiguous()
        else:  # one pinned staging copy for all host boxes
            hb = torch.cat([b.reshape(-1, 4).to(dtype=torch.float32, device="cpu") for b in boxes]).pin_memory()
            bx = hb.to(dev, non_blocking=True)
        if masks_out is None:
            masks_out = torch.empty(max(n, 1), dtype=torch.uint8, device=dev)
        elif masks_out.numel() < n or masks_out.dtype != torch.uint8 or masks_out.device != dev:
            raise RuntimeError("decode_batch: masks_out must be a uint8 tensor on the model's device with room for all masks")
        iou = torch.empty((total,), dtype=torch.float32, device=dev)
        low = torch.empty((total, 256, 256), dtype=torch.float32, device=dev)
        if total > 0:
            with torch.cuda.device(dev):
                if packed:
                    _lib.check(_lib.load().inksam_decode_batch_packed(
                        h, _ptr(emb), B, _ptr(bx), counts, in_hw, or_hw, _ptr(masks_out), _ptr(iou), _ptr(low),
                        ctypes.c_float(self.mask_threshold), _stream(dev)))
                else:
                    _lib.check(_lib.load().inksam_decode_batch(
                        h, _ptr(emb), B, _ptr(bx), counts, in_hw, or_hw, _ptr(masks_out), _ptr(iou), _ptr(low), _ptr(None),
                        ctypes.c_float(self.mask_threshold), _stream(dev)))
        return masks_out, iou, low, offsets

    @torch.no_grad()
    def forward(self, batched_input, multimask_output: bool):
        """Sam.forward (modeling/sam.py:53-131): the batched entry point.  batched_input: list of dicts with 'image'
        ([3,h,w], already resized to the model's input frame, values 0..255), 'original_size' (H, W) and 'boxes' ([K,4]
        xyxy in the input frame).  All images are encoded as one batch and every box of every image is decoded in one
        pass.  Returns one dict per image: 'masks' bool [K,1,H,W], 'iou_predictions' [K,1], 'low_res_logits'
        [K,1,256,256].  Box prompts with multimask_output=False only (the surface InkLayer uses)."""
        if multimask_output:
            raise NotImplementedError("multimask_output=True is not part of the InkLayer hot path (segmentor/sam.py:35)")
        S, dev, B = self.cfg.img_size, self.device, len(batched_input)
        for rec in batched_input:
            if "point_coords" in rec or "mask_inputs" in rec:
                raise NotImplementedError("point and mask prompts are not part of the InkLayer hot path (segmentor/sam.py:32-34)")
            if rec.get("boxes") is None:
                raise NotImplementedError("Sam.forward needs 'boxes' for every image")
            im = rec["image"]
            if im.dim() != 3 or im.shape[0] != 3 or max(im.shape[-2:]) > S:
                raise RuntimeError(f"'image' must be [3,h,w] with h, w <= {S}, got {tuple(im.shape)}")
        if B == 0:
            return []
        in_sizes = [tuple(int(v) for v in rec["image"].shape[-2:]) for rec in batched_input]
        sizes = [tuple(int(v) for v in rec["original_size"]) for rec in batched_input]
        hh, ww = max(s[0] for s in in_sizes), max(s[1] for s in in_sizes)
        stack = torch.zeros((B, 3, hh, ww), dtype=torch.float32, device=dev)
        for i, rec in enumerate(batched_input):
            stack[i, :, :in_sizes[i][0], :in_sizes[i][1]] = rec["image"].to(device=dev, dtype=torch.float32)
        feats = self.encode(stack, in_sizes)
        boxes = [rec["boxes"].reshape(-1, 4) for rec in batched_input]
        masks, iou, low, offs = self.decode_batch(feats, boxes, in_sizes, sizes)
        out, k0 = [], 0
        for i in range(B):
            K, (H, W) = int(boxes[i].shape[0]), sizes[i]
            m = masks[offs[i]:offs[i] + K * H * W].view(torch.bool).view(K, 1, H, W)
            out.append({"masks": m, "iou_predictions": iou[k0:k0 + K].view(K, 1), "low_res_logits": low[k0:k0 + K].view(K, 1, 256, 256)})
            k0 += K
        return out

    __call__ = forward


# ---------------------------------------------------------------------------------------- factories
def _build(name: str, checkpoint: Optional[str]) -> Sam:
    cfg = CONFIGS[name]
    sam = Sam(cfg)
    if checkpoint is not None:
        with open(checkpoint, "rb") as f:
            sd = torch.load(f, map_location="cpu", weights_only=True)  # a state_dict of tensors, never pickled code
        sam.load_state_dict(sd)
    else:
        # The reference returns torch's default random init here; there are no nn.Modules to initialise in this
        # implementation, so checkpoint=None gives the seeded synthetic state_dict (weights.py) instead.
        sam.load_state_dict(synthetic_state_dict(cfg, 0))
    return sam


def build_sam_vit_h(checkpoint: Optional[str] = None) -> Sam:
    return _build("vit_h", checkpoint)


def build_sam_vit_l(checkpoint: Optional[str] = None) -> Sam:
    return _build("vit_l", checkpoint)


def build_sam_vit_b(checkpoint: Optional[str] = None) -> Sam:
    return _build("vit_b", checkpoint)


build_sam = build_sam_vit_h

sam_model_registry = {
    "default": build_sam_vit_h,
    "vit_h": build_sam_vit_h,
    "vit_l": build_sam_vit_l,
    "vit_b": build_sam_vit_b,
}


def build_sam_from_state_dict(state_dict: Dict[str, torch.Tensor], cfg: Optional[SamConfig] = None) -> Sam:
    """Any encoder shape the kernels support (head_dim 64 or 80), inferred from the tensors when cfg is None."""
    sam = Sam(cfg or config_from_state_dict(state_dict))
    sam.load_state_dict(state_dict)
    return sam
