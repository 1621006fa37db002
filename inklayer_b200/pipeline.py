"""The caller side of the hot path: InkLayer's `run_SAM` adapter and a batched sketch segmenter.

`run_SAM` keeps the reference's signature and return types (InkLayer/segmentor/sam.py:16-43, including its
BGR<->RGB swap and its empty-box return) but builds the model once instead of once per image, and brings all masks
back with one pinned-memory copy instead of one `.cpu()` sync per mask.  `SketchSegmenter.segment_batch` is the
throughput form used by `main.py --dir`-style drivers: B sketches are encoded in one batch, every box of every sketch
is decoded, masks return as numpy bool arrays."""
from __future__ import annotations

from collections import deque
from concurrent.futures import ThreadPoolExecutor
from typing import Iterable, Iterator, List, Optional, Sequence, Tuple

import numpy as np
import torch

from .predictor import SamPredictor
from .resize import ResizeLongestSide
from .sam import Sam, build_sam


class _Slot:
    """One in-flight batch: pinned input staging, device input, device + pinned mask buffers, and its three events."""

    def __init__(self, sam: Sam, max_batch: int) -> None:
        S = sam.image_encoder.img_size
        self.host_img = torch.empty((max_batch, S, S, 3), dtype=torch.uint8).pin_memory()
        self.dev_img = torch.empty((max_batch, S, S, 3), dtype=torch.uint8, device=sam.device)
        self.dev_masks: Optional[torch.Tensor] = None
        self.host_masks: Optional[torch.Tensor] = None
        self.host_raw: Optional[torch.Tensor] = None   # pinned staging of sketches that still need the resize
        self.dev_raw: Optional[torch.Tensor] = None
        self.ev_in, self.ev_done, self.ev_out = torch.cuda.Event(), torch.cuda.Event(), torch.cuda.Event()
        self.views: list = []

    def raw_buffers(self, nbytes: int, device) -> None:
        if self.host_raw is None or self.host_raw.numel() < nbytes:
            self.host_raw = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
            self.dev_raw = torch.empty(nbytes, dtype=torch.uint8, device=device)

    def mask_buffers(self, nbytes: int, device) -> None:
        if self.dev_masks is None or self.dev_masks.numel() < nbytes:
            self.dev_masks = torch.empty(nbytes, dtype=torch.uint8, device=device)
            self.host_masks = torch.empty(nbytes, dtype=torch.uint8).pin_memory()


class PackedMasks:
    """The K masks of one sketch, bit-packed: `bits` is uint8 [K, H, ceil(W/8)], leftmost pixel in the MSB, rows padded to
    whole bytes -- numpy.packbits(axis=-1) and the row layout of PIL mode "1", which is what the reference turns every mask
    into right after run_SAM (InkLayer/runner.py:51, written as mask_{i}.png at :57-60)."""

    __slots__ = ("bits", "height", "width")

    def __init__(self, bits: np.ndarray, height: int, width: int) -> None:
        self.bits, self.height, self.width = bits, height, width

    def __len__(self) -> int:
        return int(self.bits.shape[0])

    def unpack(self) -> np.ndarray:
        """np.bool_ [K, H, W]: the arrays run_SAM returns."""
        if len(self) == 0:
            return np.zeros((0, self.height, self.width), np.bool_)
        return np.unpackbits(self.bits, axis=-1, count=self.width).view(np.bool_)

    def to_pil(self, i: int):
        """PIL mode "1" image of mask i, built from the packed rows without unpacking (runner.py:51)."""
        from PIL import Image

        return Image.frombytes("1", (self.width, self.height), self.bits[i].tobytes())


class SketchSegmenter:
    """Batched, pipelined form of run_SAM for `main.py --dir`-style drivers (main.py:27-32 loops run_SAM per file).

    Each batch of up to `max_batch` sketches is encoded in one pass and every box of every sketch is decoded in one
    pass.  `depth` batches are in flight: the pinned-memory H2D copy of batch i+1 and the D2H copy + host hand-off of
    batch i-1 run on their own streams / a worker thread under the kernels of batch i."""

    def __init__(self, sam: Sam, max_batch: int = 8, depth: int = 2, copy_threads: int = 4, packed: bool = False) -> None:
        """packed=True: results are `PackedMasks` (bit-packed by the post-process kernel: 8x less device->host traffic and
        no host-side packing before the PNG hand-off) instead of np.bool_ arrays."""
        if sam.device.type != "cuda":
            raise RuntimeError("SketchSegmenter needs a Sam on a CUDA device")
        self.sam = sam
        self.packed = packed
        sam.reserve(max_batch, 128)  # workspace sized once: the per-slot encoder graphs never see it move
        self.transform = ResizeLongestSide(sam.image_encoder.img_size)
        self.max_batch = max_batch
        self.depth = max(1, depth)
        self._slots = [_Slot(sam, max_batch) for _ in range(self.depth + 1)]
        with torch.cuda.device(sam.device):
            self._s_in, self._s_run, self._s_out = torch.cuda.Stream(), torch.cuda.Stream(), torch.cuda.Stream()
        self._pool = ThreadPoolExecutor(max_workers=1)        # one hand-off per batch, in order
        self._copy_pool = ThreadPoolExecutor(max_workers=max(1, copy_threads))  # per-image copies out of pinned memory (numpy drops the GIL)
        self._next = 0
        self.h2d_bytes = 0
        self.d2h_bytes = 0

    # ------------------------------------------------------------------ one batch: enqueue everything, no host sync
    def _submit(self, images: Sequence[np.ndarray], boxes: Sequence[np.ndarray]) -> _Slot:
        sam, dev, B = self.sam, self.sam.device, len(images)
        if B < 1 or B > self.max_batch:
            raise RuntimeError(f"a batch holds 1..{self.max_batch} sketches, got {B}")
        slot = self._slots[self._next]
        self._next = (self._next + 1) % len(self._slots)
        # sketches already in the model's input frame go straight into the pinned batch container; the others are staged
        # as they are and resized ON THE DEVICE (Sam.resize_longest_side: bit-exact with the reference's Pillow resize)
        sizes, in_sizes, raw = [], [], []   # raw: (slot index, byte offset, H, W)
        raw_bytes = 0
        for i, im in enumerate(images):
            H, W = int(im.shape[0]), int(im.shape[1])
            h, w = self.transform.get_preprocess_shape(H, W, self.transform.target_length)
            sizes.append((H, W))
            in_sizes.append((h, w))
            if (h, w) == (H, W):
                slot.host_img[i, :h, :w].copy_(torch.from_numpy(np.ascontiguousarray(im)))
            else:
                raw.append((i, raw_bytes, H, W))
                raw_bytes += (H * W * 3 + 255) // 256 * 256
        if raw:
            slot.raw_buffers(raw_bytes, dev)
            for i, off, H, W in raw:
                slot.host_raw[off:off + H * W * 3].view(H, W, 3).copy_(torch.from_numpy(np.ascontiguousarray(images[i])))
        tbs = [self.transform.apply_boxes_torch(torch.as_tensor(np.asarray(bx, dtype=np.float32).reshape(-1, 4)), sizes[i])
               for i, bx in enumerate(boxes)]
        hb = torch.cat(tbs).pin_memory() if sum(len(t) for t in tbs) else torch.zeros((0, 4))
        total = sum(len(t) * s[0] * ((s[1] + 7) // 8 if self.packed else s[1]) for t, s in zip(tbs, sizes))
        slot.mask_buffers(max(total, 1), dev)
        caller = torch.cuda.current_stream(dev)
        self._s_in.wait_stream(caller)
        with torch.cuda.stream(self._s_in):
            if len(raw) < B:
                if not raw:
                    slot.dev_img[:B].copy_(slot.host_img[:B], non_blocking=True)
                else:
                    skip = {r[0] for r in raw}
                    for i in range(B):
                        if i not in skip:
                            slot.dev_img[i].copy_(slot.host_img[i], non_blocking=True)
            if raw:
                slot.dev_raw[:raw_bytes].copy_(slot.host_raw[:raw_bytes], non_blocking=True)
            db = hb.to(dev, non_blocking=True)
            slot.ev_in.record()
        self.h2d_bytes += (B - len(raw)) * slot.host_img[0].numel() + sum(H * W * 3 for _, _, H, W in raw) + hb.numel() * 4
        with torch.cuda.stream(self._s_run):
            self._s_run.wait_event(slot.ev_in)
            for i, off, H, W in raw:
                sam.resize_longest_side(slot.dev_raw[off:off + H * W * 3].view(H, W, 3), out=slot.dev_img[i])
            feats = sam.encode(slot.dev_img[:B], in_sizes, graph=True)  # per-slot input buffer: replayed from a CUDA graph
            counts = [len(t) for t in tbs]
            splits = list(torch.split(db, counts)) if db.numel() else [db[:0] for _ in tbs]
            _, _, _, offs = sam.decode_batch(feats, splits, in_sizes, sizes, masks_out=slot.dev_masks, packed=self.packed)
            db.record_stream(self._s_run)
            slot.ev_done.record()
        with torch.cuda.stream(self._s_out):
            self._s_out.wait_event(slot.ev_done)
            slot.host_masks[:total].copy_(slot.dev_masks[:total], non_blocking=True)
            slot.ev_out.record()
        self.d2h_bytes += total
        slot.views = [(offs[i], counts[i], sizes[i]) for i in range(B)]
        return slot

    def _collect(self, slot: _Slot) -> List[np.ndarray]:
        """Waits for the batch's D2H copy and hands every image's masks to the caller as its own bool array."""
        slot.ev_out.synchronize()
        host = slot.host_masks.numpy()

        def own(view):
            off, k, (H, W) = view
            if self.packed:
                p = (W + 7) // 8
                return PackedMasks(host[off:off + k * H * p].reshape(k, H, p).copy(), H, W)
            return host[off:off + k * H * W].view(np.bool_).reshape(k, H, W).copy()

        return list(self._copy_pool.map(own, slot.views))

    # ------------------------------------------------------------------ public API
    @torch.no_grad()
    def segment_stream(self, batches: Iterable[Tuple[Sequence[np.ndarray], Sequence[np.ndarray]]]) -> Iterator[List[np.ndarray]]:
        """batches: iterable of (images, boxes) with up to max_batch sketches each (see segment_batch for the element
        types).  Yields, in order, one list of bool arrays [K_i,H,W] (or PackedMasks, when built with packed=True) per
        batch; results trail the input by `depth` batches so that copies overlap compute."""
        pending: deque = deque()
        for images, boxes in batches:
            assert len(images) == len(boxes)
            pending.append(self._pool.submit(self._collect, self._submit(images, boxes)))
            if len(pending) > self.depth:
                yield pending.popleft().result()
        while pending:
            yield pending.popleft().result()

    @torch.no_grad()
    def segment_batch(self, images: Sequence[np.ndarray], boxes: Sequence[np.ndarray]) -> List[np.ndarray]:
        """images: uint8 [H,W,3] each (RGB as the model should see it); boxes: float [K_i,4] pixel xyxy in the ORIGINAL
        frame of image i.  Returns one bool array [K_i,H,W] per image."""
        assert len(images) == len(boxes)
        chunks = ((images[b0:b0 + self.max_batch], boxes[b0:b0 + self.max_batch])
                  for b0 in range(0, len(images), self.max_batch))
        out: List[np.ndarray] = []
        for res in self.segment_stream(chunks):
            out.extend(res)
        return out


def get_model_path(filename: str) -> str:
    """InkLayer/utils/paths.py:4-6 resolves `<InkLayer checkout>/models/<filename>`; this package is not inside that
    checkout, so the directory is `$INKLAYER_MODELS` when set, else `./models`."""
    import os

    return os.path.join(os.environ.get("INKLAYER_MODELS", "models"), filename)


default_ckpt = get_model_path("sam_vit_h_4b8939.pth")  # InkLayer/segmentor/sam.py:13
_segmenters: dict = {}  # absolute checkpoint path -> SketchSegmenter (the reference rebuilds the model on every call)


def segmenter_for_checkpoint(sam_checkpoint: str, device="cuda", max_batch: int = 1) -> SketchSegmenter:
    """The cached segmenter of one checkpoint file.  A missing file raises: the reference stops in `breakpoint()` there
    (segmentor/sam.py:17-19); there is no silent fall-back to synthetic weights (tests and bench.py build those explicitly
    with build_sam_from_state_dict(synthetic_state_dict(...)) and pass `segmenter=`)."""
    import os

    if sam_checkpoint is None or not os.path.exists(sam_checkpoint):
        raise FileNotFoundError(f"Checkpoint not found at {sam_checkpoint}")
    key = (os.path.abspath(sam_checkpoint), str(device), max_batch)
    seg = _segmenters.get(key)
    if seg is None:
        seg = SketchSegmenter(build_sam(checkpoint=sam_checkpoint).to(device), max_batch=max_batch, packed=True)
        _segmenters[key] = seg
    return seg


def run_SAM(image_pil, boxes_filt, sam_checkpoint: Optional[str] = None, segmenter: Optional[SketchSegmenter] = None):
    """Drop-in for InkLayer/segmentor/sam.py:16-43.  image_pil: PIL RGB image; boxes_filt: float tensor/array [K,4]
    pixel xyxy (InkLayer/utils/processing.py:6-28).  Returns a list of K np.bool_ arrays [H,W], or the reference's 2-tuple
    of empty arrays for K == 0 (sam.py:20-21).  `sam_checkpoint` defaults to `default_ckpt` like the reference's; the
    model of a checkpoint is built once and kept, not rebuilt per image.  `segmenter` (an extension) supplies an already
    built SketchSegmenter instead of a checkpoint."""
    if segmenter is None:
        segmenter = segmenter_for_checkpoint(default_ckpt if sam_checkpoint is None else sam_checkpoint)
    if len(boxes_filt) == 0:
        return np.array([]), np.array([])
    image = np.array(image_pil)[..., ::-1]  # the reference applies COLOR_BGR2RGB to an RGB array (sam.py:24-25)
    bx = boxes_filt.detach().cpu().numpy() if isinstance(boxes_filt, torch.Tensor) else np.asarray(boxes_filt)
    masks = segmenter.segment_batch([np.ascontiguousarray(image)], [bx])[0]
    if isinstance(masks, PackedMasks):
        masks = masks.unpack()
    return [m for m in masks]
