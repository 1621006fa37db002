"""`main.py --dir` for the segmentation stage: the sorted sketch list sharded by image across the GPUs of one box.

The reference walks a directory as a serial loop of independent `run_inklayer_pipeline` calls (main.py:27-32), each
of which rebuilds SAM and calls `run_SAM` (InkLayer/runner.py:49-50, InkLayer/segmentor/sam.py:16-43).  Sketches are
independent, so the path shards by image with NO data-path collective:

  * one process per GPU (torchrun), every rank holds a full weight replica;
  * rank r of W takes images i == r (mod W) of the list in the reference's order (all *.png sorted, then all *.jpg
    sorted, main.py:28);
  * each rank streams its shard through `SketchSegmenter.segment_stream` in batches;
  * masks are gathered host-side only (`gather_results`, `torch.distributed.gather_object`), or every rank writes its
    own `masks/mask_{i}.png` files exactly where the reference writes them (runner.py:57-60).

The segmenter is injected, so the host-side logic (listing, sharding, batching, gathering, file hand-off) is testable
on CPU under the gloo backend with a stand-in segmenter; the product segmenter needs a B200.
"""
from __future__ import annotations

import glob
import os
from typing import Callable, Dict, Iterable, Iterator, List, Optional, Sequence, Tuple

import numpy as np

BoxesFn = Callable[[str, np.ndarray], np.ndarray]  # (path, image) -> float32 [K,4] pixel xyxy (the detector's output)


def list_sketches(directory: str) -> List[str]:
    """main.py:28: sorted *.png followed by sorted *.jpg."""
    return sorted(glob.glob(os.path.join(directory, "*.png"))) + sorted(glob.glob(os.path.join(directory, "*.jpg")))


def shard_indices(n: int, rank: int, world: int) -> List[int]:
    """Indices of the images rank `rank` of `world` processes: i == rank (mod world)."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError(f"bad rank/world {rank}/{world}")
    return list(range(rank, n, world))


def load_sketch(path: str) -> np.ndarray:
    """runner.py:23,49 + segmentor/sam.py:24-25: PIL RGB, then the reference's COLOR_BGR2RGB swap of an RGB array (the
    model sees BGR; identical for grey sketches)."""
    from PIL import Image

    rgb = np.array(Image.open(path).convert("RGB"))
    return np.ascontiguousarray(rgb[..., ::-1])


def batched(items: Sequence, size: int) -> Iterator[Sequence]:
    for i in range(0, len(items), size):
        yield items[i:i + size]


def segment_shard(paths: Sequence[str], boxes_fn: BoxesFn, segmenter, rank: int = 0, world: int = 1,
                  batch: Optional[int] = None, loader: Callable[[str], np.ndarray] = load_sketch,
                  sink: Optional[Callable[[int, object], None]] = None) -> Dict[int, List[np.ndarray]]:
    """Runs this rank's shard of `paths` through `segmenter.segment_stream`; returns {global image index: list of K bool
    masks [H,W]} in the form `run_SAM` returns them (a list per image; empty list for an image without boxes).
    With `sink`, every image's result (bool arrays, or PackedMasks from a packed segmenter) is handed to
    sink(global index, masks) as soon as its batch completes and nothing is accumulated (a 4096-sketch directory is
    64 GB of byte masks)."""
    mine = shard_indices(len(paths), rank, world)
    batch = batch or getattr(segmenter, "max_batch", 1)
    groups = list(batched(mine, batch))

    def feed() -> Iterable[Tuple[List[np.ndarray], List[np.ndarray]]]:
        for g in groups:
            imgs = [loader(paths[i]) for i in g]
            yield imgs, [np.asarray(boxes_fn(paths[i], im), dtype=np.float32).reshape(-1, 4) for i, im in zip(g, imgs)]

    out: Dict[int, List[np.ndarray]] = {}
    for g, res in zip(groups, segmenter.segment_stream(feed())):
        for i, masks in zip(g, res):
            if sink is not None:
                sink(i, masks)
            else:
                out[i] = [m for m in (masks.unpack() if hasattr(masks, "unpack") else masks)]
    return out


def gather_results(local: Dict[int, List[np.ndarray]], n: int, rank: int = 0, world: int = 1, pack: bool = True):
    """Host-side gather of every rank's masks to rank 0 (the only cross-rank traffic of the path; it is not on the GPU
    data path).  Masks travel bit-packed.  Returns the list of per-image mask lists on rank 0, None elsewhere."""
    def enc(masks):
        if not pack:
            return masks
        return [(m.shape, np.packbits(m.reshape(-1))) for m in masks]

    def dec(items):
        if not pack:
            return items
        return [np.unpackbits(b)[: int(np.prod(s))].astype(np.bool_).reshape(s) for s, b in items]

    payload = {i: enc(m) for i, m in local.items()}
    if world == 1:
        parts = [payload]
    else:
        import torch.distributed as dist

        parts = [None] * world if rank == 0 else None
        dist.gather_object(payload, parts, dst=0)
        if rank != 0:
            return None
    merged: Dict[int, list] = {}
    for p in parts:
        merged.update(p)
    missing = [i for i in range(n) if i not in merged]
    if missing:
        raise RuntimeError(f"gather_results: no masks for images {missing[:8]}")
    return [dec(merged[i]) for i in range(n)]


def write_masks(out_base_dir: str, path: str, masks) -> str:
    """The hand-off format to the refinement stage: `<out>/<name>/masks/mask_{i}.png`, PIL mode "1"
    (runner.py:22,51,55-60)."""
    from PIL import Image

    name = os.path.basename(path).split(".")[0]
    masks_dir = os.path.join(out_base_dir, name, "masks")
    os.makedirs(masks_dir, exist_ok=True)
    if hasattr(masks, "to_pil"):  # PackedMasks: the mode "1" image is built from the packed rows directly
        for i in range(len(masks)):
            masks.to_pil(i).save(os.path.join(masks_dir, f"mask_{i}.png"))
        return masks_dir
    for i, m in enumerate(masks):
        Image.fromarray(np.asarray(m, dtype=bool)).save(os.path.join(masks_dir, f"mask_{i}.png"))
    return masks_dir


def run_directory(directory: str, out_dir: str, boxes_fn: BoxesFn, segmenter=None, write: bool = True):
    """`main.py --dir DIR --out_dir OUT` for the SAM stage under torchrun (or a single process).  Reads RANK / WORLD_SIZE
    from the environment; builds the B200 segmenter on this rank's GPU when none is given (checkpoint path from
    $INKSAM_CHECKPOINT, else the reference's default models/sam_vit_h_4b8939.pth; a missing file raises).
    Returns (paths, {global index: masks directory}) when writing, else (paths, {global index: list of bool masks})."""
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    paths = list_sketches(directory)
    if segmenter is None:
        import torch

        from .pipeline import default_ckpt, segmenter_for_checkpoint

        local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(local)
        # a missing checkpoint raises (InkLayer/segmentor/sam.py:17-19 stops there too); never synthetic weights
        segmenter = segmenter_for_checkpoint(os.environ.get("INKSAM_CHECKPOINT", default_ckpt), torch.device("cuda", local),
                                             max_batch=8)
    if not write:
        return paths, segment_shard(paths, boxes_fn, segmenter, rank, world)
    written: Dict[int, str] = {}  # masks go to disk as their batch completes; nothing is accumulated in host memory

    def sink(i: int, masks) -> None:
        written[i] = write_masks(out_dir, paths[i], masks)

    segment_shard(paths, boxes_fn, segmenter, rank, world, sink=sink)
    return paths, written
