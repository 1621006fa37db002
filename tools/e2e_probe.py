"""End-to-end throughput of SketchSegmenter.segment_stream (host sketches in, host masks out) against the pipeline's host
knobs: hand-off copy threads and pipeline depth.  Same synthetic workload as bench.py's e2e leg."""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from inklayer_b200 import CONFIGS, build_sam_from_state_dict  # noqa: E402
from inklayer_b200.pipeline import SketchSegmenter  # noqa: E402
from inklayer_b200.synth import synthetic_boxes, synthetic_sketch  # noqa: E402
from inklayer_b200.weights import synthetic_state_dict  # noqa: E402

BATCH, BOXES, STEPS = 8, 16, 24


def main():
    cfg = CONFIGS["vit_h"]
    sam = build_sam_from_state_dict(synthetic_state_dict(cfg, 0), cfg).to("cuda")
    imgs = [synthetic_sketch(i, 1024, 1024) for i in range(BATCH)]
    boxes = [synthetic_boxes(i, BOXES, 1024, 1024) for i in range(BATCH)]
    for threads, depth in ((4, 2), (8, 2), (2, 2), (8, 3), (4, 2), (8, 2)):
        seg = SketchSegmenter(sam, max_batch=BATCH, depth=depth, copy_threads=threads)
        for _ in seg.segment_stream((imgs, boxes) for _ in range(2 * (depth + 1) + 1)):
            pass
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        n = sum(len(r) for r in seg.segment_stream((imgs, boxes) for _ in range(STEPS)))
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        print(f"copy_threads={threads} depth={depth}: {n / dt:7.1f} sketches/s  ({dt / STEPS * 1e3:.2f} ms per batch of {BATCH})", flush=True)
        del seg


if __name__ == "__main__":
    main()
