"""BASELINE.json config 4: `main.py --dir` over N synthetic sketches (default 4096) sharded by image across the GPUs of
one box; host images in, host masks out, through SketchSegmenter.segment_stream (the public streaming API).

    python tools/dir_bench.py [--sketches 4096]                                   # 1 GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        tools/dir_bench.py --sketches 4096                                        # N GPUs

Rank r takes images i == r (mod W) (driver.shard_indices, main.py:28 order).  Sketch i is synthetic_sketch(i % POOL)
with boxes synthetic_boxes(i, 16): the pixel pool is bounded (POOL distinct 1024x1024 canvases per rank) so that 4096
sketches need not sit in host memory; every sketch is still copied H2D, encoded, decoded and copied back.
No collective on the data path: NCCL is used for the start barrier and the max-over-ranks of the elapsed time only."""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

POOL, BATCH, BOXES = 32, 8, 16


def main():
    import torch
    import torch.distributed as dist

    from inklayer_b200 import CONFIGS, build_sam_from_state_dict
    from inklayer_b200.driver import batched, shard_indices
    from inklayer_b200.pipeline import SketchSegmenter
    from inklayer_b200.synth import synthetic_boxes, synthetic_sketch
    from inklayer_b200.weights import synthetic_state_dict

    ap = argparse.ArgumentParser()
    ap.add_argument("--sketches", type=int, default=4096)
    ap.add_argument("--model", default="vit_h")
    ap.add_argument("--hw", type=int, nargs=2, default=[1024, 1024], help="sketch size (H W); other than 1024 exercises the device resize")
    args = ap.parse_args()
    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    cfg = CONFIGS[args.model]
    sam = build_sam_from_state_dict(synthetic_state_dict(cfg, 0), cfg).to(dev)
    seg = SketchSegmenter(sam, max_batch=BATCH)
    mine = shard_indices(args.sketches, rank, world)
    pool = {}

    def sketch(i):
        k = i % POOL
        if k not in pool:
            pool[k] = synthetic_sketch(k, args.hw[0], args.hw[1])
        return pool[k]

    for i in mine[:POOL * world]:
        sketch(i)

    def feed(idx):
        for g in batched(idx, BATCH):
            yield [sketch(i) for i in g], [synthetic_boxes(i, BOXES, args.hw[0], args.hw[1]) for i in g]

    for _ in seg.segment_stream(feed(mine[:3 * BATCH])):   # warm-up: 3 batches
        pass
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    n_masks = 0
    for res in seg.segment_stream(feed(mine)):
        n_masks += sum(len(m) for m in res)
    torch.cuda.synchronize(dev)
    dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    cnt = torch.tensor([n_masks], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
    if rank == 0:
        print(json.dumps({"config": "4: --dir over synthetic sketches sharded by image", "model": args.model,
                          "sketches": args.sketches, "sketch_hw": args.hw, "n_gpus": world, "boxes_per_sketch": BOXES, "masks": int(cnt.item()),
                          "seconds": float(dt.item()), "sketches_per_s": args.sketches / float(dt.item()),
                          "h2d_bytes": seg.h2d_bytes, "d2h_bytes": seg.d2h_bytes, "host_cores": os.cpu_count()}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
