"""The N>1 path on CPU: two gloo processes shard a sketch list by image, stream their shards through a stand-in
segmenter (same `segment_stream` contract as SketchSegmenter; the real one needs a B200) and gather masks host-side
on rank 0.  There is no collective on the data path, so this covers all of the multi-GPU host logic."""
import os
import socket

import numpy as np
import torch.distributed as dist
import torch.multiprocessing as mp

from inklayer_b200 import driver

N_IMAGES = 11


class FakeSegmenter:
    """Deterministic masks from the image content and the boxes, batch by batch, yielded lazily like the real one."""

    max_batch = 3

    def __init__(self):
        self.batches = []

    def segment_stream(self, batches):
        for images, boxes in batches:
            assert 1 <= len(images) <= self.max_batch
            self.batches.append(len(images))
            out = []
            for im, bx in zip(images, boxes):
                H, W = im.shape[:2]
                yy, xx = np.mgrid[0:H, 0:W]
                out.append(np.stack([(xx >= b[0]) & (xx < b[2]) & (yy >= b[1]) & (yy < b[3]) & (im[..., 0] > 0)
                                     for b in bx]) if len(bx) else np.zeros((0, H, W), bool))
            yield out


def fake_image(path):
    i = int(os.path.basename(path).split(".")[0])
    rng = np.random.default_rng(i)
    return rng.integers(0, 2, (16 + i, 20, 3), dtype=np.uint8) * 255


def fake_boxes(path, image):
    i = int(os.path.basename(path).split(".")[0])
    if i == 4:
        return np.zeros((0, 4), np.float32)  # a sketch without detections
    rng = np.random.default_rng(100 + i)
    k = 1 + i % 3
    lo = rng.integers(0, 8, (k, 2))
    return np.concatenate([lo, lo + rng.integers(2, 10, (k, 2))], 1).astype(np.float32)


def expected():
    paths = [f"{i:03d}.png" for i in range(N_IMAGES)]
    seg = FakeSegmenter()
    res = driver.segment_shard(paths, fake_boxes, seg, 0, 1, loader=fake_image)
    return paths, [res[i] for i in range(N_IMAGES)]


def worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        paths = [f"{i:03d}.png" for i in range(N_IMAGES)]
        seg = FakeSegmenter()
        local = driver.segment_shard(paths, fake_boxes, seg, rank, world, loader=fake_image)
        assert sorted(local) == driver.shard_indices(N_IMAGES, rank, world)
        assert sum(seg.batches) == len(local) and max(seg.batches) <= seg.max_batch
        full = driver.gather_results(local, N_IMAGES, rank, world)
        if rank == 0:
            q.put([[m.copy() for m in masks] for masks in full])
        else:
            assert full is None
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_two_rank_shard_and_gather_matches_single_process():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = q.get(timeout=120)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    _, want = expected()
    assert len(got) == N_IMAGES
    for g, w in zip(got, want):
        assert len(g) == len(w)
        for a, b in zip(g, w):
            assert a.dtype == np.bool_ and a.shape == b.shape
            np.testing.assert_array_equal(a, b)


def test_gather_single_process_roundtrip_and_missing():
    paths, want = expected()
    full = driver.gather_results({i: m for i, m in enumerate(want)}, N_IMAGES)
    for g, w in zip(full, want):
        for a, b in zip(g, w):
            np.testing.assert_array_equal(a, b)
    import pytest

    with pytest.raises(RuntimeError):
        driver.gather_results({0: want[0]}, 2)
