"""CPU tests of the host side: the C-ABI library loads and exports every symbol include/*.h declares (no compute calls
without a GPU), the drop-in surface's host logic (ResizeLongestSide, state_dict checks, loud failure without a GPU),
and the directory driver (listing, mask hand-off files)."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

from conftest import ROOT
from inklayer_b200 import CONFIGS, ResizeLongestSide, Sam, _lib, build_sam_from_state_dict
from inklayer_b200 import driver
from inklayer_b200.weights import state_dict_shapes, synthetic_state_dict
from oracle import sam_oracle as O


def declared_functions():
    names = []
    for hdr in ("inksam.h", "inksam_ops.h", "inksam_refine.h"):
        src = open(os.path.join(ROOT, "include", hdr)).read()
        src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
        names += re.findall(r"\b(inksam_[a-z0-9_]+)\s*\(", src)
    return sorted(set(names))


def test_library_exports_every_declared_symbol():
    lib = ctypes.CDLL(_lib.LIB_PATH)
    names = declared_functions()
    assert {"inksam_create", "inksam_encode", "inksam_decode_boxes", "inksam_decode_batch", "inksam_destroy"} <= set(names)
    for n in names:
        assert hasattr(lib, n), f"{n} is declared in include/ but not exported by libinksam.so"
    # and the ctypes table of the shim covers exactly the declared surface
    assert set(_lib.SIGNATURES) <= set(names)
    for n in names:
        if n not in ("inksam_create", "inksam_destroy", "inksam_load_tensor"):
            assert n in _lib.SIGNATURES, f"{n} has no typed binding in inklayer_b200/_lib.py"


def test_no_cpu_fallback():
    cfg = CONFIGS["vit_tiny"]
    sam = build_sam_from_state_dict(synthetic_state_dict(cfg, 0), cfg)
    with pytest.raises(RuntimeError):
        sam.to("cpu")
    with pytest.raises(RuntimeError):
        sam.encode(torch.zeros(1, 1024, 1024, 3, dtype=torch.uint8))
    if not torch.cuda.is_available():
        with pytest.raises(Exception):
            sam.to("cuda")


def test_state_dict_checks_like_the_reference():
    cfg = CONFIGS["vit_tiny"]
    sd = synthetic_state_dict(cfg, 0)
    sam = Sam(cfg)
    bad = dict(sd)
    bad.pop("image_encoder.pos_embed")
    with pytest.raises(RuntimeError, match="missing"):
        sam.load_state_dict(bad)
    bad = dict(sd)
    bad["image_encoder.pos_embed"] = torch.zeros(1, 8, 8, cfg.embed_dim)
    with pytest.raises(RuntimeError, match="size mismatch"):
        sam.load_state_dict(bad)
    sam.load_state_dict(sd)
    assert set(sam.state_dict()) == {k for k, _, _ in state_dict_shapes(cfg)}


@pytest.mark.parametrize("hw", [(750, 750), (600, 750), (1024, 512), (333, 1000), (1024, 1024), (2000, 1500)])
def test_resize_longest_side_matches_oracle(hw):
    """SA/utils/transforms.py:26-31,67-102 through the oracle's restatement (Pillow bilinear on uint8 is exact)."""
    t = ResizeLongestSide(1024)
    assert t.get_preprocess_shape(hw[0], hw[1], 1024) == O.preprocess_shape(hw[0], hw[1], 1024)
    rng = np.random.default_rng(hw[0])
    img = rng.integers(0, 256, hw + (3,), dtype=np.uint8)
    np.testing.assert_array_equal(t.apply_image(img), O.resize_image(img))
    boxes = torch.from_numpy(rng.uniform(0, min(hw), (5, 4)).astype(np.float32))
    keep = boxes.clone()
    np.testing.assert_array_equal(t.apply_boxes_torch(boxes, hw).numpy(), O.apply_boxes(boxes, hw).numpy())
    assert torch.equal(boxes, keep)  # deep copy, the caller's boxes are untouched (transforms.py:77)
    np.testing.assert_allclose(t.apply_boxes(boxes.numpy(), hw), O.apply_boxes(boxes, hw).numpy(), rtol=1e-6)


def test_directory_listing_and_mask_files(tmp_path):
    from PIL import Image

    for name in ["b.png", "a.png", "c.jpg", "notes.txt"]:
        p = tmp_path / name
        if name.endswith("txt"):
            p.write_text("x")
        else:
            Image.fromarray(np.full((8, 8, 3), 255, np.uint8)).save(p)
    paths = driver.list_sketches(str(tmp_path))
    assert [os.path.basename(p) for p in paths] == ["a.png", "b.png", "c.jpg"]  # main.py:28 order
    img = driver.load_sketch(paths[0])
    assert img.shape == (8, 8, 3) and img.dtype == np.uint8
    rng = np.random.default_rng(0)
    masks = [rng.random((8, 8)) > 0.5 for _ in range(3)]
    d = driver.write_masks(str(tmp_path / "out"), paths[1], masks)
    assert d.endswith(os.path.join("out", "b", "masks"))
    for i, m in enumerate(masks):  # runner.py:57-60 format; mask_cleaner.py reads them back as 0/255
        back = np.array(Image.open(os.path.join(d, f"mask_{i}.png")))
        np.testing.assert_array_equal(back.astype(bool), m)


def test_shard_indices_partition():
    for n in (0, 1, 7, 4096):
        for world in (1, 2, 3, 8):
            parts = [driver.shard_indices(n, r, world) for r in range(world)]
            assert sorted(i for p in parts for i in p) == list(range(n))
            assert max(len(p) for p in parts) - min(len(p) for p in parts) <= 1
    with pytest.raises(ValueError):
        driver.shard_indices(4, 2, 2)
