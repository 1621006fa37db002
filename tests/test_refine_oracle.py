"""CPU: pins oracle/refine_oracle.py (numpy / scipy restatement) to the golden vectors produced by the reference's own
cv2-based functions (tests/golden/make_refine_golden.py), bit for bit; plus the host box rules of inklayer_b200.refine."""
import numpy as np
import pytest

from conftest import load_golden
from inklayer_b200 import refine as P
from inklayer_b200.synth import refine_case
from oracle import refine_oracle as R

CLEAN = ["300x200", "750x750", "1024x1000"]
NMS = ["512_k12", "600x750_k9", "drop_mid"]


def unpack(bits, shape):
    return np.unpackbits(bits)[: int(np.prod(shape))].reshape(shape).astype(bool)


@pytest.mark.parametrize("name", CLEAN)
def test_clean_up_mask_matches_reference(name):
    g = load_golden("refine_clean_" + name)
    K, H, W = g["shape"]
    masks = refine_case(int(g["seed"]), int(K), int(H), int(W))[0]
    np.testing.assert_array_equal(masks > 127, unpack(g["in_bits"], g["shape"]))
    want = unpack(g["out_bits"], g["shape"])
    for k in range(K):
        out = R.clean_up_mask(masks[k])
        assert set(np.unique(out)) <= {0, 255}
        np.testing.assert_array_equal(out > 0, want[k])
    assert not np.array_equal(want, masks > 127)  # the fixture does exercise the cleaning


@pytest.mark.parametrize("name", NMS)
def test_sketch_nms_matches_reference(name):
    g = load_golden("refine_nms_" + name)
    K, H, W = (int(v) for v in g["shape"])
    masks, sketch, bboxes, scores = refine_case(int(g["seed"]), K, H, W)
    if "full_box_at" in g and int(g["full_box_at"]) >= 0:  # the middle box the filter drops (make_refine_golden.py)
        bboxes[int(g["full_box_at"])] = [0, 0, W - 1, H - 1]
    keep = R.sketch_nms(sketch, bboxes, scores, masks, float(g["thr"]))
    np.testing.assert_array_equal(np.asarray(keep, dtype=np.int64), g["keep"])
    inter = R.content_iou_matrix(masks, sketch)
    for (i, j), s_iou, b_iou, better in zip(g["pairs"], g["sketch_iou"], g["bbox_iou"], g["better"]):
        got = R.pair_scores(inter, bboxes, scores, int(i), int(j), H, W)
        assert got[0] == pytest.approx(float(s_iou), abs=1e-12) and got[1] == pytest.approx(float(b_iou), abs=1e-12)
        assert got[2] == int(better)
        # the product's host rule gives the same answer from the same matrix
        assert P.content_iou(inter, bboxes, scores, int(i), int(j), H, W) == got


def test_host_box_rules_agree_with_oracle():
    rng = np.random.default_rng(5)
    sketch = refine_case(3, 1, 200, 240)[1]
    boxes = rng.uniform(0, 200, (20, 4)); boxes[:, 2:] = boxes[:, :2] + rng.uniform(2, 120, (20, 2))
    boxes = boxes.round()
    np.testing.assert_array_equal(P.filter_full_or_empty_bbox(sketch, boxes), R.filter_full_or_empty_bbox(sketch, boxes))
    for a, b in zip(boxes[:-1], boxes[1:]):
        assert P.compute_bbox_iou(a, b) == R.compute_bbox_iou(a, b)
        assert P.share_corner(a, b, 9.0) == R.share_corner(a, b, 9.0)
        assert P.is_contained_bbox(a, b, 9.0) == bool(R.is_contained_bbox(a, b, 9.0))
    assert P.calculate_kernel_size((1024, 1024)) == R.calculate_kernel_size((1024, 1024)) == (25, 25)
    assert P.calculate_kernel_size((300, 200)) == (5, 5) and P.calculate_kernel_size((750, 750)) == (19, 19)
