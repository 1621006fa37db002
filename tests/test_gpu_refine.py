"""-m gpu: the refinement kernels (bit-packed close, run-based connected components, stroke-restricted pairwise
intersections) through the C ABI against the reference's golden vectors and the CPU oracle.  Integer work: bit-exact."""
import numpy as np
import pytest
import torch

from conftest import load_golden
from inklayer_b200 import refine as P
from inklayer_b200.synth import refine_case
from oracle import refine_oracle as R

pytestmark = pytest.mark.gpu


def unpack(bits, shape):
    return np.unpackbits(bits)[: int(np.prod(shape))].reshape(shape).astype(bool)


@pytest.mark.parametrize("name", ["300x200", "750x750", "1024x1000"])
def test_clean_masks_matches_reference_golden(name):
    g = load_golden("refine_clean_" + name)
    K, H, W = (int(v) for v in g["shape"])
    masks = refine_case(int(g["seed"]), K, H, W)[0]
    out = P.clean_masks(masks, 127).cpu().numpy()
    assert set(np.unique(out)) <= {0, 255}
    np.testing.assert_array_equal(out > 0, unpack(g["out_bits"], g["shape"]))
    # bool / 0-1 input with thresh 0 is the same mask
    out01 = P.clean_masks(torch.from_numpy(masks > 127), 0).cpu().numpy()
    np.testing.assert_array_equal(out01, out)
    np.testing.assert_array_equal(P.clean_up_mask(masks[0]), out[0])


@pytest.mark.parametrize("H,W,K", [(64, 33, 2), (97, 130, 3), (1024, 1024, 40), (40, 2600, 1)])
def test_clean_masks_matches_oracle(H, W, K):
    """odd widths (row padding bits), more masks than one chunk (32), kernel radius > 32 pixels (W = 2600 -> k = 1... the
    radius comes from min(H, W)), empty and full masks."""
    masks = refine_case(100 + H + W, K, H, W)[0]
    masks[0] = 0
    if K > 1:
        masks[1] = 255
    out = P.clean_masks(masks, 127).cpu().numpy()
    for k in range(min(K, 6)):
        np.testing.assert_array_equal(out[k], R.clean_up_mask(masks[k]), err_msg=f"mask {k}")
    assert out[0].max() == 0 and (K == 1 or out[1].min() == 255)


def test_morphology_large_radius():
    masks = refine_case(7, 1, 2700, 2700)[0]  # k = 67, radius 33: the horizontal pass crosses more than one word
    out = P.clean_masks(masks, 127).cpu().numpy()
    np.testing.assert_array_equal(out[0], R.clean_up_mask(masks[0]))


@pytest.mark.parametrize("name", ["512_k12", "600x750_k9", "drop_mid"])
def test_sketch_nms_matches_reference_golden(name):
    g = load_golden("refine_nms_" + name)
    K, H, W = (int(v) for v in g["shape"])
    masks, sketch, bboxes, scores = refine_case(int(g["seed"]), K, H, W)
    if "full_box_at" in g and int(g["full_box_at"]) >= 0:  # the middle box the filter drops (make_refine_golden.py)
        bboxes[int(g["full_box_at"])] = [0, 0, W - 1, H - 1]
    inter = P.content_iou_matrix(masks, sketch).cpu().numpy()
    np.testing.assert_array_equal(inter, R.content_iou_matrix(masks, sketch))
    keep = P.sketch_nms(sketch, bboxes, scores, torch.from_numpy(masks).cuda(), float(g["thr"]))
    np.testing.assert_array_equal(np.asarray(keep, dtype=np.int64), g["keep"])
    for (i, j), s_iou in zip(g["pairs"], g["sketch_iou"]):
        assert P.content_iou(inter, bboxes, scores, int(i), int(j), H, W)[0] == pytest.approx(float(s_iou), abs=1e-12)


def test_content_iou_properties_at_full_size():
    """size-independent checks at 64 masks of 1024 x 1000 (odd row stride): symmetry, diagonal = stroke-pixel counts,
    inter <= min(area), a mask against itself."""
    masks, sketch, _, _ = refine_case(31, 64, 1024, 1000)
    inter = P.content_iou_matrix(masks, sketch).cpu().numpy().astype(np.int64)
    assert np.array_equal(inter, inter.T)
    area = ((masks > 0) & (sketch < 250)[None]).reshape(64, -1).sum(1)
    np.testing.assert_array_equal(np.diag(inter), area)
    assert (inter <= np.minimum(area[:, None], area[None, :])).all()
    none = P.content_iou_matrix(masks[:4], None).cpu().numpy()
    np.testing.assert_array_equal(np.diag(none), (masks[:4] > 0).reshape(4, -1).sum(1))
