"""Pins oracle/sam_oracle.py to the reference: golden vectors in tests/golden/*.npz were produced by the
unmodified vendored segment_anything (tests/golden/make_golden.py).  fp32 on both sides, so the only
difference is summation order: tolerances are a few ulp-scale multiples, masks must match bit for bit
except where a logit sits within that noise of zero."""
import numpy as np
import pytest
import torch

from conftest import golden_case_inputs, golden_masks, load_golden, mask_iou
from inklayer_b200.weights import synthetic_state_dict
from oracle import sam_oracle as O


ALL_CASES = ["tiny_1024", "tiny80_600x750", "vit_b_750_bunny", "vit_h_1024", "vit_h_v1_1024", "vit_h_v2_600x750",
             "vit_h_v3_1024x768", "vit_h_bench0", "vit_h_750_bunny"]


@pytest.mark.parametrize("case", ALL_CASES)
def test_oracle_matches_reference_golden(case):
    cfg_name, seed, img, boxes = golden_case_inputs(case)
    gold = load_golden(case)
    if case == "vit_h_750_bunny":
        # config 1: run_SAM hands the model the BGR view of the RGB array (InkLayer/segmentor/sam.py:24-25) and the boxes
        # are what process_boxes_ours made of the normalised json boxes (float32 cx/cy/w/h round trip)
        img = np.ascontiguousarray(img[..., ::-1])
        np.testing.assert_allclose(boxes, gold["boxes"], atol=1e-3, rtol=0)
        boxes = gold["boxes"]
    np.testing.assert_array_equal(boxes, gold["boxes"])
    cfg = O.CONFIGS[cfg_name]
    from inklayer_b200.config import CONFIGS

    pred = O.OraclePredictor(cfg, synthetic_state_dict(CONFIGS[cfg_name], seed))
    pred.set_image(img)
    tb = O.apply_boxes(torch.from_numpy(boxes), img.shape[:2])
    np.testing.assert_allclose(tb.numpy(), gold["boxes_input_frame"], rtol=0, atol=0)
    masks, iou, low = pred.predict_torch(tb)
    emb = pred.features[0].numpy()
    np.testing.assert_allclose(emb[:, ::4, ::4], gold["emb_sub"], atol=2e-4, rtol=0)
    assert abs(np.abs(emb.astype(np.float64)).sum() - gold["emb_abs_sum"]) / gold["emb_abs_sum"] < 1e-5
    # FULL low-res logits of the boxes the fixture keeps, the ::4 sub-sample of all of them, and ALL K masks
    n_low = gold["low"].shape[0]
    np.testing.assert_allclose(low.numpy()[:n_low], gold["low"], atol=2e-5, rtol=0)
    np.testing.assert_allclose(low.numpy()[:, :, ::4, ::4], gold["low_sub"], atol=2e-5, rtol=0)
    np.testing.assert_allclose(iou.numpy(), gold["iou"], atol=1e-5, rtol=0)
    ref_masks = golden_masks(gold)
    assert ref_masks.shape == tuple(masks.shape)
    assert mask_iou(masks.numpy(), ref_masks).min() > 0.9999
    np.testing.assert_allclose(masks.flatten(1).sum(1).numpy(), gold["mask_area"], rtol=2e-4)


def test_oracle_error_convention():
    """SA/predictor.py:213-214: predicting before set_image raises RuntimeError."""
    cfg = O.CONFIGS["vit_tiny"]
    from inklayer_b200.config import CONFIGS

    pred = O.OraclePredictor(cfg, synthetic_state_dict(CONFIGS["vit_tiny"], 0))
    with pytest.raises(RuntimeError):
        pred.predict_torch(torch.zeros(1, 4))


def test_preprocess_shape_rounding():
    """SA/utils/transforms.py:93-102."""
    assert O.preprocess_shape(750, 750, 1024) == (1024, 1024)
    assert O.preprocess_shape(600, 750, 1024) == (819, 1024)
    assert O.preprocess_shape(1024, 512, 1024) == (1024, 512)
    assert O.preprocess_shape(333, 1000, 1024) == (341, 1024)


def test_reference_copy_agrees_with_the_port():
    """oracle/_ref (the unmodified vendored segment_anything placed by oracle/install_ref.py; it is what bench.py's CPU arm
    times, kind "reference") and the port agree on tiny_1024: same embedding, logits and masks to fp32 summation order."""
    from oracle import ref_loader

    if not ref_loader.available():
        pytest.skip("oracle/_ref absent: run `python oracle/install_ref.py` where /root/reference exists")
    from inklayer_b200.config import CONFIGS

    cfg_name, seed, img, boxes = golden_case_inputs("tiny_1024")
    sd = synthetic_state_dict(CONFIGS[cfg_name], seed)
    ref = ref_loader.reference_predictor(CONFIGS[cfg_name], sd)
    with torch.no_grad():
        ref.set_image(img)
        tb = ref.transform.apply_boxes_torch(torch.from_numpy(boxes), img.shape[:2])
        m_ref, i_ref, l_ref = ref.predict_torch(None, None, boxes=tb, multimask_output=False)
    port = O.OraclePredictor(O.CONFIGS[cfg_name], sd)
    port.set_image(img)
    m, i, l = port.predict_torch(O.apply_boxes(torch.from_numpy(boxes), img.shape[:2]))
    np.testing.assert_allclose(port.features.numpy(), ref.features.numpy(), atol=2e-4, rtol=0)
    np.testing.assert_allclose(l.numpy(), l_ref.numpy(), atol=2e-5, rtol=0)
    np.testing.assert_allclose(i.numpy(), i_ref.numpy(), atol=1e-5, rtol=0)
    assert mask_iou(m.numpy(), m_ref.numpy()).min() > 0.9999
    gold = load_golden("tiny_1024")  # and the copy reproduces the committed vectors it was pinned by
    np.testing.assert_allclose(l_ref.numpy()[: gold["low"].shape[0]], gold["low"], atol=1e-6, rtol=0)
