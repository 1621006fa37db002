"""Property tests (hypothesis) over image sizes and box counts -- SURVEY.md 4 asks for them.
CPU: the host-side geometry of the surface (ResizeLongestSide sizes, box transform) against the oracle over
(H, W) in [1, 4096]^2.  GPU (-m gpu): predict_torch output shapes / dtypes for (H, W) in [64, 2048]^2 and K in [0, 200], and
the fused post-process against torch's two-stage F.interpolate (SA/modeling/sam.py:133-162) on random logits."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F
from hypothesis import HealthCheck, given, settings
from hypothesis import strategies as st

from inklayer_b200.resize import ResizeLongestSide
from oracle import sam_oracle as O


@settings(max_examples=300, deadline=None)
@given(H=st.integers(1, 4096), W=st.integers(1, 4096))
def test_preprocess_shape_property(H, W):
    h, w = ResizeLongestSide.get_preprocess_shape(H, W, 1024)
    assert (h, w) == O.preprocess_shape(H, W, 1024)
    assert max(h, w) == 1024 or min(H, W) * 1024 < max(H, W)  # long side lands on 1024 (transforms.py:93-102)
    assert 0 <= h <= 1024 and 0 <= w <= 1024


@settings(max_examples=100, deadline=None)
@given(H=st.integers(16, 4096), W=st.integers(16, 4096), K=st.integers(0, 50), seed=st.integers(0, 2 ** 16))
def test_apply_boxes_property(H, W, K, seed):
    rng = np.random.default_rng(seed)
    b = (rng.random((K, 4)) * np.array([W, H, W, H])).astype(np.float32)
    t = torch.from_numpy(b.copy())
    got = ResizeLongestSide(1024).apply_boxes_torch(t, (H, W))
    want = O.apply_boxes(torch.from_numpy(b.copy()), (H, W))
    assert got.dtype == torch.float32 and tuple(got.shape) == (K, 4)
    assert torch.equal(got, want)
    assert np.array_equal(t.numpy(), b)  # the input is not modified (deep copy, transforms.py:73)


# ------------------------------------------------------------------------------------------------ GPU
_sam = None


def tiny_sam():
    global _sam
    if _sam is None:
        from inklayer_b200 import CONFIGS, build_sam_from_state_dict
        from inklayer_b200.weights import synthetic_state_dict
        cfg = CONFIGS["vit_tiny"]
        _sam = build_sam_from_state_dict(synthetic_state_dict(cfg, 0), cfg).to("cuda")
    return _sam


@pytest.mark.gpu
@settings(max_examples=25, deadline=None, suppress_health_check=list(HealthCheck))
@given(H=st.integers(64, 2048), W=st.integers(64, 2048), K=st.integers(0, 200), seed=st.integers(0, 2 ** 16))
def test_predict_torch_shapes_property(H, W, K, seed):
    """Any original size and any box count (0, ragged chunks of 128, 200): shapes, dtypes, finiteness, and the masks are
    the thresholded logits of the return_logits path."""
    from inklayer_b200 import SamPredictor

    if H * W * K > 96 * 2 ** 20:  # bound the mask tensor of one example
        K = max(0, 96 * 2 ** 20 // (H * W))
    sam = tiny_sam()
    pred = SamPredictor(sam)
    g = torch.Generator().manual_seed(seed)
    pred.features = torch.randn(1, 256, 64, 64, generator=g).cuda()
    pred.original_size = (H, W)
    pred.input_size = pred.transform.get_preprocess_shape(H, W, 1024)
    pred.is_image_set = True
    b = torch.rand(K, 4, generator=g) * torch.tensor([W, H, W, H])
    tb = pred.transform.apply_boxes_torch(b, (H, W)).cuda()
    m, i, l = pred.predict_torch(None, None, boxes=tb, multimask_output=False)
    assert m.dtype == torch.bool and tuple(m.shape) == (K, 1, H, W)
    assert i.dtype == torch.float32 and tuple(i.shape) == (K, 1)
    assert l.dtype == torch.float32 and tuple(l.shape) == (K, 1, 256, 256)
    if K:
        assert torch.isfinite(l).all() and torch.isfinite(i).all()
        k = min(K, 3)
        lg, _, _ = pred.predict_torch(None, None, boxes=tb[:k], multimask_output=False, return_logits=True)
        assert lg.dtype == torch.float32 and tuple(lg.shape) == (k, 1, H, W)
        want = F.interpolate(F.interpolate(l[:k], (1024, 1024), mode="bilinear", align_corners=False)
                             [..., :pred.input_size[0], :pred.input_size[1]], (H, W), mode="bilinear", align_corners=False)
        assert (lg - want).abs().max().item() <= 1e-4 * max(1.0, want.abs().max().item())
        flips = ((lg > 0) != m[:k]).sum().item()
        assert flips == 0 or flips <= 1e-5 * m[:k].numel()  # only logits within rounding of 0 may differ between the two paths
