"""-m gpu: BASELINE.json's configurations AS STATED, against vectors made by the unmodified reference
(tests/golden/make_golden.py):

  * config 1 -- the real data/bunny_cook_sketch.png (750x750 RGBA -> RGB, InkLayer/runner.py:49), the first 16
    GroundingDINO boxes of output/bunny_cook_sketch/bboxes.json converted by InkLayer/utils/processing.py:6-28, SAM ViT-H,
    entered through `run_SAM` (InkLayer/segmentor/sam.py:16-43) -- the golden masks come from the reference's own run_SAM;
  * the configuration bench.py times -- ViT-H, 8 distinct sketches per batch through Sam.encode(graph=True) (capture and
    replay) + decode_batch with 16 boxes per sketch -- against the reference's vectors for sketch 0.

Stated tolerances (bf16 operands vs the fp32 reference): embeddings / low-res logits rel-RMS <= 1e-2, IoU >= 0.99 per box."""
import numpy as np
import pytest
import torch

from conftest import golden_case_inputs, golden_masks, load_golden, mask_iou
from gpu_util import rel_rms
from inklayer_b200 import CONFIGS, SamPredictor, build_sam_from_state_dict
from inklayer_b200.weights import synthetic_state_dict

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def vit_h_seed0():
    cfg = CONFIGS["vit_h"]
    return build_sam_from_state_dict(synthetic_state_dict(cfg, 0), cfg).to("cuda")


def test_run_sam_on_the_bunny_png_matches_the_reference_run_sam(vit_h_seed0, tmp_path):
    from PIL import Image

    from inklayer_b200 import pipeline
    from inklayer_b200.pipeline import SketchSegmenter, run_SAM

    gold = load_golden("vit_h_750_bunny")
    _, _, img_rgb, _ = golden_case_inputs("vit_h_750_bunny")
    pil = Image.fromarray(img_rgb)
    boxes_filt = torch.from_numpy(gold["run_sam_boxes_filt"])     # what process_boxes_ours hands to run_SAM
    ref = golden_masks(gold)[:, 0]                                 # the reference run_SAM's own return value
    seg = SketchSegmenter(vit_h_seed0, max_batch=1)
    masks = run_SAM(pil, boxes_filt, segmenter=seg)
    assert isinstance(masks, list) and len(masks) == 16
    assert all(m.dtype == np.bool_ and m.shape == (750, 750) for m in masks)
    ious = mask_iou(np.stack(masks), ref)
    print(f"config 1 (bunny PNG, ViT-H, 16 boxes) through run_SAM: min IoU {ious.min():.4f}, mean {ious.mean():.4f}")
    assert ious.min() >= 0.99
    # zero boxes: the reference returns a 2-tuple of empty arrays (segmentor/sam.py:20-21)
    empty = run_SAM(pil, torch.zeros(0, 4), segmenter=seg)
    assert isinstance(empty, tuple) and len(empty) == 2 and all(isinstance(e, np.ndarray) and e.size == 0 for e in empty)
    # a missing checkpoint stops the call (the reference drops into breakpoint() there, sam.py:17-19): no synthetic weights
    with pytest.raises(FileNotFoundError):
        run_SAM(pil, boxes_filt, sam_checkpoint=str(tmp_path / "absent.pth"))
    with pytest.raises(FileNotFoundError):
        run_SAM(pil, boxes_filt)  # default_ckpt (models/sam_vit_h_4b8939.pth) does not exist on this box
    assert pipeline.default_ckpt.endswith("sam_vit_h_4b8939.pth")


def test_run_sam_with_a_checkpoint_file(tmp_path):
    """build_sam(checkpoint=path) reads the reference's checkpoint format (a flat state_dict, build_sam.py:103-106); the
    model of a checkpoint is built once and cached by path.  ViT-B keeps the file small."""
    from PIL import Image

    from inklayer_b200 import build_sam_vit_b, pipeline
    from inklayer_b200.pipeline import run_SAM, segmenter_for_checkpoint

    cfg = CONFIGS["vit_b"]
    sd = synthetic_state_dict(cfg, 0)
    ck = str(tmp_path / "sam_vit_b_synth.pth")
    torch.save(sd, ck)
    gold = load_golden("vit_b_750_bunny")
    _, _, img, boxes = golden_case_inputs("vit_b_750_bunny")
    seg = pipeline.SketchSegmenter(build_sam_vit_b(checkpoint=ck).to("cuda"), max_batch=1)
    # golden was made on `img` as the model sees it: undo run_SAM's swap on the way in
    masks = run_SAM(Image.fromarray(np.ascontiguousarray(img[..., ::-1])), boxes, segmenter=seg)
    assert mask_iou(np.stack(masks), golden_masks(gold)[:, 0]).min() >= 0.99
    # the registry / cache keyed by path (vit_h builder on a vit_b file must fail loudly, not load garbage)
    with pytest.raises(RuntimeError):
        segmenter_for_checkpoint(ck)
    assert not pipeline._segmenters


def test_predict_numpy_entry(vit_h_seed0):
    """SamPredictor.predict (SA/predictor.py:92-166): one box in the ORIGINAL frame, numpy in / numpy out."""
    gold = load_golden("vit_h_750_bunny")
    _, _, img_rgb, _ = golden_case_inputs("vit_h_750_bunny")
    pred = SamPredictor(vit_h_seed0)
    with pytest.raises(RuntimeError):
        pred.predict(box=gold["boxes"][0], multimask_output=False)
    pred.set_image(np.ascontiguousarray(img_rgb[..., ::-1]))
    ref = golden_masks(gold)
    for k in (0, 5, 15):
        m, i, l = pred.predict(box=gold["boxes"][k], multimask_output=False)
        assert isinstance(m, np.ndarray) and m.dtype == np.bool_ and m.shape == (1, 750, 750)
        assert i.shape == (1,) and l.shape == (1, 256, 256)
        assert mask_iou(m, ref[k]).min() >= 0.99
        assert abs(float(i[0]) - float(gold["iou"][k, 0])) < 2e-2
        assert rel_rms(torch.from_numpy(l), torch.from_numpy(gold["low"][k])) <= 1e-2
    lg, _, _ = pred.predict(box=gold["boxes"][0], multimask_output=False, return_logits=True)
    assert lg.dtype == np.float32 and np.array_equal(lg > 0, pred.predict(box=gold["boxes"][0], multimask_output=False)[0])
    with pytest.raises(NotImplementedError):
        pred.predict(point_coords=np.zeros((1, 2)), point_labels=np.ones(1), multimask_output=False)


def test_timed_configuration_batch8_graph_matches_reference(vit_h_seed0):
    """What bench.py times: 8 distinct 1024^2 sketches, Sam.encode(graph=True) (stream pass, capture, replay) and
    decode_batch over 8 x 16 boxes.  The three encodes agree bit for bit, every image's embedding equals its batch-1
    encode bit for bit, and sketch 0 matches the unmodified reference (golden vit_h_bench0: embedding sub-sample, full
    low-res logits of 4 boxes, all 16 masks)."""
    from inklayer_b200.synth import synthetic_boxes, synthetic_sketch

    sam = vit_h_seed0
    B, K = 8, 16
    sam.reserve(B, B * K)
    imgs = torch.from_numpy(np.stack([synthetic_sketch(i) for i in range(B)])).cuda()
    boxes = [torch.from_numpy(synthetic_boxes(i, K)).cuda() for i in range(B)]
    e_stream = sam.encode(imgs, graph=True).clone()      # first sighting: stream launches
    e_capture = sam.encode(imgs, graph=True).clone()     # second: captured, then replayed once
    e_replay = sam.encode(imgs, graph=True)              # third: replay
    assert len(sam._graphs) == 1
    assert torch.equal(e_stream, e_capture) and torch.equal(e_stream, e_replay)
    for i in range(B):
        assert torch.equal(sam.encode(imgs[i:i + 1]), e_replay[i:i + 1]), f"image {i}: batch-8 differs from batch-1"
    sizes = [(1024, 1024)] * B
    flat, iou, low, offs = sam.decode_batch(e_replay, boxes, sizes, sizes)
    torch.cuda.synchronize()
    gold = load_golden("vit_h_bench0")
    np.testing.assert_array_equal(gold["boxes"], boxes[0].cpu().numpy())
    e = rel_rms(e_replay[0, :, ::4, ::4].cpu(), torch.from_numpy(gold["emb_sub"]))
    lo = rel_rms(low[:4].cpu(), torch.from_numpy(gold["low"][:, 0]))
    m0 = flat[offs[0]:offs[0] + K * 1024 * 1024].view(torch.bool).view(K, 1024, 1024).cpu().numpy()
    ious = mask_iou(m0, golden_masks(gold)[:, 0])
    print(f"timed config, sketch 0 vs reference: emb rel-RMS {e:.2e}  low-res rel-RMS {lo:.2e}  min IoU over 16 boxes {ious.min():.4f}")
    assert e <= 1e-2 and lo <= 1e-2 and ious.min() >= 0.99
    assert np.abs(iou[:K].cpu().numpy() - gold["iou"][:, 0]).max() < 2e-2
    # every other sketch of the batch: same masks as the one-image-at-a-time predictor path
    pred = SamPredictor(sam)
    pred.features, pred.is_image_set = e_replay[3:4], True
    pred.original_size = pred.input_size = (1024, 1024)
    m3, _, _ = pred.predict_torch(None, None, boxes=boxes[3], multimask_output=False)
    got3 = flat[offs[3]:offs[3] + K * 1024 * 1024].view(torch.bool).view(K, 1, 1024, 1024)
    assert torch.equal(got3, m3)


def test_graph_replay_survives_workspace_growth():
    """graphs captured by encode(graph=True) bake in workspace pointers.  Capture at B=2, then a B=4 encode
    grows the library's workspace: the old buffer is retired (not freed), the generation counter moves, and the next
    B=2 call re-captures instead of replaying into retired memory."""
    from inklayer_b200 import _lib
    from inklayer_b200.synth import synthetic_sketch

    cfg = CONFIGS["vit_tiny"]
    sam = build_sam_from_state_dict(synthetic_state_dict(cfg, 0), cfg).to("cuda")
    lib = _lib.load()
    imgs = torch.from_numpy(np.stack([synthetic_sketch(50 + i) for i in range(4)])).cuda()
    two = imgs[:2].clone()
    ref2 = sam.encode(two).clone()
    g0 = int(lib.inksam_workspace_generation(sam._handle))
    a = sam.encode(two, graph=True).clone()
    b = sam.encode(two, graph=True).clone()   # captured
    assert len(sam._graphs) == 1 and torch.equal(a, ref2) and torch.equal(b, ref2)
    ref4 = sam.encode(imgs).clone()            # larger batch: workspace grows under the captured graph
    assert int(lib.inksam_workspace_generation(sam._handle)) > g0
    c = sam.encode(two, graph=True).clone()   # must NOT replay the stale graph
    assert torch.equal(c, ref2)
    d = sam.encode(two, graph=True).clone()   # re-captured under the new generation
    e = sam.encode(two, graph=True).clone()   # replayed
    assert torch.equal(d, ref2) and torch.equal(e, ref2) and len(sam._graphs) == 1
    assert torch.equal(sam.encode(imgs), ref4)
    # reserve(): sized once, nothing inside the envelope allocates again
    need = sam.reserve(4, 64)
    assert need == int(lib.inksam_workspace_bytes(sam._handle, 4, 64)) > 0
    g1 = int(lib.inksam_workspace_generation(sam._handle))
    sam.encode(imgs); sam.encode(two)
    sam.decode_boxes(ref4[0], torch.tensor([[10., 10., 500., 600.]] * 64, device="cuda"), (1024, 1024), (1024, 1024))
    assert int(lib.inksam_workspace_generation(sam._handle)) == g1


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_two_devices_in_one_process():
    """cudaFuncSetAttribute / cluster occupancy are per device; a second handle on cuda:1 in the same process
    must get its own launch configuration."""
    from inklayer_b200.synth import synthetic_sketch

    cfg = CONFIGS["vit_tiny80"]
    sd = synthetic_state_dict(cfg, 1)
    img = torch.from_numpy(synthetic_sketch(1))[None]
    outs = []
    for d in (0, 1):
        sam = build_sam_from_state_dict(sd, cfg).to(f"cuda:{d}")
        e = sam.encode(img.to(f"cuda:{d}"))
        m, _, _ = sam.decode_boxes(e[0], torch.tensor([[100., 100., 700., 800.]], device=f"cuda:{d}"), (1024, 1024), (1024, 1024))
        outs.append((e.cpu(), m.cpu()))
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])
    moved = build_sam_from_state_dict(sd, cfg).to("cuda:0").to("cuda:1")  # Sam.to() across devices
    assert torch.equal(moved.encode(img.to("cuda:1")).cpu(), outs[0][0])
