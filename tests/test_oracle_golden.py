"""Pin the oracle (oracle/pipeline.py) against vectors produced by the unmodified
reference (tests/golden/make_golden.py).  Integer/index results must be identical;
float tensors are compared bit for bit through SHA-256 (the oracle uses the same numpy
float32 operations in the same order as the reference)."""

import numpy as np
import pytest

from conftest import sha, strided_sample
from oracle import pipeline as P
from workloads import synth


def test_letterbox_and_yolo_preprocess(golden):
    for i, case in enumerate(golden.meta["letterbox"]):
        img = synth.frames(1, case["h"], case["w"], seed=case["seed"])[0]
        target = case.get("target", 640)
        boxed, scale, pad = P.letterbox(img, target)
        assert scale == case["scale"] and list(pad) == case["pad"]
        assert sha(boxed) == case["u8_sha"], case
        if "tensor_sha" in case:
            t, scale2, pad2, shape = P.yolo_preprocess(img)
            assert t.shape == (1, 3, 640, 640) and t.dtype == np.float32 and t.flags.c_contiguous
            assert sha(t) == case["tensor_sha"], case
            assert np.array_equal(strided_sample(t), golden[f"lb{i}_tensor_sample"])
            assert shape == (case["h"], case["w"]) and scale2 == scale and pad2 == pad


def test_mobilenet_preprocess(golden):
    for i, case in enumerate(golden.meta["crops"]):
        crop = synth.frames(1, case["h"], case["w"], seed=case["seed"])[0]
        t = P.mobilenet_preprocess(crop)
        assert t.shape == (1, 3, 224, 224) and t.dtype == np.float32
        assert np.array_equal(strided_sample(t), golden[f"crop{i}_tensor_sample"]), case
        assert sha(t) == case["tensor_sha"], case


def test_imagenet_normalize(golden):
    got = P.imagenet_normalize(golden["normalize_in"])
    assert got.dtype == np.float32 and np.array_equal(got, golden["normalize_out"])


def test_scale_boxes_and_extract_crop(golden):
    boxes = golden["box0_in"]
    for j, (sc, pad, shape) in enumerate([(1.0, (0, 80), (480, 640)),
                                          (0.3333333333333333, (0, 140), (1080, 1920)),
                                          (0.5, (10, 0), (1280, 1240))]):
        before = boxes.copy()
        got = P.scale_boxes(boxes, sc, pad, shape)
        assert np.array_equal(got, golden[f"box0_scaled{j}"])
        assert np.array_equal(boxes, before)  # input not modified (transforms.py:210)
    frame = synth.frames(1, 480, 640, seed=300)[0]
    for row, shape, digest in zip(boxes, golden["box0_crop_shapes"], golden.meta["box0_crop_sha"]):
        crop = P.extract_crop(frame, row)
        assert crop.shape == tuple(shape) and sha(crop) == digest


HEADS = {
    "std": lambda: synth.head_v5(4, seed=2),
    "small": lambda: synth.head_v5(3, seed=5, anchors=1000, dup=12),
    "dense": lambda: synth.head_v5_dense(1, seed=6, candidates=2000),
    "allpass": lambda: synth.head_v5_dense(1, seed=7, candidates=3000, anchors=3000),
    "oneclass": lambda: synth.head_v5_dense(1, seed=8, candidates=1500, anchors=4000, used_classes=1),
    "empty": lambda: synth.head_v5(2, seed=9, objects=0),
}


@pytest.mark.parametrize("name", list(HEADS))
def test_v5_decode_nms(golden, name):
    head = HEADS[name]()
    for n in range(head.shape[0]):
        dets, keep = P.detect_v5(head[n], 0.25, 0.45)
        assert np.array_equal(np.array(keep, dtype=np.int64), golden[f"head_{name}{n}_keep"])
        assert dets.dtype == np.float32 and dets.shape[1] == 6
        assert np.array_equal(dets, golden[f"head_{name}{n}_dets"])


def test_v8_parse(golden):
    h8 = synth.head_v8_from_v5(synth.head_v5(2, seed=11, anchors=8400))
    for n in range(2):
        dets, _ = P.detect_v8(h8[n], 0.5, 0.45)
        assert np.array_equal(dets, golden[f"v8_{n}_dets"])


CLASSIFIER_CASES = {"cls": (lambda: synth.logits(16, seed=400), 5),
                    "cls_small": (lambda: synth.logits(4, seed=401, classes=10), 3),
                    "cls_k1": (lambda: synth.logits(16, seed=400)[:4], 1)}


@pytest.mark.parametrize("name", list(CLASSIFIER_CASES))
def test_softmax_topk(golden, name):
    """N4: the reference's classify() (run with a stubbed session) vs the oracle.  Indices
    exact; probabilities to 1e-6 relative — numpy picks its float32 exp() by CPU feature, so
    another host may differ from the committed vectors in the last bit."""
    make, k = CLASSIFIER_CASES[name]
    for r, row in enumerate(make()):
        idx, probs = P.softmax_topk(row, k)
        assert idx.tolist() == golden[f"{name}_ids"][r].tolist()
        assert probs.dtype == np.float32
        np.testing.assert_allclose(probs.astype(np.float64), golden[f"{name}_probs"][r], rtol=1e-6, atol=0)
        assert P.top1(row[None]) == (int(np.argmax(row)), float(row.max()))


def test_c1_chain(golden):
    frame = synth.frames(1, 480, 640, seed=0)[0]
    head = synth.head_v5(1, seed=1, objects=4)
    tensor, dets, keep, dets_orig, crops = P.run_image(frame, head[0], 0.25, 0.45)
    assert sha(tensor) == golden.meta["c1"]["tensor_sha"]
    assert np.array_equal(np.array(keep), golden["c1_keep"])
    assert np.array_equal(dets, golden["c1_dets"])
    assert np.array_equal(dets_orig, golden["c1_dets_orig"])
    assert [sha(c[None]) for c in crops] == golden.meta["c1"]["crops_sha"]


def test_validation_messages():
    """Same ValueError substrings as yolo_preprocess.py:185-195 / mobilenet_preprocess.py:195-210."""
    with pytest.raises(ValueError, match="3D array"):
        P.yolo_preprocess(np.zeros((4, 4), np.uint8))
    with pytest.raises(ValueError, match="3 channels"):
        P.yolo_preprocess(np.zeros((4, 4, 4), np.uint8))
    with pytest.raises(ValueError, match="uint8"):
        P.yolo_preprocess(np.zeros((4, 4, 3), np.float32))
    with pytest.raises(ValueError, match="uint8 or float32"):
        P.mobilenet_preprocess(np.zeros((4, 4, 3), np.int32))
    with pytest.raises(ValueError, match="Invalid crop dimensions"):
        P.mobilenet_preprocess(np.zeros((0, 4, 3), np.uint8))


def test_oracle_equals_the_staged_reference_modules():
    """oracle/pipeline.py against the reference's OWN modules (baseline/_ref or /root/reference, via
    oracle/refrun.py) on fresh inputs: the whole chain on uint8 frames, and MobileNetPreprocessor on
    float32 crops — there the reference's cv2 uses Intel IPP and the oracle's IPP flavour must
    give the same bits."""
    from oracle import refrun
    if refrun.find_reference() is None:
        pytest.skip("reference files not staged")
    cv2 = pytest.importorskip("cv2")
    ref = refrun.Reference()
    frames = synth.frames(2, 375, 500, seed=123)
    heads = synth.head_v5(2, seed=124)
    for f, h in zip(frames, heads):
        a, b = ref.run_image(f, h, 0.25, 0.45), P.run_image(f, h, 0.25, 0.45)
        for x, y in zip(a, b):
            assert np.array_equal(np.asarray(x), np.asarray(y))
    if not cv2.ipp.useIPP():
        pytest.skip("this cv2 build has no IPP: float crops follow the generic path")
    rng = np.random.default_rng(5)
    for shape in ((100, 150), (448, 448), (37, 301), (500, 800), (1, 9)):
        for top in (1.0, 255.0):
            crop = (rng.random((*shape, 3)) * top).astype(np.float32)
            assert np.array_equal(ref.mobilenet(crop).tensor, P.mobilenet_preprocess(crop)), (shape, top)
