"""The numpy-signature drop-in API: same names, argument meaning, ownership and
error behaviour as the reference's shared.processing / postprocess modules, results equal
to the oracle.  Mirrors the reference's own tests/shared/test_processing.py cases."""

import numpy as np
import pytest

from oracle import pipeline as O
from workloads import synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def sp():
    import shared.processing as sp
    return sp


def test_constants(sp):
    from shared.processing.mobilenet_preprocess import MOBILENET_INPUT_SIZE
    from shared.processing.transforms import IMAGENET_MEAN, IMAGENET_STD, LETTERBOX_COLOR
    from shared.processing.yolo_preprocess import YOLO_INPUT_SIZE, YOLO_NORMALIZATION_SCALE

    assert YOLO_INPUT_SIZE == 640 and YOLO_NORMALIZATION_SCALE == 255.0 and MOBILENET_INPUT_SIZE == 224
    assert LETTERBOX_COLOR == (114, 114, 114)
    assert IMAGENET_MEAN.dtype == np.float32 and np.array_equal(IMAGENET_MEAN, O.MEAN)
    assert np.array_equal(IMAGENET_STD, O.STD)


def test_yolo_preprocessor(sp):
    img = synth.frames(1, 1080, 1920, seed=42)[0]
    res = sp.YOLOPreprocessor()(img)
    want, scale, pad, shape = O.yolo_preprocess(img)
    assert res.tensor.shape == (1, 3, 640, 640) and res.tensor.dtype == np.float32
    assert res.tensor.flags["C_CONTIGUOUS"] and np.array_equal(res.tensor, want)
    assert res.scale == scale and res.padding == pad and res.original_shape == shape
    assert sp.YOLOPreprocessor.get_input_shape() == (1, 3, 640, 640)
    assert sp.YOLOPreprocessor.get_input_dtype() == np.float32
    boxes = np.array([[320, 320, 400, 400, 0.92, 0], [100, 200, 250, 350, 0.85, 1]], dtype=np.float32)
    assert np.array_equal(res.scale_boxes_to_original(boxes), O.scale_boxes(boxes, scale, pad, shape))
    batch = sp.YOLOPreprocessor().preprocess_batch(synth.frames(3, 480, 640, seed=1))
    assert batch.shape == (3, 3, 640, 640)


def test_yolo_validation(sp):
    p = sp.YOLOPreprocessor()
    with pytest.raises(ValueError, match="3D array"):
        p(np.zeros((100, 100), dtype=np.uint8))
    with pytest.raises(ValueError, match="3 channels"):
        p(np.zeros((100, 100, 4), dtype=np.uint8))
    with pytest.raises(ValueError, match="uint8"):
        p(np.zeros((100, 100, 3), dtype=np.float32))
    with pytest.raises(ValueError, match="numpy array"):
        p([[1, 2, 3]])


def test_letterbox_and_scale_boxes(sp):
    img = synth.frames(1, 1920, 1080, seed=44)[0]
    boxed, scale, pad = sp.letterbox(img, 640)
    want, wscale, wpad = O.letterbox(img, 640)
    assert boxed.dtype == np.uint8 and np.array_equal(boxed, want) and scale == wscale and pad == wpad
    boxes = np.array([[100, 100, 200, 200, 0.95, 0], [300, 150, 450, 350, 0.87, 1]], dtype=np.float32)
    before = boxes.copy()
    out = sp.scale_boxes(boxes, 0.5, (10, 20), (480, 640))
    assert np.array_equal(boxes, before) and np.array_equal(out, O.scale_boxes(boxes, 0.5, (10, 20), (480, 640)))


def test_imagenet_normalize(sp):
    for value in (0, 255):
        img = np.full((10, 10, 3), value, dtype=np.uint8)
        got = sp.imagenet_normalize(img)
        assert got.dtype == np.float32 and got.shape == (10, 10, 3)
        assert np.allclose(got[0, 0], (value / 255.0 - O.MEAN) / O.STD, rtol=1e-5)
        assert np.array_equal(got, O.imagenet_normalize(img))


def test_mobilenet_preprocessor(sp):
    from shared.processing.mobilenet_preprocess import extract_crop

    p = sp.MobileNetPreprocessor()
    for shape, seed in (((100, 150), 45), ((10, 10), 46), ((500, 800), 47)):
        crop = synth.frames(1, *shape, seed=seed)[0]
        res = p(crop)
        assert res.tensor.shape == (1, 3, 224, 224) and res.tensor.flags["C_CONTIGUOUS"]
        assert res.original_shape == shape and np.array_equal(res.tensor, O.mobilenet_preprocess(crop))
    assert p.preprocess_batch([]).shape == (0, 3, 224, 224)
    crops = [synth.frames(1, 50, 60, seed=s)[0] for s in (1, 2, 3)]
    assert p.preprocess_batch(crops).shape == (3, 3, 224, 224)
    with pytest.raises(ValueError, match="uint8 or float32"):
        p(np.zeros((10, 10, 3), dtype=np.int32))
    with pytest.raises(ValueError, match="Invalid crop dimensions"):
        p(np.zeros((0, 10, 3), dtype=np.uint8))
    with pytest.raises(ValueError, match="3 channels"):
        p(np.zeros((10, 10, 1), dtype=np.uint8))

    image = synth.frames(1, 480, 640, seed=48)[0]
    for box in ([100, 100, 300, 400, 0.9, 0], [-50, -50, 100, 100], [600, 400, 700, 500], [100, 100, 100, 100],
                [10.9, 20.9, 50.2, 60.7]):
        got = extract_crop(image, np.array(box))
        want = O.extract_crop(image, np.array(box))
        assert got.shape == want.shape and np.array_equal(got, want)
    crop = extract_crop(image, np.array([0, 0, 10, 10]))
    crop[:] = 0
    assert image[:10, :10].any()  # a copy, not a view


def test_postprocess(sp):
    from shared.processing.postprocess import apply_nms, parse_yolo_output

    head = synth.head_v5(1, seed=1, objects=4)
    got = parse_yolo_output(head, 0.25, 0.45)
    want, _ = O.detect_v5(head[0], 0.25, 0.45)
    assert got.dtype == np.float32 and np.array_equal(got, want)
    h8 = synth.head_v8_from_v5(synth.head_v5(1, seed=12, anchors=8400))
    assert np.array_equal(parse_yolo_output(h8, 0.5, 0.45), O.detect_v8(h8[0], 0.5, 0.45)[0])
    empty = synth.head_v5(1, seed=9, objects=0)
    out = parse_yolo_output(empty, 0.25, 0.45)
    assert out.shape == (0, 6) and out.dtype == np.float32
    boxes, conf, cid = O.decode_v5(head[0])
    assert apply_nms(boxes, conf, cid, 0.25, 0.45) == O.nms(boxes, conf, cid, 0.25, 0.45)
    assert apply_nms(boxes, conf, cid, 2.0, 0.45) == []
    # more detections than the first-pass row budget: wrapper retries with room for all
    dense = synth.head_v5_dense(1, seed=27, candidates=4000, anchors=6000)
    want, _ = O.detect_v5(dense[0], 0.25, 0.45)
    assert np.array_equal(parse_yolo_output(dense, 0.25, 0.45), want)


def test_end_to_end_like_reference_pipeline(sp):
    """The call sequence of monolithic/app/inference.py:161-193 through the drop-in names."""
    from shared.processing.mobilenet_preprocess import extract_crop
    from shared.processing.postprocess import parse_yolo_output

    frame = synth.frames(1, 480, 640, seed=0)[0]
    head = synth.head_v5(1, seed=1, objects=4)
    want_tensor, want_dets, _, want_orig, want_crops = O.run_image(frame, head[0], 0.25, 0.45)
    res = sp.YOLOPreprocessor()(frame)
    dets = parse_yolo_output(head, 0.25, 0.45)
    orig = res.scale_boxes_to_original(dets)
    mob = sp.MobileNetPreprocessor()
    crops = np.concatenate([mob(extract_crop(frame, d)).tensor for d in orig], axis=0)
    assert np.array_equal(res.tensor, want_tensor) and np.array_equal(dets, want_dets)
    assert np.array_equal(orig, want_orig) and np.array_equal(crops, want_crops)


def test_float32_inputs(sp):
    """The float32 branch (reference mobilenet_preprocess.py:204-206, transforms.py:264-267;
    exercised by the reference's test_processing.py:165-170 and :561-573)."""
    rng = np.random.default_rng(42)
    p = sp.MobileNetPreprocessor()
    for shape in ((100, 100), (448, 448), (224, 224), (37, 301)):
        for scale in (1.0, 255.0):  # [0,1] data stays as is, [0,255] data is divided by 255
            crop = (rng.random((*shape, 3)) * scale).astype(np.float32)
            res = p(crop)
            assert res.tensor.shape == (1, 3, 224, 224) and res.tensor.dtype == np.float32
            assert np.array_equal(res.tensor, O.mobilenet_preprocess(crop)), (shape, scale)
    img = rng.random((10, 10, 3)).astype(np.float32)
    assert np.array_equal(sp.imagenet_normalize(img), O.imagenet_normalize(img))
    img255 = (img * 255).astype(np.float32)
    assert np.array_equal(sp.imagenet_normalize(img255), O.imagenet_normalize(img255))
    with_nan = img.copy()
    with_nan[0, 0, 0] = np.nan
    got, want = sp.imagenet_normalize(with_nan), O.imagenet_normalize(with_nan)
    assert np.array_equal(got, want, equal_nan=True)


def test_scale_boxes_float64(sp):
    boxes = np.array([[320.5, 320.25, 400.75, 400.1, 0.92, 0], [-10, 700, 650, 10, 0.85, 1]], dtype=np.float64)
    for sc, pad, shape in [(0.3333333333333333, (0, 140), (1080, 1920)), (1.0, (0, 80), (480, 640))]:
        got = sp.scale_boxes(boxes, sc, pad, shape)
        assert got.dtype == np.float64 and np.array_equal(got, O.scale_boxes(boxes, sc, pad, shape))
    with pytest.raises(TypeError):
        sp.scale_boxes(boxes.astype(np.int32), 1.0, (0, 0), (10, 10))


def test_extract_crop_truncates_in_the_callers_dtype(sp):
    """Reference mobilenet_preprocess.py:256: ``map(int, box[:4])`` on the caller's own values —
    float64 coordinates that float32 cannot hold, integer boxes, and what int() refuses."""
    from shared.processing.mobilenet_preprocess import extract_crop
    img = synth.frames(1, 300, 400, seed=77)[0]
    just_below = np.array([99.99999999, 10.0, 150.0, 60.99999999, 0.9, 0], dtype=np.float64)
    assert np.float32(just_below[0]) == np.float32(100.0)  # float32 would round it up to 100
    got = extract_crop(img, just_below)
    assert got.shape == (50, 51, 3) and np.array_equal(got, img[10:60, 99:150])
    assert np.array_equal(got, O.extract_crop(img, just_below))
    big = np.array([16777217.0, 0.0, 16777300.0, 10.0], dtype=np.float64)  # > 2**24: outside the frame
    assert extract_crop(img, big).shape == (1, 1, 3)
    ints = np.array([20, 30, 120, 90], dtype=np.int64)
    assert np.array_equal(extract_crop(img, ints), img[30:90, 20:120])
    assert np.array_equal(extract_crop(img, [-5, -5, 1000, 1000]), img)
    got = extract_crop(img, ints)
    got[:] = 0
    assert img[30:90, 20:120].any()  # a copy, not a view
    with pytest.raises(ValueError):
        extract_crop(img, np.array([np.nan, 0.0, 10.0, 10.0]))
    with pytest.raises(OverflowError):
        extract_crop(img, np.array([np.inf, 0.0, 10.0, 10.0]))


def test_apply_nms_takes_any_integer_class_ids(sp):
    """postprocess.py:116 groups with np.unique, so ids beyond the kernel's 10-bit key, negative
    ids and int64 ids must behave like small ones (ascending class order, same kept boxes)."""
    from shared.processing.postprocess import apply_nms
    rng = np.random.default_rng(5)
    n = 60
    boxes = np.column_stack([rng.uniform(150, 490, n), rng.uniform(150, 490, n), rng.uniform(80, 250, n),
                             rng.uniform(80, 250, n)]).astype(np.float32)
    scores = rng.permutation(n).astype(np.float32) / n * 0.7 + 0.3
    small = rng.integers(0, 4, n)
    wild = np.array([-7, 1024, 5000, 2 ** 40], dtype=np.int64)[small]
    keep_small = apply_nms(boxes, scores, small, 0.25, 0.45)
    keep_wild = apply_nms(boxes, scores, wild, 0.25, 0.45)
    assert keep_wild == keep_small == O.nms(boxes, scores, small, 0.25, 0.45)
    assert keep_wild == O.nms(boxes, scores, wild, 0.25, 0.45)


def test_load_image_from_bytes_uses_the_device_decoder_and_keeps_cv2_pixels(sp, golden, monkeypatch):
    """The drop-in name the services call (reference transforms.py:77-110): baseline JPEGs take the
    GPU decoder, everything else cv2 on the host; pixels and error messages are the reference's."""
    cv2 = pytest.importorskip("cv2")
    from shared.processing import jpeg as jpeg_mod
    from shared.processing.transforms import load_image_from_bytes

    calls = []
    real = jpeg_mod.decode_jpeg_batch_cuda
    monkeypatch.setattr(jpeg_mod, "decode_jpeg_batch_cuda", lambda blobs, *a, **k: (calls.append(len(blobs)), real(blobs, *a, **k))[1])

    def cv2_rgb(data):
        return cv2.cvtColor(cv2.imdecode(np.frombuffer(data, np.uint8), cv2.IMREAD_COLOR), cv2.COLOR_BGR2RGB)

    for i in range(8):  # files the unmodified reference decoded when the golden vectors were made
        data = golden[f"jpeg{i}_bytes"].tobytes()
        got = load_image_from_bytes(data)
        assert got.dtype == np.uint8 and got.flags["C_CONTIGUOUS"]
        assert np.array_equal(got, golden[f"jpeg{i}_rgb"]) and np.array_equal(got, cv2_rgb(data))
    assert len(calls) == 8  # all eight went through the device
    img = synth.frames(1, 48, 64, seed=5)[0]
    smooth = (img // 8 + np.arange(64, dtype=np.uint8)[None, :, None] * 2).astype(np.uint8)
    png = cv2.imencode(".png", cv2.cvtColor(smooth, cv2.COLOR_RGB2BGR))[1].tobytes()
    assert np.array_equal(load_image_from_bytes(png), smooth) and len(calls) == 8  # not a JPEG: host only
    prog = cv2.imencode(".jpg", smooth, [cv2.IMWRITE_JPEG_QUALITY, 90, cv2.IMWRITE_JPEG_PROGRESSIVE, 1])[1].tobytes()
    assert np.array_equal(load_image_from_bytes(prog), cv2_rgb(prog))  # refused by the device, decoded by cv2
    damaged = (__import__("pathlib").Path(__file__).parent / "golden" / "jpeg_idct_overflow_case.jpg").read_bytes()
    assert np.array_equal(load_image_from_bytes(damaged), cv2_rgb(damaged))  # flagged by the device -> cv2's pixels
    assert np.array_equal(load_image_from_bytes(bytearray(golden["jpeg0_bytes"].tobytes())), golden["jpeg0_rgb"])
    with pytest.raises(ValueError, match="empty input"):
        load_image_from_bytes(b"")
    with pytest.raises(ValueError, match="Failed to decode"):
        load_image_from_bytes(b"not an image at all")
    with pytest.raises(ValueError, match="Failed to decode"):
        load_image_from_bytes(golden["jpeg0_bytes"].tobytes()[:200])  # a JPEG cut inside its headers


def test_load_image_reads_files_like_cv2_imread(sp, golden, tmp_path):
    """load_image (reference transforms.py:45-74): JPEG files through the device decoder, other
    formats, damaged files and unreadable paths exactly as cv2.imread leaves them."""
    cv2 = pytest.importorskip("cv2")
    from shared.processing.transforms import load_image

    def imread_rgb(path):
        return cv2.cvtColor(cv2.imread(str(path), cv2.IMREAD_COLOR), cv2.COLOR_BGR2RGB)

    for i in (0, 3, 7):
        path = tmp_path / f"g{i}.jpg"
        path.write_bytes(golden[f"jpeg{i}_bytes"].tobytes())
        got = load_image(str(path))
        assert np.array_equal(got, golden[f"jpeg{i}_rgb"]) and np.array_equal(got, imread_rgb(path))
    img = synth.frames(1, 40, 56, seed=9)[0]
    png = tmp_path / "x.png"
    cv2.imwrite(str(png), cv2.cvtColor(img, cv2.COLOR_RGB2BGR))
    assert np.array_equal(load_image(str(png)), img)
    damaged = tmp_path / "damaged.jpg"
    damaged.write_bytes((__import__("pathlib").Path(__file__).parent / "golden" / "jpeg_idct_overflow_case.jpg").read_bytes())
    assert np.array_equal(load_image(str(damaged)), imread_rgb(damaged))
    cut = tmp_path / "cut.jpg"
    data = golden["jpeg0_bytes"].tobytes()
    cut.write_bytes(data[: len(data) // 2])  # cv2 conceals or refuses: whatever it does is the answer
    want = cv2.imread(str(cut), cv2.IMREAD_COLOR)
    if want is None:
        with pytest.raises(ValueError, match="Failed to load image"):
            load_image(str(cut))
    else:
        assert np.array_equal(load_image(str(cut)), cv2.cvtColor(want, cv2.COLOR_BGR2RGB))
    with pytest.raises(ValueError, match="Failed to load image: /no/such/file.jpg"):
        load_image("/no/such/file.jpg")


def test_results_come_back_in_recycled_pinned_blocks(monkeypatch):
    """_hostio.to_host: large results are page-locked arrays whose blocks go back to torch's caching
    host allocator when the caller drops them; beyond the cap they are ordinary arrays; values are
    the same either way."""
    import gc
    from shared.processing import YOLOPreprocessor, _hostio
    img = synth.frames(1, 480, 640, seed=5)[0]
    want, *_ = O.yolo_preprocess(img)
    gc.collect()
    before = _hostio.pinned_result_bytes()
    res = YOLOPreprocessor()(img)
    assert np.array_equal(res.tensor, want) and res.tensor.flags.writeable
    assert _hostio.pinned_result_bytes() - before == res.tensor.nbytes
    kept = res.tensor[:, :, :8].copy()
    del res
    gc.collect()
    assert _hostio.pinned_result_bytes() == before
    assert np.array_equal(kept, want[:, :, :8])
    monkeypatch.setattr(_hostio, "_PINNED_CAP", 0)
    res = YOLOPreprocessor()(img)
    assert np.array_equal(res.tensor, want) and _hostio.pinned_result_bytes() == before
