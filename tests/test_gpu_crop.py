"""K4 parity (a6..a9): scale_boxes -> extract_crop -> resize 224 -> ImageNet normalise -> CHW
on the GPU vs the oracle and golden vectors.  Bit-exact float32 tensors and integer rects."""

import numpy as np
import pytest
import torch

from conftest import sha
from oracle import pipeline as O
from workloads import synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def D():
    from shared.processing import device
    return device


def _dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _oracle_crops(frames, dets, count, geom):
    out = []
    for n in range(frames.shape[0]):
        k = int(count[n])
        if k == 0:
            continue
        boxes = O.scale_boxes(dets[n, :k], geom.scale, (geom.pad_w, geom.pad_h), frames.shape[1:3])
        for row in boxes:
            out.append(O.mobilenet_preprocess(O.extract_crop(frames[n], row)))
    return np.concatenate(out, axis=0) if out else np.zeros((0, 3, 224, 224), np.float32)


@pytest.mark.parametrize("h,w", [(480, 640), (1080, 1920), (375, 500)])
def test_fanout_matches_oracle(D, h, w):
    n = 12
    frames = synth.frames(n, h, w, seed=3)
    dets, count = synth.detections(n, seed=31, kmax=8)
    geom = D.letterbox_geometry(h, w, 640)
    crops, offsets, rois = D.crop_batch_cuda(_dev(frames), _dev(dets), _dev(count), geom)
    torch.cuda.synchronize()
    offsets = offsets.cpu().numpy()
    assert np.array_equal(offsets, np.concatenate([[0], np.cumsum(count)]))
    total = int(offsets[-1])
    want = _oracle_crops(frames, dets, count, geom)
    assert np.array_equal(crops[:total].cpu().numpy(), want)


def test_edge_rois(D):
    """Zero-size, 1x1, exact 224 (copy), exact 448 (2x2 box), border-touching and
    out-of-frame boxes (SURVEY §8d C4 edge list)."""
    h, w = 480, 640
    frames = synth.frames(1, h, w, seed=4)
    boxes = np.array([
        [10, 90, 10, 200, 0.9, 0], [10, 90, 11.5, 91.2, 0.9, 0], [100, 180, 324, 404, 0.9, 1],
        [50, 80, 498, 528, 0.9, 2], [0, 80, 640, 560, 0.9, 3], [-50, -50, 30, 130, 0.9, 4],
        [600, 500, 700, 600, 0.9, 5], [639, 559, 640, 560, 0.9, 6], [700, 700, 800, 800, 0.9, 7],
        [30.9, 100.9, 31.1, 300.2, 0.8, 8], [5, 85, 9, 535, 0.8, 9], [0, 300, 640, 303, 0.8, 10],
    ], dtype=np.float32)[None]
    count = np.array([boxes.shape[1]], dtype=np.int32)
    geom = D.letterbox_geometry(h, w, 640)
    crops, offsets, rois = D.crop_batch_cuda(_dev(frames), _dev(boxes), _dev(count), geom)
    want = _oracle_crops(frames, boxes, count, geom)
    assert np.array_equal(crops.cpu().numpy(), want)
    rois = rois.cpu().numpy()
    scaled = O.scale_boxes(boxes[0], geom.scale, (geom.pad_w, geom.pad_h), (h, w))
    for r, row in zip(rois, scaled):
        rect = O.crop_rect(row, h, w)
        assert (r[5] == 1) == (rect is None)
        if rect is not None:
            assert tuple(r[1:5]) == rect


def test_mobilenet_golden(D, golden):
    for case in golden.meta["crops"]:
        crop = synth.frames(1, case["h"], case["w"], seed=case["seed"])
        rois = torch.tensor([[0, 0, 0, case["w"], case["h"], 0, 0, 0]], dtype=torch.int32).cuda()
        t = D.resize_normalize_rois_cuda(_dev(crop), rois).cpu().numpy()
        assert sha(t) == case["tensor_sha"], case


def test_scale_boxes_and_rects_golden(D, golden):
    boxes = golden["box0_in"]
    for j, (sc, pad, shape) in enumerate([(1.0, (0, 80), (480, 640)),
                                          (0.3333333333333333, (0, 140), (1080, 1920)),
                                          (0.5, (10, 0), (1280, 1240))]):
        got = D.scale_boxes_cuda(_dev(boxes), sc, pad, shape).cpu().numpy()
        assert np.array_equal(got, golden[f"box0_scaled{j}"])
    rects = D.extract_crop_rects_cuda(_dev(boxes), 480, 640).cpu().numpy()
    for r, shape in zip(rects, golden["box0_crop_shapes"]):
        got_shape = (1, 1, 3) if r[5] else (r[4] - r[2], r[3] - r[1], 3)
        assert tuple(shape) == got_shape


def test_imagenet_normalize_golden(D, golden):
    got = D.imagenet_normalize_cuda(_dev(golden["normalize_in"])).cpu().numpy()
    assert np.array_equal(got, golden["normalize_out"])


def test_capacity_and_empty(D):
    frames = synth.frames(4, 480, 640, seed=6)
    dets, count = synth.detections(4, seed=32, kmax=8)
    geom = D.letterbox_geometry(480, 640, 640)
    crops, offsets, _ = D.crop_batch_cuda(_dev(frames), _dev(dets), _dev(count), geom, roi_capacity=5)
    want = _oracle_crops(frames, dets, count, geom)
    assert int(offsets[-1]) == int(count.sum())
    assert np.array_equal(crops.cpu().numpy(), want[:5])
    zero = np.zeros_like(count)
    crops, offsets, _ = D.crop_batch_cuda(_dev(frames), _dev(dets), _dev(zero), geom)
    assert int(offsets[-1]) == 0


def test_full_size_c4(D):
    """C4 at full size: 512 frames, ~2048 ROIs; offsets are the exclusive scan of counts and a
    random sample of ROIs equals the oracle."""
    n = 512
    frames = synth.frames(n, 480, 640, seed=3)
    dets, count = synth.detections(n, seed=33, kmax=8)
    geom = D.letterbox_geometry(480, 640, 640)
    crops, offsets, rois = D.crop_batch_cuda(_dev(frames), _dev(dets), _dev(count), geom)
    offsets = offsets.cpu().numpy()
    assert np.array_equal(offsets, np.concatenate([[0], np.cumsum(count)]))
    rng = np.random.default_rng(0)
    for n_i in rng.choice(n, size=24, replace=False):
        want = _oracle_crops(frames[n_i : n_i + 1], dets[n_i : n_i + 1], count[n_i : n_i + 1], geom)
        got = crops[offsets[n_i] : offsets[n_i + 1]].cpu().numpy()
        assert np.array_equal(got, want)


def test_unaligned_frame_buffer_corners(D):
    """Frames that neither start nor end on a 16-byte boundary, with ROIs in the first and the
    last bytes of the buffer: the staging copies must clamp their aligned supersets to the
    buffer and fetch the cut-off bytes one by one."""
    n, h, w = 2, 37, 333
    raw = synth.frames(1, 1, n * h * w + 64, seed=8).reshape(-1)
    buf = _dev(raw)
    for off in (1, 7, 16, 29):
        frames_np = raw[off : off + n * h * w * 3].reshape(n, h, w, 3)
        frames = buf[off : off + n * h * w * 3].view(n, h, w, 3)
        g = D.letterbox_geometry(h, w, 640)
        s = np.float32(g.scale)

        def box(x1, y1, x2, y2):  # original-frame rectangle -> letterbox-space detection row
            return [x1 * s + g.pad_w, y1 * s + g.pad_h, x2 * s + g.pad_w, y2 * s + g.pad_h, 0.9, 0]

        dets = np.zeros((n, 4, 6), np.float32)
        dets[0, 0] = box(0, 0, 40, 20)        # first bytes of the buffer
        dets[0, 1] = box(0, 0, w, h)          # whole first frame
        dets[1, 0] = box(w - 50, h - 9, w, h)  # last bytes of the buffer
        dets[1, 1] = box(w - 3, h - 2, w, h)
        count = np.array([2, 2], np.int32)
        crops, offsets, _ = D.crop_batch_cuda(frames, _dev(dets), _dev(count), g)
        want = _oracle_crops(frames_np, dets, count, g)
        assert np.array_equal(crops[:4].cpu().numpy(), want), off


def test_empty_rois_interleaved_with_real_ones(D):
    """Many ROIs, every other one empty, more work items than resident CTAs: persistent CTAs
    alternate between empty items (no staging, no x table) and real ones."""
    n, h, w, kmax = 16, 480, 640, 8
    frames = synth.frames(n, h, w, seed=12)
    dets, _ = synth.detections(n, seed=35, kmax=kmax)
    rng = np.random.default_rng(3)
    for i in range(n):
        for k in range(kmax):
            if k % 2 == 1:  # empty: x2 <= x1 after truncation
                dets[i, k, :4] = [100.2, 90.0, 100.9, 300.0]
            else:
                c = rng.uniform(60, 580, size=2)
                wh = rng.uniform(30, 300, size=2)
                dets[i, k, :4] = [c[0] - wh[0] / 2, c[1] - wh[1] / 2, c[0] + wh[0] / 2, c[1] + wh[1] / 2]
    count = np.full((n,), kmax, dtype=np.int32)
    geom = D.letterbox_geometry(h, w, 640)
    for _ in range(3):
        crops, offsets, rois = D.crop_batch_cuda(_dev(frames), _dev(dets), _dev(count), geom)
    want = _oracle_crops(frames, dets, count, geom)
    assert int(rois[:, 5].sum()) >= n * kmax // 2
    assert np.array_equal(crops.cpu().numpy(), want)


@pytest.mark.parametrize("out_size", [96, 225, 299, 512])
def test_other_output_sizes(D, out_size):
    """MobileNetPreprocessor(input_size=...) other than 224 (reference mobilenet_preprocess.py:58-112):
    the tile kernel's generic-canvas instantiation, outputs wider than one 224-column tile, sizes
    that are not multiples of 4, custom statistics; rectangles that up-scale, down-scale, copy
    (= out_size) and halve exactly (2 x out_size)."""
    h, w = 1100, 1300
    frames = synth.frames(2, h, w, seed=900 + out_size)
    rects = [[0, 10, 20, 10 + 37, 20 + 53, 0, 0, 0], [1, 0, 0, w, h, 0, 0, 0],
             [0, 5, 7, 5 + out_size, 7 + out_size, 0, 0, 0], [1, 3, 1, 3 + 2 * out_size, 1 + 2 * out_size, 0, 0, 0],
             [0, 100, 100, 100, 200, 1, 0, 0], [1, w - 3, h - 2, w, h, 0, 0, 0], [0, 0, 500, w, 501, 0, 0, 0]]
    rois = np.array(rects, dtype=np.int32)
    rois[:, 3] = np.minimum(rois[:, 3], w)
    rois[:, 4] = np.minimum(rois[:, 4], h)
    mean, std = (0.4, 0.5, 0.3), (0.25, 0.2, 0.3)
    got = D.resize_normalize_rois_cuda(_dev(frames), _dev(rois), out_size, mean, std).cpu().numpy()
    m32, s32 = np.array(mean, np.float32), np.array(std, np.float32)
    for r, (img, x1, y1, x2, y2, empty, *_r) in enumerate(rois):
        crop = np.zeros((1, 1, 3), np.uint8) if empty else np.ascontiguousarray(frames[img, y1:y2, x1:x2])
        resized = O.resize_linear_u8(crop, out_size, out_size)
        want = ((resized.astype(np.float32) / np.float32(255.0) - m32) / s32).transpose(2, 0, 1)
        assert np.array_equal(got[r], want), (out_size, rois[r].tolist())


@pytest.mark.parametrize("n", [1, 3, 70, 300])
def test_k1_linear_mode_batch_sizes(D, n):
    """K1's general linear mode runs in the tile kernel, whose tile height depends on the batch
    size (small batches take shorter tiles): 1, 3, 70 and 300 frames of a COCO shape, every frame
    of the small batches and a sample of the large ones against the oracle."""
    h, w = 333, 500
    frames = synth.frames(min(n, 8), h, w, seed=40 + n)
    batch = np.ascontiguousarray(np.concatenate([frames] * ((n + len(frames) - 1) // len(frames)))[:n])
    got = D.preprocess_batch_cuda(_dev(batch))
    for i in sorted({0, n // 2, n - 1, *range(min(n, 4))}):
        want, *_ = O.yolo_preprocess(batch[i])
        assert np.array_equal(got[i : i + 1].cpu().numpy(), want), (n, i)


def test_widest_frames(D):
    """include/ia_b200.h: two staged rows of the widest crop must fit shared memory.  A 10 000-pixel-wide
    frame (one team per SM, the largest staging buffers) is still bit-exact; 12 000 is refused with
    IA_ERR_UNSUPPORTED before any launch, and the context stays usable."""
    frames = synth.frames(1, 6, 10000, seed=77)
    rois = np.array([[0, 0, 0, 10000, 6, 0, 0, 0], [0, 17, 1, 9999, 5, 0, 0, 0], [0, 9000, 0, 9001, 6, 0, 0, 0]], np.int32)
    got = D.resize_normalize_rois_cuda(torch.from_numpy(frames).cuda(), torch.from_numpy(rois).cuda(), 224).cpu().numpy()
    for r, (_, x1, y1, x2, y2, *_r) in enumerate(rois):
        want = O.mobilenet_preprocess(np.ascontiguousarray(frames[0, y1:y2, x1:x2]))
        assert np.array_equal(got[r:r + 1], want), r
    wide = torch.zeros((1, 2, 12000, 3), dtype=torch.uint8, device="cuda")
    with pytest.raises(RuntimeError, match="unsupported"):
        D.resize_normalize_rois_cuda(wide, torch.tensor([[0, 0, 0, 12000, 2, 0, 0, 0]], dtype=torch.int32, device="cuda"), 224)
    again = D.resize_normalize_rois_cuda(torch.from_numpy(frames).cuda(), torch.from_numpy(rois[:1]).cuda(), 224).cpu().numpy()
    assert np.array_equal(again, got[:1])
