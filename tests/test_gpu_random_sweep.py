"""Seeded random sweeps: frame shapes / targets for K1, rectangles for K4, head sizes and
thresholds for K2+K3 — every case compared bit for bit with the oracle.  These hunt for
geometry edge cases (extreme aspect ratios, one-pixel sides, slot/step boundaries in the
staging logic) that the hand-picked cases might miss."""

import numpy as np
import pytest
import torch

from oracle import pipeline as O
from workloads import synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def D():
    from shared.processing import device
    return device


def _dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def test_k1_random_shapes(D):
    rng = np.random.default_rng(1234)
    done = 0
    while done < 48:
        h, w = int(rng.integers(1, 1400)), int(rng.integers(1, 1400))
        target = int(rng.choice([640, 640, 320, 416, 97, 224]))
        scale = min(target / h, target / w)
        if int(h * scale) < 1 or int(w * scale) < 1:
            continue  # cv2 rejects an empty resize
        frames = synth.frames(2, h, w, seed=1000 + done)
        got = D.preprocess_batch_cuda(_dev(frames), target=target).cpu().numpy()
        want, *_ = O.yolo_preprocess(frames[1], target=target)
        assert np.array_equal(got[1:2], want), (h, w, target)
        u8 = D.letterbox_batch_cuda(_dev(frames[:1]), target).cpu().numpy()[0]
        assert np.array_equal(u8, O.letterbox(frames[0], target)[0]), (h, w, target)
        done += 1


def test_k4_random_rectangles(D):
    rng = np.random.default_rng(4321)
    for case, (h, w) in enumerate([(480, 640), (1080, 1920), (97, 1500), (1500, 61), (2160, 3840)]):
        n = 2
        frames = synth.frames(n, h, w, seed=2000 + case)
        rects = []
        for _ in range(24):
            x1, y1 = rng.integers(0, w), rng.integers(0, h)
            kind = rng.integers(0, 5)
            if kind == 0:  # full-width / full-height slabs
                x1, x2, y2 = 0, w, min(h, y1 + rng.integers(1, 60))
            elif kind == 1:
                y1, y2, x2 = 0, h, min(w, x1 + rng.integers(1, 60))
            elif kind == 2:  # tiny
                x2, y2 = min(w, x1 + rng.integers(1, 4)), min(h, y1 + rng.integers(1, 4))
            elif kind == 3:  # exact special sizes when they fit
                s = int(rng.choice([224, 448]))
                x1, y1 = min(x1, max(0, w - s)), min(y1, max(0, h - s))
                x2, y2 = min(w, x1 + s), min(h, y1 + s)
            else:
                x2, y2 = min(w, x1 + rng.integers(1, w)), min(h, y1 + rng.integers(1, h))
            rects.append([rng.integers(0, n), x1, y1, max(x2, x1 + 1), max(y2, y1 + 1), 0, 0, 0])
        rois = np.array(rects, dtype=np.int32)
        rois[:, 3] = np.minimum(rois[:, 3], w)
        rois[:, 4] = np.minimum(rois[:, 4], h)
        got = D.resize_normalize_rois_cuda(_dev(frames), _dev(rois)).cpu().numpy()
        for r, (img, x1, y1, x2, y2, *_rest) in enumerate(rois):
            want = O.mobilenet_preprocess(np.ascontiguousarray(frames[img, y1:y2, x1:x2]))
            assert np.array_equal(got[r : r + 1], want), (h, w, rois[r].tolist())


def test_k2k3_random_heads(D):
    rng = np.random.default_rng(99)
    for case in range(10):
        anchors = int(rng.integers(1, 5000))
        classes = int(rng.choice([1, 2, 7, 80]))
        cand = int(rng.integers(0, anchors + 1))
        conf, iou = float(rng.choice([0.25, 0.5, 0.05])), float(rng.choice([0.45, 0.2, 0.7]))
        head = synth.head_v5_dense(2, seed=3000 + case, candidates=cand, anchors=anchors, classes=classes,
                                   used_classes=min(classes, int(rng.integers(1, 9))))
        synth.assert_tie_free(head, conf)
        dets, keep, count, status = D.decode_nms_cuda(_dev(head), conf, iou, max_det=anchors)
        for i in range(2):
            want_dets, want_keep = O.detect_v5(head[i], conf, iou)
            k = int(count[i])
            assert int(status[i]) == 0 and keep[i, :k].cpu().tolist() == want_keep, (anchors, classes, cand)
            assert np.array_equal(dets[i, :k].cpu().numpy(), want_dets)


def test_k2k3_random_heads_many_classes(D):
    """As above with 9..80 classes in play: blocks of 64 sorted candidates then hold more class runs
    than the lazy rows of K3's resolve, which goes straight to the row-parallel path, and most
    rows suppress nothing (the serial walk visits only those that do)."""
    rng = np.random.default_rng(777)
    for case in range(10):
        anchors = int(rng.integers(64, 4000))
        cand = int(rng.integers(1, min(anchors, 700) + 1))
        conf, iou = float(rng.choice([0.25, 0.5])), float(rng.choice([0.45, 0.2, 0.7]))
        head = synth.head_v5_dense(2, seed=3500 + case, candidates=cand, anchors=anchors, classes=80,
                                   used_classes=int(rng.integers(9, 81)))
        synth.assert_tie_free(head, conf)
        dets, keep, count, status = D.decode_nms_cuda(_dev(head), conf, iou, max_det=anchors)
        for i in range(2):
            want_dets, want_keep = O.detect_v5(head[i], conf, iou)
            k = int(count[i])
            assert int(status[i]) == 0 and keep[i, :k].cpu().tolist() == want_keep, (anchors, cand, case)
            assert np.array_equal(dets[i, :k].cpu().numpy(), want_dets)


def test_jpeg_random_files():
    """Random sizes, sampling modes, qualities, restart intervals, table optimisation, content
    (photo-like, noise, flat), orientation tags — one mixed batch per round, against cv2.imdecode
    (the call load_image_from_bytes makes)."""
    cv2 = pytest.importorskip("cv2")  # noqa: F841
    from shared.processing import jpeg as J
    from workloads import jpeg_cases as W

    rng = np.random.default_rng(4321)
    for rnd in range(3):
        blobs = []
        for i in range(20):
            h, w = int(rng.integers(1, 400)), int(rng.integers(1, 600))
            kind = int(rng.integers(0, 4))
            if kind == 0:
                img = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
            elif kind == 1:
                img = np.full((h, w, 3), rng.integers(0, 256, 3), dtype=np.uint8)
            else:
                img = W.photo(h, w, seed=rnd * 100 + i)
            q = int(rng.integers(1, 101))
            if rng.random() < 0.15:
                data = W.encode(img[..., 0], q, restart=int(rng.integers(0, 9)))
            else:
                data = W.encode(img, q, W.SAMPLINGS[int(rng.integers(0, 5))], restart=int(rng.integers(0, 9)),
                                optimize=bool(rng.integers(0, 2)))
            if rng.random() < 0.3:
                data = W.with_exif_orientation(data, int(rng.integers(1, 9)), big_endian=bool(rng.integers(0, 2)))
            blobs.append(data)
        batch = J.decode_jpeg_batch_cuda(blobs)
        torch.cuda.synchronize()
        assert batch.status.cpu().tolist() == [0] * len(blobs)
        for k, (frame, data) in enumerate(zip(batch.frames, blobs)):
            want = W.reference_decode(data)
            assert tuple(frame.shape) == want.shape, (rnd, k)
            assert np.array_equal(frame.cpu().numpy(), want), (rnd, k)
