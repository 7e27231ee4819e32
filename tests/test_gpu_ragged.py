"""Ragged batches: frames of different sizes through K1 and K4 in one call,
vs the oracle applied per image.  Bit-exact."""

import numpy as np
import pytest
import torch

from oracle import pipeline as O
from workloads import synth

pytestmark = pytest.mark.gpu

SHAPES = [(480, 640), (1080, 1920), (427, 640), (480, 640), (500, 375), (1280, 1280), (333, 500), (480, 640),
          (640, 480), (1, 1), (37, 640)]


@pytest.fixture(scope="module")
def D():
    from shared.processing import device
    return device


def test_preprocess_ragged(D):
    frames = [synth.frames(1, h, w, seed=200 + i)[0] for i, (h, w) in enumerate(SHAPES)]
    out = D.preprocess_ragged_cuda([torch.from_numpy(f).cuda() for f in frames]).cpu().numpy()
    assert out.shape == (len(frames), 3, 640, 640)
    for i, f in enumerate(frames):
        want, *_ = O.yolo_preprocess(f)
        assert np.array_equal(out[i : i + 1], want), SHAPES[i]


def test_preprocess_ragged_many_of_one_shape(D):
    """More images of one shape than fit one launch's pointer table (256)."""
    base = synth.frames(3, 60, 80, seed=7)
    frames = [torch.from_numpy(base[i % 3]).cuda() for i in range(300)]
    out = D.preprocess_ragged_cuda(frames, target=96).cpu().numpy()
    for i in (0, 1, 2, 255, 256, 257, 299):
        want, *_ = O.yolo_preprocess(base[i % 3], target=96)
        assert np.array_equal(out[i : i + 1], want)


def test_preprocess_ragged_launch_paths(D):
    """Up to 32 shapes and 256 frames go out as one launch (geometry per shape in the kernel
    parameters); more shapes, or more frames, fall back to one launch per shape.  Same pixels."""
    rng = np.random.default_rng(5)
    shapes = [(int(rng.integers(20, 90)), int(rng.integers(20, 90))) for _ in range(40)]
    frames = [synth.frames(1, h, w, seed=300 + i)[0] for i, (h, w) in enumerate(shapes)]
    dev = [torch.from_numpy(f).cuda() for f in frames]
    many = D.preprocess_ragged_cuda(dev, target=64).cpu().numpy()  # 40 shapes: per-shape launches
    few = D.preprocess_ragged_cuda(dev[:30], target=64).cpu().numpy()  # 30 shapes: one launch
    assert np.array_equal(many[:30], few)
    for i in (0, 7, 29, 39):
        want, *_ = O.yolo_preprocess(frames[i], target=64)
        assert np.array_equal(many[i : i + 1], want), shapes[i]
    big = D.preprocess_ragged_cuda([dev[i % 5] for i in range(260)], target=64).cpu().numpy()  # > 256 frames
    assert np.array_equal(big[255:260], few[[0, 1, 2, 3, 4]])


def test_crop_ragged(D):
    shapes = [(480, 640), (1080, 1920), (427, 640), (500, 375), (720, 1280), (333, 500)]
    frames = [synth.frames(1, h, w, seed=300 + i)[0] for i, (h, w) in enumerate(shapes)]
    dets, count = synth.detections(len(shapes), seed=34, kmax=8)
    crops, offsets, rois = D.crop_ragged_cuda([torch.from_numpy(f).cuda() for f in frames],
                                              torch.from_numpy(dets).cuda(), torch.from_numpy(count).cuda())
    offsets = offsets.cpu().numpy()
    assert np.array_equal(offsets, np.concatenate([[0], np.cumsum(count)]))
    for i, f in enumerate(frames):
        g = O.letterbox_geometry(f.shape[0], f.shape[1], 640)
        boxes = O.scale_boxes(dets[i, : count[i]], g.scale, (g.pad_w, g.pad_h), f.shape[:2])
        want = np.concatenate([O.mobilenet_preprocess(O.extract_crop(f, b)) for b in boxes], axis=0)
        assert np.array_equal(crops[offsets[i] : offsets[i + 1]].cpu().numpy(), want), shapes[i]
