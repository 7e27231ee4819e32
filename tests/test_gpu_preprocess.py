"""K1 parity (a1, a2): CUDA letterbox / YOLO preprocess through the C-ABI vs the oracle and
the reference-generated golden vectors.  Bit-exact (uint8 pixels AND float32 tensors)."""

import numpy as np
import pytest
import torch

from conftest import sha
from oracle import pipeline as O
from workloads import synth

pytestmark = pytest.mark.gpu

SHAPES = [(480, 640), (1080, 1920), (1920, 1080), (500, 375), (333, 500), (427, 640), (640, 640),
          (100, 50), (1281, 1279), (720, 1280), (1280, 1280), (37, 640), (641, 643), (5, 7), (1, 1)]


@pytest.fixture(scope="module")
def D():
    from shared.processing import device
    return device


def _dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.mark.parametrize("h,w", SHAPES)
def test_preprocess_matches_oracle(D, h, w):
    frames = synth.frames(2, h, w, seed=h * 7 + w)
    got = D.preprocess_batch_cuda(_dev(frames)).cpu().numpy()
    for i in range(2):
        want, *_ = O.yolo_preprocess(frames[i])
        assert np.array_equal(got[i : i + 1], want)


@pytest.mark.parametrize("h,w", SHAPES)
@pytest.mark.parametrize("target", [640, 320])
def test_letterbox_u8_matches_oracle(D, h, w, target):
    if min(int(h * min(target / h, target / w)), int(w * min(target / h, target / w))) < 1:
        pytest.skip("degenerate: cv2 rejects an empty resize")
    frames = synth.frames(1, h, w, seed=h * 7 + w + 1)
    got = D.letterbox_batch_cuda(_dev(frames), target).cpu().numpy()
    want, scale, pad = O.letterbox(frames[0], target)
    assert np.array_equal(got[0], want)
    g = D.letterbox_geometry(h, w, target)
    assert g.scale == scale and g.padding == pad


def test_golden_vectors(D, golden):
    for case in golden.meta["letterbox"]:
        img = synth.frames(1, case["h"], case["w"], seed=case["seed"])
        target = case.get("target", 640)
        boxed = D.letterbox_batch_cuda(_dev(img), target).cpu().numpy()[0]
        assert sha(boxed) == case["u8_sha"], case
        if "tensor_sha" in case:
            t = D.preprocess_batch_cuda(_dev(img)).cpu().numpy()
            assert sha(t) == case["tensor_sha"], case


def test_odd_target_and_padding_colour(D):
    frames = synth.frames(1, 123, 77, seed=5)
    for target in (333, 97):
        got = D.letterbox_batch_cuda(_dev(frames), target, color=(1, 2, 3)).cpu().numpy()[0]
        want, _, _ = O.letterbox(frames[0], target, color=(1, 2, 3))
        assert np.array_equal(got, want)
        t = D.preprocess_batch_cuda(_dev(frames), target=target).cpu().numpy()
        want_t, *_ = O.yolo_preprocess(frames[0], target=target)
        assert np.array_equal(t, want_t)


def test_unaligned_base_pointer(D):
    """Frames that do not start on a 16-byte boundary (bulk copy head/tail path)."""
    h, w = 250, 333
    raw = synth.frames(1, h + 1, w, seed=9).reshape(-1)
    buf = _dev(raw)
    for off in (1, 3, 5, 13):
        view = buf[off : off + h * w * 3].view(1, h, w, 3)
        got = D.preprocess_batch_cuda(view).cpu().numpy()
        want, *_ = O.yolo_preprocess(raw[off : off + h * w * 3].reshape(h, w, 3))
        assert np.array_equal(got, want)


def test_full_size_batch_properties(D):
    """C2 at full size (64 x 640x480): every image equals the oracle on a sample, padding
    rows hold 114/255 and the image band equals frame/255 exactly."""
    frames = synth.frames(64, 480, 640, seed=1)
    out = D.preprocess_batch_cuda(_dev(frames))
    assert out.shape == (64, 3, 640, 640) and out.dtype == torch.float32 and out.is_contiguous()
    pad = np.float32(114) / np.float32(255)
    assert bool((out[:, :, :80, :] == float(pad)).all()) and bool((out[:, :, 560:, :] == float(pad)).all())
    band = out[:, :, 80:560, :].cpu().numpy()
    want = (frames.astype(np.float32) / np.float32(255.0)).transpose(0, 3, 1, 2)
    assert np.array_equal(band, want)


def test_empty_batch(D):
    out = D.preprocess_batch_cuda(torch.empty((0, 480, 640, 3), dtype=torch.uint8, device="cuda"))
    assert out.shape == (0, 3, 640, 640)


def test_two_host_threads_with_different_shapes(D):
    """include/ia_b200.h promises that calls on different streams may run concurrently from
    different host threads.  Two threads preprocess frames of different shapes (different resize
    modes, different shared-memory sizes: the per-kernel shared-memory cap is raised once, never
    re-set per launch) and crop different ROI sets, each on its own stream, many times over."""
    import threading

    cases = [(synth.frames(3, 480, 640, seed=1), None), (synth.frames(3, 1080, 1920, seed=2), None),
             (synth.frames(2, 375, 500, seed=3), None), (synth.frames(2, 2160, 3840, seed=4), None)]
    cases = [(f, np.concatenate([O.yolo_preprocess(x)[0] for x in f])) for f, _ in cases]
    errors = []

    def worker(k):
        try:
            stream = torch.cuda.Stream()
            with torch.cuda.stream(stream):
                for it in range(40):
                    frames, want = cases[(k + it) % len(cases)] if it % 2 else cases[k]
                    got = D.preprocess_batch_cuda(_dev(frames))
                    rois = torch.tensor([[0, 0, 0, frames.shape[2], frames.shape[1], 0, 0, 0]], dtype=torch.int32).cuda()
                    crop = D.resize_normalize_rois_cuda(_dev(frames), rois)
                    stream.synchronize()
                    if not np.array_equal(got.cpu().numpy(), want):
                        errors.append(("K1", k, it))
                    if it == 0 and not np.array_equal(crop.cpu().numpy(), O.mobilenet_preprocess(frames[0])):
                        errors.append(("K4", k, it))
        except Exception as e:  # noqa: BLE001
            errors.append((k, repr(e)))

    threads = [threading.Thread(target=worker, args=(k,)) for k in range(4)]
    [t.start() for t in threads]
    [t.join() for t in threads]
    assert not errors, errors[:5]
