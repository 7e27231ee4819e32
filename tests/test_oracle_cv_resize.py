"""The numpy restatement of OpenCV's 8-bit INTER_LINEAR must equal cv2.resize bit for bit
(third-party arithmetic behind transforms.py:160-164 and mobilenet_preprocess.py:139-143)."""

import numpy as np
import pytest

from oracle.cv_resize import axis_table, is_box2x, resize_linear_u8

cv2 = pytest.importorskip("cv2")

CROP_SHAPES = [(h, w) for h in (1, 2, 3, 17, 100, 223, 224, 225, 447, 448, 449, 720)
               for w in (1, 2, 33, 150, 224, 448, 640, 1279)]
FRAME_SHAPES = [(480, 640), (1080, 1920), (1920, 1080), (500, 375), (333, 500), (427, 640),
                (640, 640), (1280, 1280), (1281, 1279), (100, 50), (720, 1280), (640, 480)]


@pytest.mark.parametrize("h,w", CROP_SHAPES)
def test_crop_to_224_matches_cv2(h, w):
    img = np.random.default_rng(h * 4096 + w).integers(0, 256, (h, w, 3), dtype=np.uint8)
    want = cv2.resize(img, (224, 224), interpolation=cv2.INTER_LINEAR)
    assert np.array_equal(resize_linear_u8(img, 224, 224), want)


@pytest.mark.parametrize("h,w", FRAME_SHAPES)
@pytest.mark.parametrize("target", [320, 640])
def test_letterbox_resize_matches_cv2(h, w, target):
    scale = min(target / h, target / w)
    nw, nh = int(w * scale), int(h * scale)
    img = np.random.default_rng(h * 4096 + w).integers(0, 256, (h, w, 3), dtype=np.uint8)
    want = cv2.resize(img, (nw, nh), interpolation=cv2.INTER_LINEAR)
    assert np.array_equal(resize_linear_u8(img, nw, nh), want)


def test_special_cases():
    assert is_box2x(224, 224, 448, 448)
    assert not is_box2x(224, 224, 448, 449)
    assert not is_box2x(224, 224, 672, 672)
    # exact 3x downscale lands on pixel centres: pure subsample
    x0, a0, a1 = axis_table(640, 1920, True)
    assert np.array_equal(x0, 3 * np.arange(640) + 1) and (a1 == 0).all() and (a0 == 2048).all()


@pytest.mark.parametrize("h,w", [(100, 100), (10, 10), (500, 800), (224, 224), (448, 448), (1, 1), (300, 17), (223, 225)])
def test_float32_resize_matches_cv2_generic_path(h, w):
    """Bit-exact against cv2 with IPP off; within 1e-4 of the IPP-accelerated default (the
    pip wheels' float resize is Intel IPP, whose arithmetic is not OpenCV's published one)."""
    from oracle.cv_resize import resize_linear_f32

    img = np.random.default_rng(h * 4096 + w).random((h, w, 3)).astype(np.float32)
    got = resize_linear_f32(img, 224, 224)
    with_ipp = cv2.resize(img, (224, 224), interpolation=cv2.INTER_LINEAR)
    was = cv2.ipp.useIPP()
    try:
        cv2.ipp.setUseIPP(False)
        generic = cv2.resize(img, (224, 224), interpolation=cv2.INTER_LINEAR)
    finally:
        cv2.ipp.setUseIPP(was)
    assert np.array_equal(got, generic)
    assert np.allclose(got, with_ipp, rtol=0, atol=1e-4)
