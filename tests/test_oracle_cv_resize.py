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


F32_SHAPES = [(100, 100), (10, 10), (500, 800), (224, 224), (448, 448), (448, 224), (1, 1), (1, 300), (300, 1),
              (2, 2), (3, 3), (300, 17), (17, 300), (223, 225), (449, 447), (672, 672), (896, 896), (64, 1000),
              (31, 33), (57, 91), (225, 225), (2, 500), (112, 112), (1080, 1920)]


@pytest.mark.parametrize("h,w", F32_SHAPES)
def test_float32_resize_matches_cv2_with_ipp(h, w):
    """The opencv-python wheels (what the reference pins and what this image has) hand float32
    resizes to Intel IPP.  oracle.cv_resize.resize_linear_f32_ipp reproduces that result bit for bit,
    on [0,1] and on [0,255] data."""
    from oracle.cv_resize import resize_linear_f32_ipp

    if not cv2.ipp.useIPP():
        pytest.skip("this cv2 build has no IPP")
    rng = np.random.default_rng(h * 4096 + w)
    for top in (1.0, 255.0):
        img = (rng.random((h, w, 3)) * top).astype(np.float32)
        want = cv2.resize(img, (224, 224), interpolation=cv2.INTER_LINEAR)
        assert np.array_equal(resize_linear_f32_ipp(img, 224, 224), want)


def test_float32_resize_ipp_random_shapes():
    from oracle.cv_resize import resize_linear_f32_ipp

    if not cv2.ipp.useIPP():
        pytest.skip("this cv2 build has no IPP")
    rng = np.random.default_rng(77)
    for _ in range(60):
        h, w = (int(v) for v in rng.integers(1, 700, 2))
        img = rng.random((h, w, 3)).astype(np.float32)
        want = cv2.resize(img, (224, 224), interpolation=cv2.INTER_LINEAR)
        assert np.array_equal(resize_linear_f32_ipp(img, 224, 224), want), (h, w)
