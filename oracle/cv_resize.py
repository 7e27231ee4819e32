"""numpy restatement of OpenCV's ``cv2.resize(..., INTER_LINEAR)`` for CV_8UC3.

TEST INFRASTRUCTURE (see ``oracle/__init__.py``).

The reference does its resize arithmetic inside a third-party dependency that
is not vendored under ``/root/reference``: **opencv-python** (pinned
4.11.0.86 in the reference's ``uv.lock:2549-2550``; 4.13.0 in this image).
Reference call sites: ``src/shared/processing/transforms.py:160-164``
(letterbox) and ``src/shared/processing/mobilenet_preprocess.py:139-143``
(crop -> 224x224).  This file restates the published algorithm of OpenCV's
``modules/imgproc/src/resize.cpp`` 8-bit linear path
(``HResizeLinear<uchar,int,short,2048>`` + ``VResizeLinear<uchar,int,short,
FixedPtCast<int,uchar,22>>``):

* ``inv = (double)dst / src``; ``sc = 1.0 / inv``; per destination index ``i``
  ``f = (float)((i + 0.5) * sc - 0.5)``; ``s0 = floor(f)``; ``f -= s0``.
* X axis only: ``s0 < 0 -> (s0, f) = (0, 0)``; ``s0 >= S-1 -> (S-1, 0)``.
* Y axis: the fraction is kept; the two rows are ``clamp(s0)``, ``clamp(s0+1)``.
* 11-bit coefficients ``a1 = rint(f * 2048)``, ``a0 = rint((1 - f) * 2048)``
  (float32 maths, round-half-even, stored as int16).
* horizontal pass ``h = p[x0] * a0 + p[x1] * a1`` (int32);
* vertical pass ``((b0 * (h0 >> 4)) >> 16) + ((b1 * (h1 >> 4)) >> 16) + 2) >> 2``.
* ``dst size == src size`` is a plain copy; an exact 2x downscale on both axes
  is silently turned into a 2x2 box average ``(p00 + p01 + p10 + p11 + 2) >> 2``.

It is pinned against ``cv2.resize`` itself in ``tests/test_oracle_cv_resize.py``
(runs wherever cv2 is importable) and against reference-generated golden
vectors in ``tests/golden/``.
"""

from __future__ import annotations

import numpy as np

COEF_BITS = 11
COEF_ONE = 1 << COEF_BITS  # 2048


def axis_table(dst: int, src: int, zero_frac_at_border: bool):
    """Return ``(s0, a0, a1)`` for one axis; ``s0`` is unclamped on the Y axis."""
    inv = float(dst) / float(src)
    sc = 1.0 / inv
    i = np.arange(dst, dtype=np.float64)
    f = ((i + 0.5) * sc - 0.5).astype(np.float32)
    s0 = np.floor(f).astype(np.int32)
    f = (f - s0.astype(np.float32)).astype(np.float32)
    if zero_frac_at_border:
        lo = s0 < 0
        s0 = np.where(lo, 0, s0)
        f = np.where(lo, np.float32(0), f)
        hi = s0 >= src - 1
        s0 = np.where(hi, src - 1, s0)
        f = np.where(hi, np.float32(0), f)
    one = np.float32(1.0)
    scale = np.float32(COEF_ONE)
    a1 = np.rint(f.astype(np.float32) * scale).astype(np.int32)
    a0 = np.rint((one - f.astype(np.float32)) * scale).astype(np.int32)
    return s0.astype(np.int32), a0, a1


def is_box2x(dst_h: int, dst_w: int, src_h: int, src_w: int) -> bool:
    """OpenCV's ``is_area_fast && iscale == 2`` test (resize.cpp), in doubles."""
    sx = 1.0 / (float(dst_w) / float(src_w))
    sy = 1.0 / (float(dst_h) / float(src_h))
    eps = np.finfo(np.float64).eps
    return (
        int(np.rint(sx)) == 2
        and int(np.rint(sy)) == 2
        and abs(sx - 2) < eps
        and abs(sy - 2) < eps
    )


def resize_linear_u8(img: np.ndarray, dst_w: int, dst_h: int) -> np.ndarray:
    """``cv2.resize(img, (dst_w, dst_h), interpolation=cv2.INTER_LINEAR)`` for uint8 HWC."""
    assert img.dtype == np.uint8 and img.ndim == 3
    src_h, src_w = img.shape[:2]
    if dst_w <= 0 or dst_h <= 0 or src_w <= 0 or src_h <= 0:
        raise ValueError("empty size in resize")
    if (dst_h, dst_w) == (src_h, src_w):
        return img.copy()
    if is_box2x(dst_h, dst_w, src_h, src_w):
        p = img.astype(np.int32)
        s = p[0::2, 0::2] + p[0::2, 1::2] + p[1::2, 0::2] + p[1::2, 1::2]
        return ((s[:dst_h, :dst_w] + 2) >> 2).astype(np.uint8)

    x0, a0, a1 = axis_table(dst_w, src_w, True)
    y0, b0, b1 = axis_table(dst_h, src_h, False)
    x1 = np.minimum(x0 + 1, src_w - 1)
    r0 = np.clip(y0, 0, src_h - 1)
    r1 = np.clip(y0 + 1, 0, src_h - 1)

    p = img.astype(np.int32)
    # horizontal pass on every source row that is used
    rows = np.unique(np.concatenate([r0, r1]))
    hrow = np.empty((src_h, dst_w, img.shape[2]), dtype=np.int32)
    hrow[rows] = p[rows][:, x0] * a0[None, :, None] + p[rows][:, x1] * a1[None, :, None]
    h0 = hrow[r0] >> 4
    h1 = hrow[r1] >> 4
    out = (((b0[:, None, None] * h0) >> 16) + ((b1[:, None, None] * h1) >> 16) + 2) >> 2
    return np.clip(out, 0, 255).astype(np.uint8)


def resize_linear_f32(img: np.ndarray, dst_w: int, dst_h: int) -> np.ndarray:
    """``cv2.resize(img, (dst_w, dst_h), INTER_LINEAR)`` for float32 HWC — OpenCV's GENERIC
    float path (resize.cpp ``HResizeLinear<float>`` / ``VResizeLinear<float>``): the same
    coordinate maths as the 8-bit path, float coefficients ``1 - f`` and ``f``, every product
    and every sum rounded to float32 on its own (no FMA), horizontal pass first.  An exact
    2x downscale is the box average ``(((p00 + p01) + p10) + p11) * 0.25``.

    Bit-identical to ``cv2.resize`` with ``cv2.ipp.setUseIPP(False)``.  The pip builds of
    OpenCV route float resizes through Intel IPP, whose (unpublished) arithmetic differs from
    this by up to ~3e-5 on [0, 1] data — see tests/test_oracle_cv_resize.py."""
    assert img.dtype == np.float32 and img.ndim == 3
    src_h, src_w = img.shape[:2]
    if (dst_h, dst_w) == (src_h, src_w):
        return img.copy()
    if is_box2x(dst_h, dst_w, src_h, src_w):
        s = img[0::2, 0::2] + img[0::2, 1::2]
        s = s + img[1::2, 0::2]
        s = s + img[1::2, 1::2]
        return (s[:dst_h, :dst_w] * np.float32(0.25)).astype(np.float32)

    def table(dst, src, zero_frac_at_border):
        sc = 1.0 / (float(dst) / float(src))
        i = np.arange(dst, dtype=np.float64)
        f = ((i + 0.5) * sc - 0.5).astype(np.float32)
        s0 = np.floor(f).astype(np.int32)
        f = (f - s0.astype(np.float32)).astype(np.float32)
        if zero_frac_at_border:
            lo = s0 < 0
            s0, f = np.where(lo, 0, s0), np.where(lo, np.float32(0), f)
            hi = s0 >= src - 1
            s0, f = np.where(hi, src - 1, s0), np.where(hi, np.float32(0), f)
        return s0, (np.float32(1) - f).astype(np.float32), f.astype(np.float32)

    x0, a0, a1 = table(dst_w, src_w, True)
    y0, b0, b1 = table(dst_h, src_h, False)
    x1 = np.minimum(x0 + 1, src_w - 1)
    r0, r1 = np.clip(y0, 0, src_h - 1), np.clip(y0 + 1, 0, src_h - 1)
    h = (img[:, x0] * a0[None, :, None]).astype(np.float32) + (img[:, x1] * a1[None, :, None]).astype(np.float32)
    out = (h[r0] * b0[:, None, None]).astype(np.float32) + (h[r1] * b1[:, None, None]).astype(np.float32)
    return out.astype(np.float32)


def resize_linear_f32_ipp(img: np.ndarray, dst_w: int, dst_h: int) -> np.ndarray:
    """``cv2.resize(img, (dst_w, dst_h), INTER_LINEAR)`` for float32 HWC as the opencv-python
    WHEELS compute it (the reference pins opencv-python 4.11.0.86, ``uv.lock:2549-2550``): those
    builds hand float resizes to Intel IPP (``ipp_resize`` in resize.cpp, ``ippiResizeLinear_32f``).
    IPP's arithmetic is not published; the following was found by differential testing against
    cv2 4.13.0 (IPP 2022.2) and is bit-identical to it on every shape tried
    (tests/test_oracle_cv_resize.py):

    * coordinates in double, ``pos = (i + 0.5) * (src / dst) - 0.5``; the tap is ``floor(pos)``
      and the weight ``float32(pos - floor(pos))`` — the fraction is NOT rounded to float32
      before the floor is subtracted (OpenCV's own tables do that);
    * both taps clamped to the image (replicated border);
    * each interpolation is ``fma(p1 - p0, w, p0)`` (difference rounded to float32, then one
      fused multiply-add), horizontal pass first, then vertical;
    * an exact 2x downscale is interpolated like any other size (no box average);
    * equal sizes are a copy; an image with a side of one pixel is refused by IPP and takes
      OpenCV's generic path (``resize_linear_f32``)."""
    assert img.dtype == np.float32 and img.ndim == 3
    src_h, src_w = img.shape[:2]
    if (dst_h, dst_w) == (src_h, src_w):
        return img.copy()
    if src_h < 2 or src_w < 2:
        return resize_linear_f32(img, dst_w, dst_h)

    def fma(a, b, c):  # exact: the product of two float32 is exact in float64
        return (a.astype(np.float64) * b.astype(np.float64) + c.astype(np.float64)).astype(np.float32)

    def table(dst, src):
        pos = (np.arange(dst, dtype=np.float64) + 0.5) * (float(src) / float(dst)) - 0.5
        s0 = np.floor(pos)
        w = (pos - s0).astype(np.float32)
        s0 = s0.astype(np.int64)
        return np.clip(s0, 0, src - 1), np.clip(s0 + 1, 0, src - 1), w

    x0, x1, wx = table(dst_w, src_w)
    y0, y1, wy = table(dst_h, src_h)
    p0, p1 = img[:, x0], img[:, x1]
    h = fma((p1 - p0).astype(np.float32), np.broadcast_to(wx[None, :, None], p0.shape), p0)
    q0, q1 = h[y0], h[y1]
    return fma((q1 - q0).astype(np.float32), np.broadcast_to(wy[:, None, None], q0.shape), q0)
