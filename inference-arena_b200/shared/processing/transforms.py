"""Atomic image transforms — same names and signatures as the reference's
``src/shared/processing/transforms.py``; arithmetic on the GPU through libia_b200.so.

    letterbox            transforms.py:118-180   -> ia_letterbox_u8
    scale_boxes          transforms.py:183-230   -> ia_scale_boxes
    imagenet_normalize   transforms.py:238-272   -> ia_imagenet_normalize_u8
    load_image(_from_bytes) transforms.py:45-110 -> baseline JPEGs: shared.processing.jpeg
                         (device decoder); everything else, and damaged files:
                         cv2 on the host, as in the reference
"""

from __future__ import annotations

import numpy as np

from shared.config import get_controlled_variable
from shared.processing import _hostio, device as _dev

_mobilenet_config = get_controlled_variable("preprocessing", "mobilenet")
IMAGENET_MEAN: np.ndarray = np.array(_mobilenet_config["mean"], dtype=np.float32)
IMAGENET_STD: np.ndarray = np.array(_mobilenet_config["std"], dtype=np.float32)
LETTERBOX_COLOR: tuple[int, int, int] = (114, 114, 114)


def _decode_jpeg_on_device(data: bytes) -> tuple[np.ndarray | None, bool]:
    """(RGB pixels, False) for a baseline JPEG decoded on the GPU.  (None, False): the host decoder
    has to take the file — not a JPEG the device decoder handles, or a damaged one (libjpeg's
    concealment is cv2's to apply).  (None, True): the data ends before the last block, which is
    where ``cv2.imdecode`` gives up and returns None."""
    if data[:2] != b"\xff\xd8":
        return None, False
    from shared.processing import jpeg as _jpeg
    try:
        batch = _jpeg.decode_jpeg_batch_cuda([data])
    except _jpeg.UnsupportedJpeg:
        return None, False
    status = int(batch.status[0])  # synchronises
    if status == 0:
        return _hostio.to_host(batch.frames[0]), False
    return None, bool(status & 2)


def load_image(image_path: str) -> np.ndarray:
    """Decode an image file to RGB uint8 [H, W, 3] (reference transforms.py:45-74): an intact
    baseline JPEG on the GPU like ``load_image_from_bytes``, everything else — and every path
    that cannot be read — through ``cv2.imread`` as in the reference."""
    try:
        with open(image_path, "rb") as fh:
            data = fh.read()
    except (OSError, TypeError, ValueError):
        data = b""
    pixels, _ = _decode_jpeg_on_device(data)
    if pixels is not None:
        return pixels
    import cv2

    bgr = cv2.imread(image_path, cv2.IMREAD_COLOR)
    if bgr is None:
        raise ValueError(f"Failed to load image: {image_path}")
    return cv2.cvtColor(bgr, cv2.COLOR_BGR2RGB)


def load_image_from_bytes(image_bytes: bytes) -> np.ndarray:
    """Decode encoded image bytes to RGB uint8 [H, W, 3] (reference transforms.py:77-110).

    A baseline JPEG — what the services receive (monolithic/app/inference.py:153) — is decoded on
    the GPU (``shared.processing.jpeg``, bit-identical to ``cv2.imdecode``): only the compressed
    bytes cross PCIe.  Everything the device decoder does not take (PNG and other formats,
    progressive / arithmetic / CMYK JPEGs) and any file it reports as damaged goes through
    ``cv2.imdecode`` on the host exactly as in the reference, so the pixels and the two
    ``ValueError`` messages are the reference's for every input."""
    if not image_bytes:
        raise ValueError("Failed to decode image from bytes: empty input")
    data = bytes(image_bytes)
    pixels, truncated = _decode_jpeg_on_device(data)
    if pixels is not None:
        return pixels
    if truncated:
        raise ValueError("Failed to decode image from bytes")
    import cv2

    bgr = cv2.imdecode(np.frombuffer(data, np.uint8), cv2.IMREAD_COLOR)
    if bgr is None:
        raise ValueError("Failed to decode image from bytes")
    return cv2.cvtColor(bgr, cv2.COLOR_BGR2RGB)


def letterbox(image: np.ndarray, target_size: int,
              color: tuple[int, int, int] = LETTERBOX_COLOR) -> tuple[np.ndarray, float, tuple[int, int]]:
    """Aspect-preserving resize into a ``target_size`` square padded with ``color``.
    Returns ``(uint8[T,T,3], scale, (pad_w, pad_h))``."""
    height, width = image.shape[:2]
    geom = _dev.letterbox_geometry(height, width, target_size)
    frames = _hostio.to_device(image.astype(np.uint8, copy=False))[None]
    boxed = _dev.letterbox_batch_cuda(frames, target_size, color)
    return _hostio.to_host(boxed[0]), geom.scale, geom.padding


def scale_boxes(boxes: np.ndarray, scale: float, padding: tuple[int, int],
                original_shape: tuple[int, int]) -> np.ndarray:
    """Map ``[N, 4+]`` boxes from letterbox space back to the original frame (input is not
    modified; columns past the fourth are passed through)."""
    if boxes.dtype not in (np.float32, np.float64):
        raise TypeError(f"scale_boxes: float32 or float64 boxes expected, got {boxes.dtype}")
    if boxes.shape[0] == 0:
        return boxes.copy()
    out = _dev.scale_boxes_cuda(_hostio.to_device(boxes), scale, padding, original_shape)
    return _hostio.to_host(out)


def imagenet_normalize(image: np.ndarray) -> np.ndarray:
    """``(pixel / 255 - mean) / std`` per channel, float32 HWC out."""
    if image.dtype == np.uint8:
        out = _dev.imagenet_normalize_cuda(_hostio.to_device(image), IMAGENET_MEAN, IMAGENET_STD)
    else:  # reference: image.astype(np.float32), rescaled by 1/255 only when its maximum exceeds 1
        out = _dev.imagenet_normalize_f32_cuda(_hostio.to_device(image.astype(np.float32)), IMAGENET_MEAN,
                                               IMAGENET_STD)
    return _hostio.to_host(out)
