"""Seeded JPEG test inputs (host side): photo-like synthetic images and the encoder settings
the parity tests sweep.  Encoding uses cv2 (libjpeg-turbo), the encoder the reference's data
came through; the decoders under test never see these helpers."""

from __future__ import annotations

import struct

import numpy as np

SAMPLINGS = ("444", "420", "422", "440", "411")


def photo(h: int, w: int, seed: int) -> np.ndarray:
    """Smooth gradients + texture + flat rectangles: exercises DC chains, EOBs, long AC runs."""
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:h, 0:w]
    img = np.stack([128 + 100 * np.sin(xx / (7 + seed % 5) + yy / 13.0),
                    128 + 90 * np.cos(yy / (5 + seed % 3) - xx / 17.0),
                    (xx * 3 + yy * 2) % 256], axis=-1).astype(np.float64)
    img += rng.normal(0, 12, img.shape)
    for _ in range(4):
        x0, y0 = int(rng.integers(0, max(w - 1, 1))), int(rng.integers(0, max(h - 1, 1)))
        img[y0:y0 + int(rng.integers(1, h // 2 + 2)), x0:x0 + int(rng.integers(1, w // 2 + 2))] = \
            rng.integers(0, 256, 3)
    return np.clip(img, 0, 255).astype(np.uint8)


def encode(img: np.ndarray, quality: int = 90, sampling: str = "420", restart: int = 0,
           optimize: bool = False, progressive: bool = False) -> bytes:
    import cv2

    flags = {"444": cv2.IMWRITE_JPEG_SAMPLING_FACTOR_444, "420": cv2.IMWRITE_JPEG_SAMPLING_FACTOR_420,
             "422": cv2.IMWRITE_JPEG_SAMPLING_FACTOR_422, "440": cv2.IMWRITE_JPEG_SAMPLING_FACTOR_440,
             "411": cv2.IMWRITE_JPEG_SAMPLING_FACTOR_411}
    params = [cv2.IMWRITE_JPEG_QUALITY, int(quality)]
    if img.ndim == 3:
        params += [cv2.IMWRITE_JPEG_SAMPLING_FACTOR, flags[sampling]]
    if restart:
        params += [cv2.IMWRITE_JPEG_RST_INTERVAL, int(restart)]
    if optimize:
        params += [cv2.IMWRITE_JPEG_OPTIMIZE, 1]
    if progressive:
        params += [cv2.IMWRITE_JPEG_PROGRESSIVE, 1]
    src = img[..., ::-1] if img.ndim == 3 else img  # cv2 encodes BGR
    ok, buf = cv2.imencode(".jpg", np.ascontiguousarray(src), params)
    assert ok
    return buf.tobytes()


def with_exif_orientation(jpeg: bytes, orientation: int, big_endian: bool = False) -> bytes:
    """Insert an APP1 Exif segment carrying only the orientation tag right after SOI."""
    e = ">" if big_endian else "<"
    tiff = (b"MM" if big_endian else b"II") + struct.pack(e + "HI", 42, 8)
    tiff += struct.pack(e + "H", 1) + struct.pack(e + "HHIHH", 0x0112, 3, 1, orientation, 0) + struct.pack(e + "I", 0)
    payload = b"Exif\x00\x00" + tiff
    return jpeg[:2] + b"\xff\xe1" + struct.pack(">H", len(payload) + 2) + payload + jpeg[2:]


def reference_decode(jpeg: bytes) -> np.ndarray:
    """cv2.imdecode + BGR2RGB: the two calls of load_image_from_bytes (transforms.py:102-108)."""
    import cv2

    bgr = cv2.imdecode(np.frombuffer(jpeg, np.uint8), cv2.IMREAD_COLOR)
    if bgr is None:
        raise ValueError("Failed to decode image from bytes")
    return cv2.cvtColor(bgr, cv2.COLOR_BGR2RGB)


def with_16bit_quant_tables(jpeg: bytes) -> bytes:
    """Rewrite every 8-bit DQT table of the header as a 16-bit one (Pq = 1) with the same values:
    legal, rarely produced by encoders, and a separate branch in every decoder."""
    out = bytearray(jpeg[:2])
    i = 2
    while True:
        assert jpeg[i] == 0xFF
        m = jpeg[i + 1]
        length = (jpeg[i + 2] << 8) | jpeg[i + 3]
        seg = jpeg[i + 4:i + 2 + length]
        if m == 0xDB:
            body = bytearray()
            k = 0
            while k < len(seg):
                pq, tq = seg[k] >> 4, seg[k] & 15
                k += 1
                if pq:
                    body += bytes([seg[k - 1]]) + seg[k:k + 128]
                    k += 128
                else:
                    body.append(0x10 | tq)
                    for v in seg[k:k + 64]:
                        body += bytes([0, v])
                    k += 64
            out += b"\xff\xdb" + struct.pack(">H", len(body) + 2) + body
        else:
            out += jpeg[i:i + 2 + length]
        i += 2 + length
        if m == 0xDA:
            break
    return bytes(out) + jpeg[i:]
