"""CPU restatement of what ``load_image_from_bytes`` does to a baseline JPEG.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

The reference calls ``cv2.imdecode(buf, IMREAD_COLOR)`` + ``cvtColor(BGR2RGB)``
(src/shared/processing/transforms.py:77-110).  The arithmetic lives in a third-party
dependency that is not under /root/reference: OpenCV's bundled libjpeg-turbo (opencv-python
4.13.0 here bundles 3.1.2; the reference pins opencv-python 4.11.0.86).  This module restates
libjpeg(-turbo)'s published decoder with its default settings, the ones OpenCV uses:

  * sequential Huffman decoding, ITU T.81 Annex F.2.2 (jdhuff.c);
  * dequantisation + the "slow-but-accurate" integer inverse DCT, jidctint.c
    (CONST_BITS 13, PASS1_BITS 2, Loeffler-Ligtenberg-Moschytz), JDCT_ISLOW;
  * "fancy" (triangle-filter) chroma upsampling, jdsample.c: h2v1, h2v2, h1v2; plain
    replication when the subsampled width is <= 2 or for other integral factors;
  * YCbCr -> RGB with the 16-bit fixed-point tables of jdcolor.c; grayscale replicated;
  * the EXIF orientation OpenCV applies after decoding (modules/imgcodecs/src/exif.cpp,
    loadsave.cpp ApplyExifOrientation).

It is pinned by tests/test_oracle_jpeg.py against ``cv2.imdecode`` itself (every sampling
mode, odd sizes, restart intervals, optimised tables, qualities 5..100) and against the
committed fixtures produced through the reference's own ``load_image_from_bytes``
(tests/golden/make_golden.py).  Pure-Python loops: meant for small images.
"""

from __future__ import annotations

import numpy as np

ZIGZAG = np.array([0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34,
                   27, 20, 13, 6, 7, 14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44,
                   51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63], dtype=np.int64)


class UnsupportedJpeg(ValueError):
    """Valid JPEG, but a flavour the GPU decoder does not take (progressive, arithmetic,
    12-bit, CMYK, multi-scan)."""


# ------------------------------------------------------------------------------ parsing
def _u16(b: bytes, i: int) -> int:
    return (b[i] << 8) | b[i + 1]


def _exif_orientation(seg: bytes) -> int:
    """Orientation tag 0x0112 of an APP1 'Exif' segment payload (after the length); 1 if absent."""
    if len(seg) < 14 or seg[:6] != b"Exif\x00\x00":
        return 1
    t = seg[6:]
    if t[:2] == b"II":
        rd16 = lambda o: t[o] | (t[o + 1] << 8)  # noqa: E731
        rd32 = lambda o: t[o] | (t[o + 1] << 8) | (t[o + 2] << 16) | (t[o + 3] << 24)  # noqa: E731
    elif t[:2] == b"MM":
        rd16 = lambda o: (t[o] << 8) | t[o + 1]  # noqa: E731
        rd32 = lambda o: (t[o] << 24) | (t[o + 1] << 16) | (t[o + 2] << 8) | t[o + 3]  # noqa: E731
    else:
        return 1
    if len(t) < 8 or rd16(2) != 42:
        return 1
    ifd = rd32(4)
    if ifd + 2 > len(t):
        return 1
    n = rd16(ifd)
    for e in range(n):
        o = ifd + 2 + 12 * e
        if o + 12 > len(t):
            break
        if rd16(o) == 0x0112:
            v = rd16(o + 8)
            return v if 1 <= v <= 8 else 1
    return 1


def parse(data: bytes) -> dict:
    """Headers of a baseline / extended-sequential Huffman JPEG with one interleaved scan."""
    if len(data) < 4 or data[0] != 0xFF or data[1] != 0xD8:
        raise ValueError("not a JPEG stream")
    info = {"qt": {}, "dc": {}, "ac": {}, "restart": 0, "orientation": 1, "adobe_transform": None}
    i = 2
    saw_exif = False
    while True:
        while i < len(data) and data[i] != 0xFF:
            i += 1
        while i < len(data) and data[i] == 0xFF:
            i += 1
        if i >= len(data):
            raise ValueError("truncated JPEG (no SOS)")
        m = data[i]
        i += 1
        if m == 0xD8 or 0xD0 <= m <= 0xD7 or m == 0x01:
            continue
        if m == 0xD9:
            raise ValueError("EOI before SOS")
        L = _u16(data, i)
        seg = data[i + 2:i + L]
        if m == 0xDB:  # DQT
            k = 0
            while k < len(seg):
                pq, tq = seg[k] >> 4, seg[k] & 15
                k += 1
                if pq:
                    vals = [(seg[k + 2 * j] << 8) | seg[k + 2 * j + 1] for j in range(64)]
                    k += 128
                else:
                    vals = list(seg[k:k + 64])
                    k += 64
                q = np.zeros(64, dtype=np.int64)
                q[ZIGZAG] = vals  # stored in zigzag order -> natural order
                info["qt"][tq] = q
        elif m in (0xC0, 0xC1):  # SOF0 / SOF1 (Huffman, sequential)
            if seg[0] != 8:
                raise UnsupportedJpeg("sample precision != 8")
            info["height"], info["width"] = _u16(seg, 1), _u16(seg, 3)
            nc = seg[5]
            info["comps"] = [{"id": seg[6 + 3 * c], "h": seg[7 + 3 * c] >> 4, "v": seg[7 + 3 * c] & 15,
                              "tq": seg[8 + 3 * c]} for c in range(nc)]
        elif m in (0xC2, 0xC3, 0xC5, 0xC6, 0xC7, 0xC9, 0xCA, 0xCB, 0xCD, 0xCE, 0xCF):
            raise UnsupportedJpeg(f"SOF marker 0x{m:02X} (progressive / lossless / arithmetic)")
        elif m == 0xC4:  # DHT
            k = 0
            while k < len(seg):
                tc, th = seg[k] >> 4, seg[k] & 15
                counts = list(seg[k + 1:k + 17])
                n = sum(counts)
                vals = list(seg[k + 17:k + 17 + n])
                k += 17 + n
                info["ac" if tc else "dc"][th] = (counts, vals)
        elif m == 0xDD:
            info["restart"] = _u16(seg, 0)
        elif m == 0xE1 and not saw_exif and seg[:6] == b"Exif\x00\x00":
            saw_exif = True
            info["orientation"] = _exif_orientation(seg)
        elif m == 0xEE and seg[:5] == b"Adobe" and len(seg) >= 12:
            info["adobe_transform"] = seg[11]
        elif m == 0xE0 and seg[:5] == b"JFIF\x00":
            info["jfif"] = True
        elif m == 0xDA:  # SOS
            ns = seg[0]
            if "comps" not in info:
                raise ValueError("SOS before SOF")
            if ns != len(info["comps"]):
                raise UnsupportedJpeg("non-interleaved (multi-scan) JPEG")
            for s in range(ns):
                cid, tabs = seg[1 + 2 * s], seg[2 + 2 * s]
                comp = info["comps"][s]
                if comp["id"] != cid:
                    raise UnsupportedJpeg("scan component order differs from frame order")
                comp["td"], comp["ta"] = tabs >> 4, tabs & 15
            info["scan_start"] = i + L
            break
        i += L
    nc = len(info["comps"])
    if nc not in (1, 3):
        raise UnsupportedJpeg(f"{nc} components")
    # libjpeg's colour-space guess (jdapimin.c default_decompress_parms): JFIF -> YCbCr; else the
    # Adobe marker decides; else component ids 'R','G','B' -> RGB
    if nc == 3 and not info.get("jfif"):
        ids_rgb = [c["id"] for c in info["comps"]] == [ord("R"), ord("G"), ord("B")]
        if info["adobe_transform"] == 0 or (info["adobe_transform"] is None and ids_rgb):
            raise UnsupportedJpeg("stored as RGB (no YCbCr transform)")
    if nc == 1:
        info["comps"][0]["h"] = info["comps"][0]["v"] = 1  # sampling factors mean nothing here
    if info["width"] == 0 or info["height"] == 0:
        raise ValueError("empty image")
    return info


# ------------------------------------------------------------------------------ Huffman
class _Huff:
    """Decoding tables of T.81 F.2.2.3 (mincode / maxcode / valptr)."""

    def __init__(self, counts, vals):
        self.vals = vals
        self.mincode, self.maxcode, self.valptr = [0] * 17, [-1] * 17, [0] * 17
        code = k = 0
        for length in range(1, 17):
            self.valptr[length] = k
            self.mincode[length] = code
            code += counts[length - 1]
            k += counts[length - 1]
            self.maxcode[length] = code - 1 if counts[length - 1] else -1
            code <<= 1


class _Bits:
    def __init__(self, data: bytes, pos: int):
        self.d, self.p, self.acc, self.n = data, pos, 0, 0

    def _fill(self):
        d = self.d
        if self.p >= len(d):
            b = 0  # past the end: libjpeg feeds zeros after warning
        else:
            b = d[self.p]
            if b == 0xFF:
                nxt = d[self.p + 1] if self.p + 1 < len(d) else 0xD9
                if nxt == 0:
                    self.p += 2
                else:
                    b = 0  # a marker: do not consume, feed zeros
                    self.acc = (self.acc << 8)
                    self.n += 8
                    return
            else:
                self.p += 1
        self.acc = (self.acc << 8) | b
        self.n += 8

    def bit(self) -> int:
        if self.n == 0:
            self._fill()
        self.n -= 1
        return (self.acc >> self.n) & 1

    def bits(self, k: int) -> int:
        while self.n < k:
            self._fill()
        self.n -= k
        return (self.acc >> self.n) & ((1 << k) - 1)

    def restart(self):
        """Drop the partial byte and consume the RSTn marker."""
        self.acc = self.n = 0
        d = self.d
        while self.p + 1 < len(d) and not (d[self.p] == 0xFF and 0xD0 <= d[self.p + 1] <= 0xD7):
            self.p += 1
        self.p += 2


def _decode_symbol(br: _Bits, h: _Huff) -> int:
    code = 0
    for length in range(1, 17):
        code = (code << 1) | br.bit()
        if h.maxcode[length] >= 0 and code <= h.maxcode[length] and code >= h.mincode[length]:
            return h.vals[h.valptr[length] + code - h.mincode[length]]
    return 0  # invalid code: libjpeg warns and returns 0


def _extend(v: int, s: int) -> int:
    return v - ((1 << s) - 1) if s and v < (1 << (s - 1)) else v


def decode_coefficients(data: bytes, info: dict):
    """Entropy decode -> per component int array [blocks_y, blocks_x, 64] (natural order,
    not dequantised).  blocks_* are padded to whole MCUs."""
    comps = info["comps"]
    hmax, vmax = max(c["h"] for c in comps), max(c["v"] for c in comps)
    mcux = -(-info["width"] // (8 * hmax))
    mcuy = -(-info["height"] // (8 * vmax))
    coefs = [np.zeros((mcuy * c["v"], mcux * c["h"], 64), dtype=np.int64) for c in comps]
    dct = {k: _Huff(*v) for k, v in info["dc"].items()}
    act = {k: _Huff(*v) for k, v in info["ac"].items()}
    br = _Bits(data, info["scan_start"])
    pred = [0] * len(comps)
    ri = info["restart"]
    for m in range(mcux * mcuy):
        if ri and m and m % ri == 0:
            br.restart()
            pred = [0] * len(comps)
        my, mx = divmod(m, mcux)
        for ci, c in enumerate(comps):
            for by in range(c["v"]):
                for bx in range(c["h"]):
                    blk = coefs[ci][my * c["v"] + by, mx * c["h"] + bx]
                    s = _decode_symbol(br, dct[c["td"]])
                    diff = _extend(br.bits(s), s) if s else 0
                    pred[ci] += diff
                    blk[0] = pred[ci]
                    k = 1
                    while k < 64:
                        rs = _decode_symbol(br, act[c["ta"]])
                        r, s = rs >> 4, rs & 15
                        if s == 0:
                            if r != 15:
                                break
                            k += 16
                            continue
                        k += r
                        v = _extend(br.bits(s), s)
                        if k < 64:
                            blk[ZIGZAG[k]] = v
                        k += 1
    return coefs


# ------------------------------------------------------------------------------- IDCT
def _range_limit(x: np.ndarray) -> np.ndarray:
    """libjpeg's ``range_limit[x & RANGE_MASK]`` (table centred on 128): x in [-128, 127] ->
    x + 128, saturating outside, wrapping modulo 1024 like the table lookup."""
    idx = (x.astype(np.int64)) & 1023
    out = np.where(idx < 128, idx + 128, np.where(idx < 512, 255, np.where(idx < 896, 0, idx - 896)))
    return out.astype(np.uint8)


F = dict(f0298=2446, f0390=3196, f0541=4433, f0765=6270, f0899=7373, f1175=9633, f1501=12299, f1847=15137,
         f1961=16069, f2053=16819, f2562=20995, f3072=25172)


def _idct_1d(d0, d1, d2, d3, d4, d5, d6, d7):
    """One LL&M pass on int64 arrays; returns the 8 un-descaled outputs (even part carries
    << CONST_BITS)."""
    z1 = (d2 + d6) * F["f0541"]
    tmp2 = z1 + d6 * (-F["f1847"])
    tmp3 = z1 + d2 * F["f0765"]
    tmp0 = (d0 + d4) << 13
    tmp1 = (d0 - d4) << 13
    tmp10, tmp13, tmp11, tmp12 = tmp0 + tmp3, tmp0 - tmp3, tmp1 + tmp2, tmp1 - tmp2
    t0, t1, t2, t3 = d7, d5, d3, d1
    z1, z2, z3, z4 = t0 + t3, t1 + t2, t0 + t2, t1 + t3
    z5 = (z3 + z4) * F["f1175"]
    t0, t1, t2, t3 = t0 * F["f0298"], t1 * F["f2053"], t2 * F["f3072"], t3 * F["f1501"]
    z1, z2, z3, z4 = z1 * (-F["f0899"]), z2 * (-F["f2562"]), z3 * (-F["f1961"]), z4 * (-F["f0390"])
    z3, z4 = z3 + z5, z4 + z5
    t0, t1, t2, t3 = t0 + z1 + z3, t1 + z2 + z4, t2 + z2 + z3, t3 + z1 + z4
    return (tmp10 + t3, tmp11 + t2, tmp12 + t1, tmp13 + t0, tmp13 - t0, tmp12 - t1, tmp11 - t2, tmp10 - t3)


def idct_islow(coef: np.ndarray, quant: np.ndarray) -> np.ndarray:
    """coef [..., 64] natural order (int), quant [64] -> uint8 [..., 8, 8] (jidctint.c
    jpeg_idct_islow: columns first with DESCALE by 11, then rows with DESCALE by 18)."""
    d = (coef.astype(np.int64) * quant.astype(np.int64)).reshape(coef.shape[:-1] + (8, 8))
    cols = _idct_1d(*[d[..., r, :] for r in range(8)])  # each [..., 8]: index = column
    ws = np.stack([(c + (1 << 10)) >> 11 for c in cols], axis=-2)  # [..., row, col]
    rows = _idct_1d(*[ws[..., :, c] for c in range(8)])  # each [..., 8]: index = row
    out = np.stack([(r + (1 << 17)) >> 18 for r in rows], axis=-1)  # [..., row, col]
    return _range_limit(out)


# ---------------------------------------------------------------------------- upsampling
def _h2v1_fancy(p: np.ndarray) -> np.ndarray:
    a = p.astype(np.int64)
    n = a.shape[1]
    out = np.empty((a.shape[0], 2 * n), dtype=np.int64)
    left = np.concatenate([a[:, :1], a[:, :-1]], axis=1)
    right = np.concatenate([a[:, 1:], a[:, -1:]], axis=1)
    out[:, 0::2] = (3 * a + left + 1) >> 2
    out[:, 1::2] = (3 * a + right + 2) >> 2
    out[:, 0] = a[:, 0]
    out[:, -1] = a[:, -1]
    return out.astype(np.uint8)


def _h2v2_fancy(p: np.ndarray) -> np.ndarray:
    a = p.astype(np.int64)
    rows, n = a.shape
    above = np.concatenate([a[:1], a[:-1]], axis=0)
    below = np.concatenate([a[1:], a[-1:]], axis=0)
    out = np.empty((2 * rows, 2 * n), dtype=np.int64)
    for v, other in ((0, above), (1, below)):
        s = 3 * a + other  # "thiscolsum"
        last = np.concatenate([s[:, :1], s[:, :-1]], axis=1)
        nxt = np.concatenate([s[:, 1:], s[:, -1:]], axis=1)
        even = (3 * s + last + 8) >> 4
        odd = (3 * s + nxt + 7) >> 4
        even[:, 0] = (4 * s[:, 0] + 8) >> 4
        odd[:, -1] = (4 * s[:, -1] + 7) >> 4
        out[v::2, 0::2] = even
        out[v::2, 1::2] = odd
    return out.astype(np.uint8)


def _h1v2_fancy(p: np.ndarray) -> np.ndarray:
    a = p.astype(np.int64)
    above = np.concatenate([a[:1], a[:-1]], axis=0)
    below = np.concatenate([a[1:], a[-1:]], axis=0)
    out = np.empty((2 * a.shape[0], a.shape[1]), dtype=np.int64)
    out[0::2] = (3 * a + above + 1) >> 2
    out[1::2] = (3 * a + below + 2) >> 2
    return out.astype(np.uint8)


def upsample(plane: np.ndarray, hexp: int, vexp: int) -> np.ndarray:
    """plane = the component's real samples [downsampled_height, downsampled_width]."""
    if hexp == 1 and vexp == 1:
        return plane
    if hexp == 2 and vexp == 1:
        return _h2v1_fancy(plane) if plane.shape[1] > 2 else np.repeat(plane, 2, axis=1)
    if hexp == 2 and vexp == 2:
        return _h2v2_fancy(plane) if plane.shape[1] > 2 else np.repeat(np.repeat(plane, 2, axis=0), 2, axis=1)
    if hexp == 1 and vexp == 2:
        return _h1v2_fancy(plane)
    return np.repeat(np.repeat(plane, vexp, axis=0), hexp, axis=1)


# --------------------------------------------------------------------------------- colour
def ycc_to_rgb(y: np.ndarray, cb: np.ndarray, cr: np.ndarray) -> np.ndarray:
    def fix(x):
        return int(x * 65536 + 0.5)

    x = np.arange(256, dtype=np.int64) - 128
    cr_r = (fix(1.40200) * x + 32768) >> 16
    cb_b = (fix(1.77200) * x + 32768) >> 16
    cr_g = -fix(0.71414) * x
    cb_g = -fix(0.34414) * x + 32768
    yy = y.astype(np.int64)
    r = np.clip(yy + cr_r[cr], 0, 255)
    g = np.clip(yy + ((cb_g[cb] + cr_g[cr]) >> 16), 0, 255)
    b = np.clip(yy + cb_b[cb], 0, 255)
    return np.stack([r, g, b], axis=-1).astype(np.uint8)


def apply_orientation(img: np.ndarray, orientation: int) -> np.ndarray:
    """OpenCV's ExifTransform (loadsave.cpp): 1 none, 2 flip-h, 3 rot180, 4 flip-v,
    5 transpose, 6 rot90 cw, 7 transverse, 8 rot90 ccw."""
    if orientation == 2:
        return img[:, ::-1]
    if orientation == 3:
        return img[::-1, ::-1]
    if orientation == 4:
        return img[::-1]
    if orientation == 5:
        return img.transpose(1, 0, 2)
    if orientation == 6:
        return img.transpose(1, 0, 2)[:, ::-1]
    if orientation == 7:
        return img.transpose(1, 0, 2)[::-1, ::-1]
    if orientation == 8:
        return img.transpose(1, 0, 2)[::-1]
    return img


# ------------------------------------------------------------------------------ the chain
def decode_planes(data: bytes, info: dict | None = None):
    """-> (info, [uint8 plane per component, cropped to its real downsampled size])."""
    info = info or parse(data)
    coefs = decode_coefficients(data, info)
    comps = info["comps"]
    hmax, vmax = max(c["h"] for c in comps), max(c["v"] for c in comps)
    planes = []
    for c, cf in zip(comps, coefs):
        px = idct_islow(cf, info["qt"][c["tq"]])  # [by, bx, 8, 8]
        full = px.transpose(0, 2, 1, 3).reshape(cf.shape[0] * 8, cf.shape[1] * 8)
        dw = -(-info["width"] * c["h"] // hmax)
        dh = -(-info["height"] * c["v"] // vmax)
        planes.append(full[:dh, :dw])
    return info, planes


def load_image_from_bytes(data: bytes) -> np.ndarray:
    """RGB uint8 [H, W, 3] — transforms.py:77-110 for a baseline JPEG."""
    if not data:
        raise ValueError("Failed to decode image from bytes: empty input")
    info, planes = decode_planes(data)
    comps = info["comps"]
    hmax, vmax = max(c["h"] for c in comps), max(c["v"] for c in comps)
    H, W = info["height"], info["width"]
    up = []
    for c, p in zip(comps, planes):
        if hmax % c["h"] or vmax % c["v"]:
            raise UnsupportedJpeg("fractional sampling ratio")
        up.append(upsample(p, hmax // c["h"], vmax // c["v"])[:H, :W])
    if len(up) == 1:
        rgb = np.repeat(up[0][..., None], 3, axis=-1)
    else:
        rgb = ycc_to_rgb(*up)
    return np.ascontiguousarray(apply_orientation(rgb, info["orientation"]))
