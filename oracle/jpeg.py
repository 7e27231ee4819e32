"""CPU restatement of what ``load_image_from_bytes`` does to a baseline JPEG.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

The reference calls ``cv2.imdecode(buf, IMREAD_COLOR)`` + ``cvtColor(BGR2RGB)``
(src/shared/processing/transforms.py:77-110).  The arithmetic lives in a third-party
dependency that is not under /root/reference: OpenCV's bundled libjpeg-turbo (opencv-python
4.13.0 here bundles 3.1.2; the reference pins opencv-python 4.11.0.86).  This module restates
libjpeg(-turbo)'s published decoder with its default settings, the ones OpenCV uses:

  * the header checks that make libjpeg give up (jdmarker.c, jdinput.c, jdhuff.c table
    validation) and the standard Huffman tables it installs when a file has none;
  * sequential Huffman decoding, ITU T.81 Annex F.2.2 (jdhuff.c), including what it does with
    damaged data: invalid codes, runs past the end of a block, zero bits behind a marker and
    the MCUs left zero after them, restarts by MCU count and the marker resynchronisation of
    jdmarker.c jpeg_resync_to_restart; data that ends without a marker raises, as OpenCV's
    source manager makes libjpeg fail there;
  * dequantisation + the "slow-but-accurate" integer inverse DCT, jidctint.c
    (CONST_BITS 13, PASS1_BITS 2, Loeffler-Ligtenberg-Moschytz), JDCT_ISLOW, evaluated the way
    libjpeg-turbo's SIMD code does (16-bit lanes, saturating packs): identical to the C code
    for every block an encoder produces, and what cv2 returns for garbage coefficients;
  * "fancy" (triangle-filter) chroma upsampling, jdsample.c: h2v1, h2v2, h1v2; plain
    replication when the subsampled width is <= 2 or for other integral factors;
  * YCbCr -> RGB with the 16-bit fixed-point tables of jdcolor.c; grayscale replicated;
  * the EXIF orientation OpenCV applies after decoding (modules/imgcodecs/src/exif.cpp,
    loadsave.cpp ApplyExifOrientation).

It is pinned by tests/test_oracle_jpeg.py against ``cv2.imdecode`` itself (every sampling
mode, odd sizes, restart intervals, optimised tables, qualities 5..100; damaged headers through
scripts/jpeg_header_fuzz.py; cut, renumbered and damaged scans) and against the committed
fixtures produced through the reference's own ``load_image_from_bytes``
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


# ITU T.81 Annex K.3-K.6.  libjpeg-turbo installs them for table ids 0 (luminance) and 1
# (chrominance) when the file defines none (jdhuff.c jinit_huff_decoder -> std_huff_tables:
# "Motion JPEG frames typically do not include the Huffman tables").
_STD_DC = {
    0: ([0, 1, 5, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0, 0, 0], list(range(12))),
    1: ([0, 3, 1, 1, 1, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0], list(range(12))),
}
_STD_AC = {
    0: ([0, 2, 1, 3, 3, 2, 4, 3, 5, 5, 4, 4, 0, 0, 1, 125],
        [0x01, 0x02, 0x03, 0x00, 0x04, 0x11, 0x05, 0x12, 0x21, 0x31, 0x41, 0x06, 0x13, 0x51, 0x61, 0x07,
         0x22, 0x71, 0x14, 0x32, 0x81, 0x91, 0xA1, 0x08, 0x23, 0x42, 0xB1, 0xC1, 0x15, 0x52, 0xD1, 0xF0,
         0x24, 0x33, 0x62, 0x72, 0x82, 0x09, 0x0A, 0x16, 0x17, 0x18, 0x19, 0x1A, 0x25, 0x26, 0x27, 0x28,
         0x29, 0x2A, 0x34, 0x35, 0x36, 0x37, 0x38, 0x39, 0x3A, 0x43, 0x44, 0x45, 0x46, 0x47, 0x48, 0x49,
         0x4A, 0x53, 0x54, 0x55, 0x56, 0x57, 0x58, 0x59, 0x5A, 0x63, 0x64, 0x65, 0x66, 0x67, 0x68, 0x69,
         0x6A, 0x73, 0x74, 0x75, 0x76, 0x77, 0x78, 0x79, 0x7A, 0x83, 0x84, 0x85, 0x86, 0x87, 0x88, 0x89,
         0x8A, 0x92, 0x93, 0x94, 0x95, 0x96, 0x97, 0x98, 0x99, 0x9A, 0xA2, 0xA3, 0xA4, 0xA5, 0xA6, 0xA7,
         0xA8, 0xA9, 0xAA, 0xB2, 0xB3, 0xB4, 0xB5, 0xB6, 0xB7, 0xB8, 0xB9, 0xBA, 0xC2, 0xC3, 0xC4, 0xC5,
         0xC6, 0xC7, 0xC8, 0xC9, 0xCA, 0xD2, 0xD3, 0xD4, 0xD5, 0xD6, 0xD7, 0xD8, 0xD9, 0xDA, 0xE1, 0xE2,
         0xE3, 0xE4, 0xE5, 0xE6, 0xE7, 0xE8, 0xE9, 0xEA, 0xF1, 0xF2, 0xF3, 0xF4, 0xF5, 0xF6, 0xF7, 0xF8,
         0xF9, 0xFA]),
    1: ([0, 2, 1, 2, 4, 4, 3, 4, 7, 5, 4, 4, 0, 1, 2, 119],
        [0x00, 0x01, 0x02, 0x03, 0x11, 0x04, 0x05, 0x21, 0x31, 0x06, 0x12, 0x41, 0x51, 0x07, 0x61, 0x71,
         0x13, 0x22, 0x32, 0x81, 0x08, 0x14, 0x42, 0x91, 0xA1, 0xB1, 0xC1, 0x09, 0x23, 0x33, 0x52, 0xF0,
         0x15, 0x62, 0x72, 0xD1, 0x0A, 0x16, 0x24, 0x34, 0xE1, 0x25, 0xF1, 0x17, 0x18, 0x19, 0x1A, 0x26,
         0x27, 0x28, 0x29, 0x2A, 0x35, 0x36, 0x37, 0x38, 0x39, 0x3A, 0x43, 0x44, 0x45, 0x46, 0x47, 0x48,
         0x49, 0x4A, 0x53, 0x54, 0x55, 0x56, 0x57, 0x58, 0x59, 0x5A, 0x63, 0x64, 0x65, 0x66, 0x67, 0x68,
         0x69, 0x6A, 0x73, 0x74, 0x75, 0x76, 0x77, 0x78, 0x79, 0x7A, 0x82, 0x83, 0x84, 0x85, 0x86, 0x87,
         0x88, 0x89, 0x8A, 0x92, 0x93, 0x94, 0x95, 0x96, 0x97, 0x98, 0x99, 0x9A, 0xA2, 0xA3, 0xA4, 0xA5,
         0xA6, 0xA7, 0xA8, 0xA9, 0xAA, 0xB2, 0xB3, 0xB4, 0xB5, 0xB6, 0xB7, 0xB8, 0xB9, 0xBA, 0xC2, 0xC3,
         0xC4, 0xC5, 0xC6, 0xC7, 0xC8, 0xC9, 0xCA, 0xD2, 0xD3, 0xD4, 0xD5, 0xD6, 0xD7, 0xD8, 0xD9, 0xDA,
         0xE2, 0xE3, 0xE4, 0xE5, 0xE6, 0xE7, 0xE8, 0xE9, 0xEA, 0xF2, 0xF3, 0xF4, 0xF5, 0xF6, 0xF7, 0xF8,
         0xF9, 0xFA]),
}

_BAD = "Failed to decode image: "

def _next_marker(data: bytes, i: int):
    """jdmarker.c next_marker: skip to an FF, skip fill FFs; FF00 is stuffing, not a marker."""
    n = len(data)
    while True:
        while i < n and data[i] != 0xFF:
            i += 1
        while i < n and data[i] == 0xFF:
            i += 1
        if i >= n:
            raise ValueError(_BAD + "no scan before the end of the data")
        if data[i] != 0:
            return data[i], i + 1
        i += 1


def parse(data: bytes) -> dict:
    """Headers of a baseline / extended-sequential Huffman JPEG with one interleaved scan.

    What makes libjpeg give up (ERREXIT -> cv2.imdecode returns None -> the reference raises
    ValueError) raises ValueError here too; jdmarker.c read_markers / get_sof / get_sos / get_dht /
    get_dqt / get_dri, jdinput.c initial_setup / per_scan_setup / latch_quant_tables.
    """
    # OpenCV picks its decoder by the three-byte signature FF D8 FF
    if len(data) < 4 or data[0] != 0xFF or data[1] != 0xD8 or data[2] != 0xFF:
        raise ValueError(_BAD + "not a JPEG stream")
    info = {"qt": {}, "dc": {}, "ac": {}, "restart": 0, "orientation": 1, "adobe_transform": None}
    i = 2
    saw_exif = False
    while True:
        m, i = _next_marker(data, i)
        if 0xD0 <= m <= 0xD7 or m == 0x01:
            continue  # parameterless
        if m == 0xD8:
            raise ValueError(_BAD + "second SOI")
        if m == 0xD9:
            raise ValueError(_BAD + "EOI before SOS")
        if not (0xC0 <= m <= 0xCF or m in (0xDA, 0xDB, 0xDC, 0xDD, 0xFE) or 0xE0 <= m <= 0xEF):
            raise ValueError(_BAD + f"unknown marker 0x{m:02X}")
        if i + 2 > len(data):
            raise ValueError(_BAD + "truncated")
        L = _u16(data, i)
        if L < 2 and (0xE0 <= m <= 0xEF or m in (0xFE, 0xDC)):
            L = 2  # skip_variable / get_interesting_appn skip nothing for a length below 2
        if L < 2 or i + L > len(data):
            raise ValueError(_BAD + "bad segment length")
        seg = data[i + 2:i + L]
        if m == 0xCC:  # DAC: pointless in a Huffman file, but get_dac still checks it
            if len(seg) % 2:
                raise ValueError(_BAD + "bad DAC length")
            for k in range(0, len(seg), 2):
                index, val = seg[k], seg[k + 1]
                if index >= 32 or (index < 16 and (val & 15) > (val >> 4)):
                    raise ValueError(_BAD + "bad DAC")
        if m == 0xDB:  # DQT
            k = 0
            while k < len(seg):
                pq, tq = seg[k] >> 4, seg[k] & 15
                k += 1
                size = 128 if pq else 64  # any non-zero precision nibble means 16-bit entries
                if tq > 3 or k + size > len(seg):
                    raise ValueError(_BAD + "bad DQT")
                if pq:
                    vals = [(seg[k + 2 * j] << 8) | seg[k + 2 * j + 1] for j in range(64)]
                else:
                    vals = list(seg[k:k + 64])
                k += size
                q = np.zeros(64, dtype=np.int64)
                q[ZIGZAG] = vals  # stored in zigzag order -> natural order
                info["qt"][tq] = q
        elif m in (0xC0, 0xC1):  # SOF0 / SOF1 (Huffman, sequential)
            if "comps" in info:
                raise ValueError(_BAD + "second SOF")
            if len(seg) < 6:
                raise ValueError(_BAD + "bad SOF length")
            info["height"], info["width"] = _u16(seg, 1), _u16(seg, 3)
            nc = seg[5]
            if info["height"] == 0 or info["width"] == 0 or nc == 0:
                raise ValueError(_BAD + "empty image")
            if len(seg) != 6 + 3 * nc:
                raise ValueError(_BAD + "bad SOF length")
            if seg[0] != 8:
                raise UnsupportedJpeg("sample precision != 8")
            if nc not in (1, 3):
                raise UnsupportedJpeg(f"{nc} components")
            info["comps"] =[{"id": seg[6 + 3 * c], "h": seg[7 + 3 * c] >> 4, "v": seg[7 + 3 * c] & 15,
                              "tq": seg[8 + 3 * c]} for c in range(nc)]
        elif 0xC0 <= m <= 0xCF and m not in (0xC4, 0xCC):
            raise UnsupportedJpeg(f"SOF marker 0x{m:02X} (progressive / lossless / arithmetic)")
        elif m == 0xC4:  # DHT
            k = 0
            while k < len(seg):
                if k + 17 > len(seg):
                    raise ValueError(_BAD + "bad DHT length")
                tc, th = seg[k] >> 4, seg[k] & 15
                counts = list(seg[k + 1:k + 17])
                n = sum(counts)
                if tc > 1 or th > 3 or n > 256 or k + 17 + n > len(seg):
                    raise ValueError(_BAD + "bad DHT")
                vals = list(seg[k + 17:k + 17 + n])
                k += 17 + n
                info["ac" if tc else "dc"][th] = (counts, vals)
        elif m == 0xDD:
            if len(seg) != 2:
                raise ValueError(_BAD + "bad DRI length")
            info["restart"] = _u16(seg, 0)
        elif m == 0xE1 and not saw_exif and seg[:6] == b"Exif\x00\x00":
            saw_exif = True
            info["orientation"] = _exif_orientation(seg)
        elif m == 0xEE and seg[:5] == b"Adobe" and len(seg) >= 12:
            info["adobe_transform"] = seg[11]
        elif m == 0xE0 and seg[:5] == b"JFIF\x00" and len(seg) >= 14:
            info["jfif"] = True
        elif m == 0xDA:  # SOS
            if "comps" not in info:
                raise ValueError(_BAD + "SOS before SOF")
            ns = seg[0] if seg else 0
            if ns < 1 or ns > 4 or len(seg) != 2 * ns + 4:
                raise ValueError(_BAD + "bad SOS length")
            ids = [c["id"] for c in info["comps"]]
            for s in range(ns):
                if seg[1 + 2 * s] not in ids:
                    raise ValueError(_BAD + "scan names a component the frame does not have")
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
    comps = info["comps"]
    nc = len(comps)
    if info["width"] > 65500 or info["height"] > 65500:
        raise ValueError(_BAD + "image too big")
    for c in comps:
        if not (1 <= c["h"] <= 4 and 1 <= c["v"] <= 4):
            raise ValueError(_BAD + "bad sampling factors")
    # libjpeg's colour-space guess (jdapimin.c default_decompress_parms): JFIF -> YCbCr; else the
    # Adobe marker decides; else component ids 'R','G','B' -> RGB
    if nc == 3 and not info.get("jfif"):
        ids_rgb = [c["id"] for c in comps] == [ord("R"), ord("G"), ord("B")]
        if info["adobe_transform"] == 0 or (info["adobe_transform"] is None and ids_rgb):
            raise UnsupportedJpeg("stored as RGB (no YCbCr transform)")
    if nc == 1:
        comps[0]["h"] = comps[0]["v"] = 1  # sampling factors mean nothing here
    if sum(c["h"] * c["v"] for c in comps) > 10:
        raise ValueError(_BAD + "more than 10 blocks per MCU")
    for c in comps:
        if c["tq"] > 3 or c["tq"] not in info["qt"]:
            raise ValueError(_BAD + "quantisation table not defined")
        if c["td"] > 3 or c["ta"] > 3:
            raise ValueError(_BAD + "Huffman table not defined")
    for t in (0, 1):  # the standard tables stand in for undefined ids 0 and 1
        info["dc"].setdefault(t, _STD_DC[t])
        info["ac"].setdefault(t, _STD_AC[t])
    for c in comps:
        if c["td"] not in info["dc"] or c["ta"] not in info["ac"]:
            raise ValueError(_BAD + "Huffman table not defined")
        _Huff(*info["dc"][c["td"]], is_dc=True)  # raises on a table libjpeg refuses
        _Huff(*info["ac"][c["ta"]])
    return info


# ------------------------------------------------------------------------------ Huffman
class _Huff:
    """Decoding tables of T.81 F.2.2.3 (mincode / maxcode / valptr)."""

    def __init__(self, counts, vals, is_dc=False):
        self.vals = vals
        self.mincode, self.maxcode, self.valptr = [0] * 17, [-1] * 17, [0] * 17
        code = k = 0
        for length in range(1, 17):
            self.valptr[length] = k
            self.mincode[length] = code
            code += counts[length - 1]
            k += counts[length - 1]
            self.maxcode[length] = code - 1 if counts[length - 1] else -1
            # jdhuff.c jpeg_make_d_derived_tbl: the codes of a length may not reach the all-ones
            # code (JERR_BAD_HUFF_TABLE is fatal: cv2.imdecode returns None)
            if code >= (1 << length):
                raise ValueError("Failed to decode image: bogus Huffman table")
            code <<= 1
        # same function: DC symbols are categories 0..15 (more than 16 of them: one is repeated or
        # out of range; the device's tables hold 16)
        if is_dc and (k > 16 or any(v > 15 for v in vals[:k])):
            raise ValueError("Failed to decode image: bogus Huffman table")


class _Bits:
    """libjpeg's bit reader (jdhuff.c jpeg_fill_bit_buffer, the non-"fast" path): bytes are
    fetched ahead until at least 57 bits are buffered; FF00 is a stuffed FF; any other FFxx is
    a marker, which stays unread and ends the data: from then on zeros are supplied, and the
    first request that really needs one of them sets ``insufficient``.  Running out of bytes
    WITHOUT having met a marker is what OpenCV's memory source turns into a failure."""

    MIN_GET_BITS = 57

    def __init__(self, data: bytes, pos: int):
        self.d, self.p, self.acc, self.n = data, pos, 0, 0
        self.marker = 0
        self.insufficient = False

    def _byte(self) -> int:
        if self.p >= len(self.d):
            raise ValueError(_BAD + "the data ends inside the scan")
        self.p += 1
        return self.d[self.p - 1]

    def fill(self, nbits: int):
        if not self.marker:
            while self.n < self.MIN_GET_BITS:
                c = self._byte()
                if c == 0xFF:
                    c = self._byte()
                    while c == 0xFF:
                        c = self._byte()
                    if c == 0:
                        c = 0xFF
                    else:
                        self.marker = c
                        break
                self.acc = ((self.acc << 8) | c) & ((1 << 64) - 1)
                self.n += 8
        if self.marker and nbits > self.n:
            self.insufficient = True
            self.acc = (self.acc << (self.MIN_GET_BITS - self.n)) & ((1 << 64) - 1)
            self.n = self.MIN_GET_BITS

    def bits(self, k: int) -> int:
        if k == 0:
            return 0
        if self.n < k:
            self.fill(k)
        self.n -= k
        return (self.acc >> self.n) & ((1 << k) - 1)

    def next_marker(self) -> int:
        """jdmarker.c next_marker from the current input position."""
        while True:
            c = self._byte()
            while c != 0xFF:
                c = self._byte()
            while c == 0xFF:
                c = self._byte()
            if c != 0:
                return c

    def restart(self, desired: int):
        """jdhuff.c process_restart + jdmarker.c read_restart_marker / jpeg_resync_to_restart."""
        self.acc = self.n = 0  # whole bytes still buffered are dropped with the partial one
        if not self.marker:
            self.marker = self.next_marker()
        if self.marker == 0xD0 + desired:
            self.marker = 0
        else:
            while True:
                m = self.marker
                if m < 0xC0:
                    action = 2  # not a marker libjpeg knows: look for the next one
                elif m < 0xD0 or m > 0xD7:
                    action = 3  # a real marker: leave it, the interval decodes as zeros
                elif m in (0xD0 + ((desired + 1) & 7), 0xD0 + ((desired + 2) & 7)):
                    action = 3  # one of the next two restarts: this interval was lost
                elif m in (0xD0 + ((desired - 1) & 7), 0xD0 + ((desired - 2) & 7)):
                    action = 2  # an earlier restart: skip it
                else:
                    action = 1  # too far off to tell: take it as the wanted one
                if action == 1:
                    self.marker = 0
                    break
                if action == 3:
                    break
                self.marker = self.next_marker()
        if not self.marker:
            self.insufficient = False


def _decode_symbol(br: _Bits, h: _Huff) -> int:
    if br.n < 8:
        br.fill(0)  # HUFF_DECODE looks 8 bits ahead
    code = 0
    for length in range(1, 17):
        code = (code << 1) | br.bits(1)
        if h.maxcode[length] >= 0 and code <= h.maxcode[length] and code >= h.mincode[length]:
            return h.vals[h.valptr[length] + code - h.mincode[length]]
    # invalid code: libjpeg warns and returns 0 — after a 17th bit, because its search loop is
    # stopped by a sentinel maxcode[17]
    br.bits(1)
    return 0


def _extend(v: int, s: int) -> int:
    return v - ((1 << s) - 1) if s and v < (1 << (s - 1)) else v


def decode_coefficients(data: bytes, info: dict, report: dict | None = None):
    """Entropy decode -> per component int array [blocks_y, blocks_x, 64] (natural order,
    not dequantised).  blocks_* are padded to whole MCUs.

    Damaged streams are concealed the way jdhuff.c decode_mcu does: once a request needed bits
    behind a marker, the MCU in hand is finished from zero bits and the following ones stay
    zero until a restart marker is consumed.  ``report["damaged"]`` tells whether that (or a
    restart resynchronisation) happened: the device decoder flags such images instead of
    reproducing the concealment."""
    comps = info["comps"]
    hmax, vmax = max(c["h"] for c in comps), max(c["v"] for c in comps)
    mcux = -(-info["width"] // (8 * hmax))
    mcuy = -(-info["height"] // (8 * vmax))
    coefs = [np.zeros((mcuy * c["v"], mcux * c["h"], 64), dtype=np.int64) for c in comps]
    # only the tables the scan names are derived (and validated), as in start_pass_huff_decoder
    dct = {k: _Huff(*info["dc"][k], is_dc=True) for k in {c["td"] for c in comps}}
    act = {k: _Huff(*info["ac"][k]) for k in {c["ta"] for c in comps}}
    br = _Bits(data, info["scan_start"])
    pred = [0] * len(comps)
    ri = info["restart"]
    to_go, next_rst = ri, 0
    damaged = False
    for m in range(mcux * mcuy):
        if ri:
            if to_go == 0:
                # an intact segment leaves the expected marker pending behind less than a byte of padding
                clean = br.marker == 0xD0 + next_rst and not br.insufficient and br.n < 8
                br.restart(next_rst)
                damaged |= not clean or br.marker != 0
                next_rst = (next_rst + 1) & 7
                pred = [0] * len(comps)
                to_go = ri
            to_go -= 1
        if br.insufficient:
            damaged = True
            continue  # the MCU stays zero
        my, mx = divmod(m, mcux)
        for ci, c in enumerate(comps):
            for by in range(c["v"]):
                for bx in range(c["h"]):
                    blk = coefs[ci][my * c["v"] + by, mx * c["h"] + bx]
                    s = _decode_symbol(br, dct[c["td"]])
                    diff = _extend(br.bits(s), s) if s else 0
                    pred[ci] += diff
                    blk[0] = _w16(pred[ci])  # the predictor is an int, the stored JCOEF a short
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
                        # jpeg_natural_order has 16 extra entries, all 63: a run past the end of
                        # the block (corrupt data) lands on the last coefficient
                        blk[ZIGZAG[min(k, 63)]] = v
                        k += 1
    if report is not None:
        report["damaged"] = damaged or br.insufficient
    return coefs


# ------------------------------------------------------------------------------- IDCT
def _range_limit(x: np.ndarray) -> np.ndarray:
    """libjpeg's ``range_limit[x & RANGE_MASK]`` (table centred on 128): x in [-128, 127] ->
    x + 128, saturating outside, wrapping modulo 1024 like the table lookup."""
    idx = (x.astype(np.int64)) & 1023
    out = np.where(idx < 128, idx + 128, np.where(idx < 512, 255, np.where(idx < 896, 0, idx - 896)))
    return out.astype(np.uint8)


def _w16(x):
    return ((x + 32768) & 0xFFFF) - 32768


def _w32(x):
    return ((x + (1 << 31)) & 0xFFFFFFFF) - (1 << 31)


def _sat16(x):
    return np.clip(x, -32768, 32767)


def _idct_1d(d0, d1, d2, d3, d4, d5, d6, d7, shift):
    """One LL&M pass of jidctint.c as libjpeg-turbo's SIMD code evaluates it (jidctint-sse2 /
    -avx2): the same integers as the C code for every block an encoder can produce; for
    out-of-range (corrupt) data the 16-bit lanes wrap in the sums in0 +- in4, in7 + in3, in5 +
    in1, the rotations are sums of two 16x16 products (pmaddwd) with the constants pre-combined,
    32-bit sums wrap, and the descaled results are packed to 16 bits with saturation.
    Inputs: int64 arrays holding 16-bit values."""
    tmp3 = _w32(d2 * (4433 + 6270) + d6 * 4433)
    tmp2 = _w32(d2 * 4433 + d6 * (4433 - 15137))
    tmp0 = _w16(d0 + d4) << 13
    tmp1 = _w16(d0 - d4) << 13
    tmp10, tmp13 = _w32(tmp0 + tmp3), _w32(tmp0 - tmp3)
    tmp11, tmp12 = _w32(tmp1 + tmp2), _w32(tmp1 - tmp2)
    z3, z4 = _w16(d7 + d3), _w16(d5 + d1)
    z3, z4 = _w32(z3 * (9633 - 16069) + z4 * 9633), _w32(z3 * 9633 + z4 * (9633 - 3196))
    t0 = _w32(_w32(d7 * (2446 - 7373) + d1 * -7373) + z3)
    t3 = _w32(_w32(d7 * -7373 + d1 * (12299 - 7373)) + z4)
    t1 = _w32(_w32(d5 * (16819 - 20995) + d3 * -20995) + z4)
    t2 = _w32(_w32(d5 * -20995 + d3 * (25172 - 20995)) + z3)
    outs = (tmp10 + t3, tmp11 + t2, tmp12 + t1, tmp13 + t0, tmp13 - t0, tmp12 - t1, tmp11 - t2, tmp10 - t3)
    return [_sat16(_w32(_w32(o) + (1 << (shift - 1))) >> shift) for o in outs]


def idct_left_16bit_range(coef: np.ndarray, quant: np.ndarray) -> np.ndarray:
    """bool [...]: True for a block in which a 16-bit lane of the SIMD evaluation wraps or a
    descaled result saturates (dequantised coefficient, the sums in0 +- in4 / in7 + in3 / in5 +
    in1 of either pass, the packed results of either pass).  While it is False the SIMD code is
    plain jidctint.c integer arithmetic — the only case in which the pixels are certain to be
    cv2's; the device reports frames containing such a block as damaged (status bit 0)."""
    c = coef.astype(np.int64)
    dq = c * quant.astype(np.int64)
    bad = np.any(dq != _w16(dq), axis=-1)
    d = _w16(dq).reshape(c.shape[:-1] + (8, 8))

    def lanes(rows):  # rows: list of 8 arrays (the 8 inputs of a 1-D pass)
        d0, d1, d2, d3, d4, d5, d6, d7 = rows
        out = np.zeros(d0.shape, dtype=bool)
        for v in (d0 + d4, d0 - d4, d7 + d3, d5 + d1):
            out |= v != _w16(v)
        return out

    def unsat(rows, shift):
        d0, d1, d2, d3, d4, d5, d6, d7 = rows
        tmp3 = d2 * (4433 + 6270) + d6 * 4433
        tmp2 = d2 * 4433 + d6 * (4433 - 15137)
        tmp0, tmp1 = _w16(d0 + d4) << 13, _w16(d0 - d4) << 13
        z3, z4 = _w16(d7 + d3), _w16(d5 + d1)
        z3, z4 = z3 * (9633 - 16069) + z4 * 9633, z3 * 9633 + z4 * (9633 - 3196)
        t0 = d7 * (2446 - 7373) + d1 * -7373 + z3
        t3 = d7 * -7373 + d1 * (12299 - 7373) + z4
        t1 = d5 * (16819 - 20995) + d3 * -20995 + z4
        t2 = d5 * -20995 + d3 * (25172 - 20995) + z3
        outs = (tmp0 + tmp3 + t3, tmp1 + tmp2 + t2, tmp1 - tmp2 + t1, tmp0 - tmp3 + t0,
                tmp0 - tmp3 - t0, tmp1 - tmp2 - t1, tmp1 + tmp2 - t2, tmp0 + tmp3 - t3)
        res = np.zeros(d0.shape, dtype=bool)
        for o in outs:
            v = (_w32(o) + (1 << (shift - 1))) >> shift
            res |= v != _sat16(v)
        return res

    flat = ~np.any(c[..., 8:] != 0, axis=-1)
    rows1 = [d[..., r, :] for r in range(8)]
    pass1 = np.any(lanes(rows1) | unsat(rows1, 11), axis=-1)
    short = np.any((d[..., 0, :] << 2) != _w16(d[..., 0, :] << 2), axis=-1)
    bad |= np.where(flat, short, pass1)
    ws = np.stack(_idct_1d(*rows1, shift=11), axis=-2)
    ws = np.where(flat[..., None, None], _w16(d[..., :1, :] << 2), ws)
    rows2 = [ws[..., :, k] for k in range(8)]
    bad |= np.any(lanes(rows2) | unsat(rows2, 18), axis=-1)
    return bad


def idct_islow(coef: np.ndarray, quant: np.ndarray) -> np.ndarray:
    """coef [..., 64] natural order (int16 values), quant [64] -> uint8 [..., 8, 8] (jidctint.c
    jpeg_idct_islow: columns first with DESCALE by 11, then rows with DESCALE by 18, the
    result clamped to [-128, 127] + 128), evaluated the way the SIMD code does — see _idct_1d."""
    c = coef.astype(np.int64)
    d = _w16(c * quant.astype(np.int64)).reshape(c.shape[:-1] + (8, 8))  # pmullw
    cols = _idct_1d(*[d[..., r, :] for r in range(8)], shift=11)  # each [..., 8]: index = column
    ws = np.stack(cols, axis=-2)  # [..., row, col]
    # a block whose rows 1..7 are all zero takes a shortcut: row 0 * q << PASS1_BITS in 16-bit lanes
    flat = ~np.any(c[..., 8:] != 0, axis=-1)
    ws = np.where(flat[..., None, None], _w16(d[..., :1, :] << 2), ws)
    rows = _idct_1d(*[ws[..., :, k] for k in range(8)], shift=18)  # each [..., 8]: index = row
    out = np.stack(rows, axis=-1)  # [..., row, col]
    return (np.clip(out, -128, 127) + 128).astype(np.uint8)


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
