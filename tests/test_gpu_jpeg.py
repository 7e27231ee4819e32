"""N3 parity: baseline JPEG decode on the GPU vs the oracle (oracle/jpeg.py, pinned against
cv2.imdecode) and vs cv2.imdecode itself — the call load_image_from_bytes makes
(transforms.py:102).  Pixels, coefficients and shapes: bit-exact."""

import numpy as np
import pytest
import torch

from oracle import jpeg as J
from workloads import jpeg_cases as W

pytestmark = pytest.mark.gpu
cv2 = pytest.importorskip("cv2")


@pytest.fixture(scope="module")
def M():
    from shared.processing import jpeg
    return jpeg


def _scan_order_coefs(data):
    """Oracle coefficients in the device layout: blocks in MCU scan order, natural order."""
    info = J.parse(data)
    coefs = J.decode_coefficients(data, info)
    comps = info["comps"]
    hmax, vmax = max(c["h"] for c in comps), max(c["v"] for c in comps)
    mcux, mcuy = -(-info["width"] // (8 * hmax)), -(-info["height"] // (8 * vmax))
    out = []
    for my in range(mcuy):
        for mx in range(mcux):
            for c, cf in zip(comps, coefs):
                for by in range(c["v"]):
                    for bx in range(c["h"]):
                        out.append(cf[my * c["v"] + by, mx * c["h"] + bx])
    return np.stack(out)


@pytest.mark.parametrize("sampling", W.SAMPLINGS)
def test_coefficients_match_oracle(M, sampling):
    data = W.encode(W.photo(40, 56, seed=21), 85, sampling)
    batch = M.decode_jpeg_batch_cuda([data])
    torch.cuda.synchronize()
    assert batch.status.cpu().tolist() == [0]
    assert np.array_equal(batch.coefficients(0).cpu().numpy().astype(np.int64), _scan_order_coefs(data))


@pytest.mark.parametrize("sampling", W.SAMPLINGS)
@pytest.mark.parametrize("shape", [(16, 16), (48, 64), (33, 17), (17, 33), (1, 1), (2, 3), (8, 5), (5, 4), (64, 6)])
def test_pixels_match_oracle(M, shape, sampling):
    blobs, wants = [], []
    for q, kw in ((5, {}), (50, {"restart": 3}), (90, {"optimize": True}), (100, {})):
        data = W.encode(W.photo(*shape, seed=shape[0] * shape[1] + q), q, sampling, **kw)
        blobs.append(data)
        wants.append(J.load_image_from_bytes(data))
    batch = M.decode_jpeg_batch_cuda(blobs)
    torch.cuda.synchronize()
    assert batch.status.cpu().tolist() == [0] * len(blobs)
    for got, want in zip(batch.frames, wants):
        assert got.shape == want.shape
        assert np.array_equal(got.cpu().numpy(), want)


def test_reference_fixtures(M, golden):
    """The committed files, decoded by the unmodified reference when the fixtures were made."""
    cases = golden.meta["jpeg"]
    blobs = [golden[f"jpeg{i}_bytes"].tobytes() for i in range(len(cases))]
    batch = M.decode_jpeg_batch_cuda(blobs)
    torch.cuda.synchronize()
    assert batch.status.cpu().tolist() == [0] * len(cases)
    for i, frame in enumerate(batch.frames):
        assert np.array_equal(frame.cpu().numpy(), golden[f"jpeg{i}_rgb"]), cases[i]


def test_grayscale(M):
    blobs = [W.encode(W.photo(*shape, seed=3)[..., 0], 80) for shape in ((20, 30), (7, 9), (31, 250))]
    batch = M.decode_jpeg_batch_cuda(blobs)
    for got, data in zip(batch.frames, blobs):
        assert np.array_equal(got.cpu().numpy(), W.reference_decode(data))


@pytest.mark.parametrize("orientation", range(1, 9))
def test_exif_orientation(M, orientation):
    data = W.with_exif_orientation(W.encode(W.photo(24, 40, seed=9), 85, "420"), orientation)
    got = M.load_jpeg_from_bytes(data)
    want = W.reference_decode(data)
    assert got.shape == want.shape and np.array_equal(got, want)


@pytest.mark.parametrize("shape,sampling,q,restart", [
    ((480, 640), "420", 90, 0), ((480, 640), "444", 75, 0), ((1080, 1920), "420", 90, 0),
    ((1080, 1920), "422", 50, 120), ((1080, 1920), "420", 100, 0), ((721, 1283), "420", 95, 7),
    ((300, 5000), "440", 80, 0)])
def test_large_frames_match_cv2(M, shape, sampling, q, restart):
    """Many 64 KB chunks, thousands of subsequences, the carried state between chunks."""
    data = W.encode(W.photo(*shape, seed=shape[0] + q), q, sampling, restart=restart)
    batch = M.decode_jpeg_batch_cuda([data])
    torch.cuda.synchronize()
    assert batch.status.cpu().tolist() == [0]
    assert np.array_equal(batch.frames[0].cpu().numpy(), W.reference_decode(data))


def test_noise_image_long_codes(M):
    """Pure noise at quality 100: long Huffman codes (slow-path look-ups), few end-of-blocks."""
    rng = np.random.default_rng(5)
    img = rng.integers(0, 256, (96, 128, 3), dtype=np.uint8)
    for sampling in ("444", "420"):
        data = W.encode(img, 100, sampling)
        assert np.array_equal(M.load_jpeg_from_bytes(data), W.reference_decode(data))
    flat = np.full((64, 64, 3), 77, np.uint8)  # DC-only blocks: almost no bits per block
    data = W.encode(flat, 90, "420")
    assert np.array_equal(M.load_jpeg_from_bytes(data), W.reference_decode(data))


def test_one_byte_restart_segments(M):
    """Flat images with a restart marker after every MCU: segments of one or two bytes, so a
    symbol that starts in the padding in front of a marker can reach past the whole next
    segment — also past the last one."""
    blobs = []
    for value in (0, 77, 128, 255):
        flat = np.full((40, 72, 3), value, np.uint8)
        flat[..., 1] = 255 - value
        for sampling in W.SAMPLINGS:
            blobs.append(W.encode(flat, 90, sampling, restart=1))
            blobs.append(W.encode(flat, 30, sampling, restart=2, optimize=True))
        blobs.append(W.encode(flat[..., 0], 90, restart=1))
    ramp = np.zeros((24, 64, 3), np.uint8)
    ramp[:, 32:] = 200  # a single edge: most segments stay tiny, a few are not
    blobs += [W.encode(ramp, 95, s, restart=1) for s in W.SAMPLINGS]
    batch = M.decode_jpeg_batch_cuda(blobs)
    torch.cuda.synchronize()
    assert batch.status.cpu().tolist() == [0] * len(blobs)
    for got, data in zip(batch.frames, blobs):
        assert np.array_equal(got.cpu().numpy(), W.reference_decode(data))


def test_mixed_batch(M):
    """Different sizes, samplings and table sets in one call; two calls in a row."""
    rng = np.random.default_rng(7)
    blobs = []
    for i in range(24):
        h, w = int(rng.integers(1, 200)), int(rng.integers(1, 300))
        img = W.photo(h, w, seed=100 + i)
        if i % 5 == 4:
            blobs.append(W.encode(img[..., 0], int(rng.integers(20, 100))))
        else:
            blobs.append(W.encode(img, int(rng.integers(20, 100)), W.SAMPLINGS[i % 5], restart=int(i % 3) * 2,
                                  optimize=bool(i & 1)))
    for _ in range(2):
        batch = M.decode_jpeg_batch_cuda(blobs)
        torch.cuda.synchronize()
        assert batch.status.cpu().tolist() == [0] * len(blobs)
        for got, data in zip(batch.frames, blobs):
            assert np.array_equal(got.cpu().numpy(), W.reference_decode(data))


def test_16bit_quantisation_tables_and_comment_segments(M):
    """Header variants encoders rarely emit: 16-bit DQT entries, a COM segment full of 0xFF
    bytes and marker look-alikes in front of the scan."""
    data = W.encode(W.photo(120, 200, seed=12), 40, "420", restart=3)
    wide = W.with_16bit_quant_tables(data)
    junk = bytes([0xFF, 0xD8, 0xFF, 0xDA, 0xFF, 0x00, 0xFF, 0xD9] * 9)
    noisy = wide[:2] + b"\xff\xfe" + (len(junk) + 2).to_bytes(2, "big") + junk + wide[2:]
    sof = noisy.index(b"\xff\xc0", 2 + 4 + len(junk))
    extended = noisy[:sof] + b"\xff\xc1" + noisy[sof + 2:]  # SOF1: extended sequential, same decoding
    import re
    start = J.parse(data)["scan_start"]  # fill bytes (extra 0xFF) in front of every restart marker
    filled = data[:start] + re.sub(rb"\xff([\xd0-\xd7])", lambda m: b"\xff\xff\xff" + m.group(1), data[start:])
    assert len(filled) > len(data)
    for blob in (wide, noisy, extended, filled):
        assert np.array_equal(M.load_jpeg_from_bytes(blob), W.reference_decode(blob))


def test_errors(M):
    with pytest.raises(M.UnsupportedJpeg):
        M.decode_jpeg_batch_cuda([W.encode(W.photo(16, 16, 1), 90, "420", progressive=True)])
    with pytest.raises(ValueError, match="Failed to decode"):
        M.load_jpeg_from_bytes(b"garbage")
    assert M.decode_jpeg_batch_cuda([]).frames == []


@pytest.mark.timeout(300)
def test_corrupt_streams_terminate(M):
    """Damaged entropy data (flipped bytes, inserted markers, cuts) must neither hang nor fault:
    every call returns frames of the header's size, and a clean file decodes exactly afterwards.
    (The pixels of a damaged file are not compared: libjpeg's error concealment is out of scope.)"""
    rng = np.random.default_rng(11)
    good = W.encode(W.photo(120, 200, seed=4), 85, "420", restart=5)
    start = J.parse(good)["scan_start"]
    blobs = []
    for k in range(24):
        b = bytearray(good)
        for _ in range(1 + k % 7):
            pos = int(rng.integers(start, len(b) - 2))
            if k % 3 == 0:
                b[pos] = int(rng.integers(0, 256))
            elif k % 3 == 1:
                b[pos:pos + 2] = bytes([0xFF, int(rng.integers(0xD0, 0xD8))])  # stray restart marker
            else:
                b[pos] = 0xFF  # FF followed by whatever comes next
        if k % 4 == 3:
            b = b[: int(rng.integers(start + 1, len(b)))]
        blobs.append(bytes(b))
    blobs.append(good[:start])  # headers only, empty scan
    batch = M.decode_jpeg_batch_cuda(blobs)
    torch.cuda.synchronize()
    for f in batch.frames:
        assert tuple(f.shape) == (120, 200, 3)
    assert len(batch.status.cpu().tolist()) == len(blobs)
    assert np.array_equal(M.load_jpeg_from_bytes(good), W.reference_decode(good))


def test_truncated_stream_is_flagged(M):
    """Data that ends mid-scan without a marker: libjpeg runs out of input, cv2.imdecode returns
    None, the reference raises — status bit 1, and load_jpeg_from_bytes raises too.  The same cut
    closed by a marker is a damaged but decodable file (bit 0 only): libjpeg finishes the MCU in
    hand from zero bits and leaves the rest grey, and so does the device."""
    data = W.encode(W.photo(64, 64, seed=2), 90, "420")
    cut = data[: len(data) * 2 // 3]
    assert cv2.imdecode(np.frombuffer(cut, np.uint8), cv2.IMREAD_COLOR) is None
    closed = cut + b"\xff\xd9"
    want = W.reference_decode(closed)
    batch = M.decode_jpeg_batch_cuda([cut, closed, data])
    torch.cuda.synchronize()
    assert batch.status.cpu().tolist() == [3, 1, 0]
    assert np.array_equal(batch.frames[1].cpu().numpy(), want)
    assert (want[-8:] == 128).all()
    with pytest.raises(ValueError, match="Failed to decode"):
        M.load_jpeg_from_bytes(cut)
    assert np.array_equal(M.load_jpeg_from_bytes(closed), want)


@pytest.mark.parametrize("sampling,kw", [("420", {}), ("444", {"optimize": True}), ("422", {"restart": 3}),
                                         ("gray", {}), ("411", {"restart": 1})])
def test_scans_cut_short_are_concealed_like_libjpeg(M, sampling, kw):
    """Every way a scan can stop early at a marker — inside a code, inside the magnitude bits, at
    a block or MCU boundary, after whole restart segments — against cv2: pixels equal, status 1."""
    img = W.photo(56, 88, seed=31)
    data = W.encode(img[..., 0], 85) if sampling == "gray" else W.encode(img, 85, sampling, **kw)
    scan = J.parse(data)["scan_start"]
    body = len(data) - 2 - scan
    blobs, wants = [data[:scan] + b"\xff\xd9"], []  # no entropy-coded data at all
    for k in range(40):
        cut = scan + 1 + (k * 37) % (body - 1) if k < 30 else scan + body - (k - 29)
        tail = (b"\xff\xd9", b"\xff\xc4\x00\x02", b"\xff\xfe")[k % 3]
        blobs.append(data[:cut] + tail)
    for blob in blobs:
        ref = cv2.imdecode(np.frombuffer(blob, np.uint8), cv2.IMREAD_COLOR)
        wants.append(None if ref is None else ref[..., ::-1])
    batch = M.decode_jpeg_batch_cuda(blobs)
    torch.cuda.synchronize()
    status = batch.status.cpu().tolist()
    compared = 0
    for blob, st, got, want in zip(blobs, status, batch.frames, wants):
        if want is None:
            continue  # a restart came due and libjpeg's search for its marker ran off the end
        assert st in (0, 1)
        if st == 0:  # the cut fell into the padding of the last byte
            assert np.array_equal(got.cpu().numpy(), want)
            continue
        assert np.array_equal(got.cpu().numpy(), want), (sampling, len(blob))
        compared += 1
    assert compared >= 25


def test_bit_errors_without_restart_markers_decode_like_libjpeg(M):
    """Flipped bytes, stray FFs and stray markers (RSTn too: without a restart interval libjpeg
    never consumes one, it ends the data) in the entropy-coded bytes of files without restart
    markers.  The garbage that follows — runs past the end of a block, invalid codes, predictors
    that leave 16 bits, too few or too many blocks — comes out as libjpeg's, pixel for pixel."""
    rng = np.random.default_rng(3)
    blobs = []
    for i, sampling in enumerate(("420", "444", "422", "gray", "411")):
        img = W.photo(72, 120, seed=50 + i)
        good = W.encode(img[..., 0], 80) if sampling == "gray" else W.encode(img, 80, sampling, optimize=bool(i & 1))
        start = J.parse(good)["scan_start"]
        for k in range(20):
            b = bytearray(good)
            for _ in range(1 + k % 3):
                pos = int(rng.integers(start, len(b) - 3))
                if k % 4 == 0:
                    b[pos] = int(rng.integers(0, 256))
                elif k % 4 == 1:
                    b[pos] ^= 1 << int(rng.integers(0, 8))
                elif k % 4 == 2:
                    b[pos] = 0xFF
                else:
                    b[pos:pos + 2] = bytes([0xFF, int(rng.integers(0xD0, 0xD8))])
            blobs.append(bytes(b))
    wants = []
    for blob in blobs:
        ref = cv2.imdecode(np.frombuffer(blob, np.uint8), cv2.IMREAD_COLOR)
        wants.append(None if ref is None else ref[..., ::-1])
    batch = M.decode_jpeg_batch_cuda(blobs)
    torch.cuda.synchronize()
    status = batch.status.cpu().tolist()
    compared = damaged = 0
    for st, got, want in zip(status, batch.frames, wants):
        if want is None:
            continue
        assert np.array_equal(got.cpu().numpy(), want)
        compared += 1
        damaged += st != 0
    assert compared >= 90 and damaged >= 30


def _bit_errors_inside_restart_segments(seed=9, per_file=16):
    """Files with restart intervals whose entropy-coded bytes carry bit errors while every
    marker stays where it was: some segments come out short, some long."""
    rng = np.random.default_rng(seed)
    blobs = []
    for i, (sampling, ri) in enumerate((("420", 1), ("444", 3), ("422", 5), ("gray", 2), ("411", 4), ("420", 11))):
        img = W.photo(88, 136, seed=60 + i)
        good = W.encode(img[..., 0], 85, restart=ri) if sampling == "gray" else \
            W.encode(img, 85, sampling, restart=ri, optimize=bool(i & 1))
        start = J.parse(good)["scan_start"]
        for k in range(per_file):
            b = bytearray(good)
            hits = 0
            while hits < 1 + k % 4:
                pos = int(rng.integers(start + 1, len(b) - 3))
                new = int(rng.integers(0, 255)) if k % 2 else b[pos] ^ (1 << int(rng.integers(0, 8)))
                if 0xFF in (b[pos - 1], b[pos], b[pos + 1], new):
                    continue  # leave markers and stuffing alone
                b[pos] = new
                hits += 1
            blobs.append(bytes(b))
    return blobs


def test_bit_errors_inside_restart_segments_decode_like_libjpeg(M):
    """libjpeg restarts by MCU count: a segment that runs out of data early is finished from
    zero bits and padded with zero MCUs, one that has data left over loses it, and the next
    segment starts at its own MCU either way.  Same pixels as cv2, status 1."""
    blobs = _bit_errors_inside_restart_segments()
    wants = [W.reference_decode(b) for b in blobs]
    batch = M.decode_jpeg_batch_cuda(blobs)
    torch.cuda.synchronize()
    status = batch.status.cpu().tolist()
    # (Huffman codes resynchronise: most bit errors leave the block count of their segment intact)
    assert set(status) <= {0, 1} and sum(status) >= 5, status
    for n, (got, want) in enumerate(zip(batch.frames, wants)):
        assert np.array_equal(got.cpu().numpy(), want), n


def test_scan_fuzz_contract(M):
    """scripts/jpeg_scan_fuzz.py: eleven classes of damage to the entropy-coded bytes and to the
    restart markers.  Across all of them: cv2 returns None exactly when status bit 1 is set, and
    every file cv2 decodes comes back with cv2's pixels, flagged (status 1) when damaged."""
    import importlib.util
    from pathlib import Path

    spec = importlib.util.spec_from_file_location(
        "jpeg_scan_fuzz", Path(__file__).resolve().parent.parent / "scripts" / "jpeg_scan_fuzz.py")
    fuzz = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(fuzz)
    table = fuzz.run(per_class=60, seed=5, verbose=False)
    assert set(table) == set(fuzz.KINDS)
    for kind, row in table.items():
        assert row["verdict_differs"] == 0 and row["differs_unflagged"] == 0, (kind, row)
        # libjpeg's whole damage handling is reproduced: concealment of short segments, restart
        # by MCU count, resynchronisation by marker number (jpeg_resync_to_restart)
        assert row["same_pixels"] == row["cv2_decodes"], (kind, row)
    assert table["cut"]["refused"] == 60 and table["cut+eoi"]["same_pixels"] == 60
    assert sum(row["flagged"] for row in table.values()) > 300


def _strip_dht(data):
    out, i = bytearray(data[:2]), 2
    while True:
        m, L = data[i + 1], (data[i + 2] << 8) | data[i + 3]
        if m == 0xDA:
            return bytes(out + data[i:])
        if m != 0xC4:
            out += data[i:i + 2 + L]
        i += 2 + L


def test_motion_jpeg_frames_without_huffman_tables(M):
    """No DHT segment: the standard tables of T.81 Annex K apply (libjpeg's std_huff_tables)."""
    blobs, wants = [], []
    for i, sampling in enumerate(("420", "422", "444")):
        data = W.encode(W.photo(120 + 8 * i, 200, seed=70 + i), 75, sampling, restart=4 * i)
        bare = _strip_dht(data)
        assert len(bare) == len(data) - 4 * 2 - (31 + 181 + 31 + 181)
        blobs.append(bare)
        wants.append(W.reference_decode(bare))
        assert np.array_equal(wants[-1], W.reference_decode(data))
    gray = _strip_dht(W.encode(W.photo(90, 130, seed=77)[..., 0], 80))
    blobs.append(gray)
    wants.append(W.reference_decode(gray))
    batch = M.decode_jpeg_batch_cuda(blobs)
    torch.cuda.synchronize()
    assert batch.status.cpu().tolist() == [0] * len(blobs)
    for got, want in zip(batch.frames, wants):
        assert np.array_equal(got.cpu().numpy(), want)


def test_out_of_range_coefficients_wrap_like_the_simd_idct(M):
    """Quantisation tables far beyond what the coefficients were coded for: the dequantised values
    leave 16 bits, and what comes out is defined by libjpeg-turbo's SIMD IDCT (16-bit lanes that
    wrap, saturating packs), which is what OpenCV runs.  Every file must come back with cv2's
    pixels, and — the entropy decode being untouched — flagged (status 1) exactly when the oracle's
    twin of the range check finds a block that left the 16-bit range."""
    rng = np.random.default_rng(5)
    blobs = []
    for i, sampling in enumerate(("420", "444", "422", "411")):
        base = W.encode(W.photo(72, 104, seed=80 + i), 30 + 20 * i, sampling)
        wide = W.with_16bit_quant_tables(base)
        for variant in range(4):
            b = bytearray(wide if variant >= 2 else base)
            i0 = 2
            while b[i0 + 1] != 0xDA:
                L = (b[i0 + 2] << 8) | b[i0 + 3]
                if b[i0 + 1] == 0xDB:
                    lo, hi = i0 + 5, i0 + 2 + L  # one table per segment: values follow the Pq/Tq byte
                    assert hi - lo == (128 if variant >= 2 else 64)
                    vals = rng.integers(0, 256, hi - lo, dtype=np.uint8)
                    if variant % 2 == 0:
                        vals[:] = 0xFF
                    b[lo:hi] = vals.tobytes()
                i0 += 2 + L
            blobs.append(bytes(b))
    wants = [W.reference_decode(b) for b in blobs]
    batch = M.decode_jpeg_batch_cuda(blobs)
    torch.cuda.synchronize()
    expect = []
    for b in blobs:
        info = J.parse(b)
        expect.append(int(any(J.idct_left_16bit_range(cf, info["qt"][c["tq"]]).any()
                              for c, cf in zip(info["comps"], J.decode_coefficients(b, info)))))
    assert batch.status.cpu().tolist() == expect and sum(expect) >= 8
    for got, want in zip(batch.frames, wants):
        assert np.array_equal(got.cpu().numpy(), want)
    assert len(blobs) == 16


def test_damaged_headers_decode_like_cv2(M):
    """scripts/jpeg_header_fuzz.py mutants (random damage to tables and parameters before the
    scan).  Verdicts: the library refuses exactly the files cv2 refuses (or calls them
    unsupported); pixels: whenever the device reports status 0 they are cv2's — garbage
    coefficients from damaged Huffman or quantisation tables included."""
    import importlib.util
    import random
    from pathlib import Path

    spec = importlib.util.spec_from_file_location(
        "jpeg_header_fuzz", Path(__file__).resolve().parent.parent / "scripts" / "jpeg_header_fuzz.py")
    fuzz = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(fuzz)
    rng = random.Random(23)
    files = [(f, M.parse_jpeg(f).scan_offset) for f in fuzz.bases()]
    blobs, wants = [], []
    for _ in range(1500):
        base, scan = rng.choice(files)
        data = fuzz.mutate(rng, base, scan)
        ref = cv2.imdecode(np.frombuffer(data, np.uint8), cv2.IMREAD_COLOR)
        try:
            M.parse_jpeg(data)
        except M.UnsupportedJpeg:
            continue
        except ValueError:
            assert ref is None
            continue
        assert ref is not None
        blobs.append(data)
        wants.append(ref[..., ::-1])
    assert len(blobs) > 700
    exact = 0
    for lo in range(0, len(blobs), 256):
        batch = M.decode_jpeg_batch_cuda(blobs[lo:lo + 256])
        torch.cuda.synchronize()
        status = batch.status.cpu().tolist()
        for st, got, want in zip(status, batch.frames, wants[lo:lo + 256]):
            assert tuple(got.shape) == want.shape
            if st == 0:
                assert np.array_equal(got.cpu().numpy(), want)
                exact += 1
    assert exact > 600


def test_two_decoders_on_two_streams(M):
    """Two JpegDecoder objects, each on its own stream, launched back to back so that their
    kernels (and the Huffman kernel's look-back chains) overlap on the device."""
    a = [W.encode(W.photo(300, 400, seed=40 + i), 90, "420") for i in range(12)]
    b = [W.encode(W.photo(200, 520, seed=60 + i), 60, "444", restart=5) for i in range(12)]
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    d1, d2 = M.JpegDecoder("cuda"), M.JpegDecoder("cuda")
    outs = []
    for _ in range(3):
        with torch.cuda.stream(s1):
            o1 = d1.decode(a)
        with torch.cuda.stream(s2):
            o2 = d2.decode(b)
        outs.append((o1, o2))
    torch.cuda.synchronize()
    for o1, o2 in outs[-1:]:
        assert o1.status.cpu().tolist() == [0] * 12 and o2.status.cpu().tolist() == [0] * 12
        for f, data in zip(o1.frames + o2.frames, a + b):
            assert np.array_equal(f.cpu().numpy(), W.reference_decode(data))
    for o1, o2 in outs[:-1]:  # frames of earlier calls are the callers' own tensors: still intact
        assert np.array_equal(o1.frames[0].cpu().numpy(), W.reference_decode(a[0]))
        assert np.array_equal(o2.frames[5].cpu().numpy(), W.reference_decode(b[5]))


def test_out_of_range_idct_block_is_flagged_not_clean(M):
    """tests/golden/jpeg_idct_overflow_case.jpg: a damaged file whose luma block (23, 55) drives the
    16-bit lanes of libjpeg-turbo's SIMD IDCT over their range; cv2 decodes it (8 pixels of that
    block differ from the modelled wrap-arounds).  Such a frame must not come back as clean:
    status bit 0 is set, the oracle's twin of the check names exactly that block, and every
    other pixel is cv2's."""
    import cv2
    from pathlib import Path
    data = (Path(__file__).parent / "golden" / "jpeg_idct_overflow_case.jpg").read_bytes()
    want = cv2.cvtColor(cv2.imdecode(np.frombuffer(data, np.uint8), cv2.IMREAD_COLOR), cv2.COLOR_BGR2RGB)
    info = J.parse(data)
    coefs = J.decode_coefficients(data, info)
    flagged = [np.argwhere(J.idct_left_16bit_range(cf, info["qt"][c["tq"]])).tolist()
               for c, cf in zip(info["comps"], coefs)]
    assert flagged == [[[23, 55]], [], []]
    clean = W.encode(W.photo(600, 800, seed=3), 90, "420")
    batch = M.decode_jpeg_batch_cuda([clean, data, clean])
    torch.cuda.synchronize()
    assert batch.status.cpu().tolist() == [0, 1, 0]  # flagged as damaged; the neighbours in the batch are not
    got = batch.frames[1].cpu().numpy()
    differs = np.argwhere((got != want).any(axis=-1))
    assert len(differs) == 0 or (differs.min(0) >= [23 * 8 - 1, 55 * 8 - 1]).all() and (differs.max(0) <= [23 * 8 + 8, 55 * 8 + 8]).all()
