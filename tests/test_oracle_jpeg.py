"""Pin oracle/jpeg.py (the restated libjpeg-turbo decoder) against cv2.imdecode — the call
the reference makes in load_image_from_bytes (transforms.py:102) — and check the library's
host-side header parser (ia_jpeg_parse, no CUDA involved) against the oracle's."""

import ctypes

import numpy as np
import pytest

from oracle import jpeg as J
from workloads import jpeg_cases as W

cv2 = pytest.importorskip("cv2")


@pytest.mark.parametrize("sampling", W.SAMPLINGS)
@pytest.mark.parametrize("shape", [(16, 16), (48, 64), (33, 17), (17, 33), (1, 1), (2, 3), (8, 5), (5, 4), (64, 6)])
def test_matches_cv2(shape, sampling):
    for q, kw in ((5, {}), (50, {"restart": 3}), (90, {"optimize": True}), (100, {})):
        data = W.encode(W.photo(*shape, seed=shape[0] * shape[1] + q), q, sampling, **kw)
        assert np.array_equal(J.load_image_from_bytes(data), W.reference_decode(data)), (shape, sampling, q, kw)


def test_grayscale_and_wide():
    for shape in ((20, 30), (7, 9), (31, 250)):
        data = W.encode(W.photo(*shape, seed=3)[..., 0], 80)
        assert np.array_equal(J.load_image_from_bytes(data), W.reference_decode(data))
    data = W.encode(W.photo(31, 250, seed=4), 75, "420", restart=1)
    assert np.array_equal(J.load_image_from_bytes(data), W.reference_decode(data))


@pytest.mark.parametrize("orientation", range(1, 9))
def test_exif_orientation(orientation):
    base = W.encode(W.photo(24, 40, seed=9), 85, "420")
    for big in (False, True):
        data = W.with_exif_orientation(base, orientation, big_endian=big)
        want = W.reference_decode(data)
        got = J.load_image_from_bytes(data)
        assert got.shape == want.shape and np.array_equal(got, want)
    if orientation >= 5:
        assert want.shape[:2] == (40, 24)


def test_reference_fixtures(golden):
    """Files and pixels committed by tests/golden/make_golden.py: the unmodified reference's
    load_image_from_bytes (transforms.py:77-110) decoded these bytes to these arrays."""
    cases = golden.meta["jpeg"]
    assert len(cases) >= 8
    for i, case in enumerate(cases):
        data = golden[f"jpeg{i}_bytes"].tobytes()
        got = J.load_image_from_bytes(data)
        assert list(got.shape) == case["shape"] and got.dtype == np.uint8 and got.flags.c_contiguous
        assert np.array_equal(got, golden[f"jpeg{i}_rgb"]), case


def test_16bit_quantisation_tables(jpeg_mod):
    """DQT with Pq = 1: same pixels from cv2, the oracle and the same tables from the host parser."""
    data = W.encode(W.photo(40, 56, seed=1), 60, "420", optimize=True)
    wide = W.with_16bit_quant_tables(data)
    assert len(wide) == len(data) + 128
    want = W.reference_decode(wide)
    assert np.array_equal(want, W.reference_decode(data))
    assert np.array_equal(J.load_image_from_bytes(wide), want)
    a, b = jpeg_mod.parse_jpeg(data), jpeg_mod.parse_jpeg(wide)
    for c in range(3):
        assert list(a.quant[c]) == list(b.quant[c])
    assert b.scan_offset == a.scan_offset + 128


def test_colour_space_guess_follows_libjpeg(jpeg_mod):
    """jdapimin.c default_decompress_parms: without a JFIF or Adobe marker three components with
    ids 1,2,3 are still YCbCr; ids 'R','G','B' mean stored RGB (cv2 then returns other pixels), which
    the device decoder refuses instead of converting wrongly."""
    data = W.encode(W.photo(40, 56, seed=1), 80, "444")
    assert data[2:4] == b"\xff\xe0"
    app0 = (data[4] << 8) | data[5]
    bare = data[:2] + data[4 + app0:]  # no JFIF segment
    want = W.reference_decode(data)
    assert np.array_equal(W.reference_decode(bare), want)
    assert np.array_equal(J.load_image_from_bytes(bare), want)
    assert jpeg_mod.parse_jpeg(bare).ncomp == 3
    rgb = bytearray(bare)
    sof, sos = rgb.index(b"\xff\xc0"), rgb.index(b"\xff\xda")
    for c, ch in enumerate(b"RGB"):
        rgb[sof + 10 + 3 * c] = ch
        rgb[sos + 5 + 2 * c] = ch
    rgb = bytes(rgb)
    assert not np.array_equal(W.reference_decode(rgb), want)
    with pytest.raises(J.UnsupportedJpeg):
        J.parse(rgb)
    with pytest.raises(jpeg_mod.UnsupportedJpeg):
        jpeg_mod.parse_jpeg(rgb)
    odd = bytearray(data)  # luma subsampled below chroma: legal, never produced; refused
    sof = odd.index(b"\xff\xc0")
    odd[sof + 11], odd[sof + 14], odd[sof + 17] = 0x11, 0x22, 0x22
    with pytest.raises(jpeg_mod.UnsupportedJpeg):
        jpeg_mod.parse_jpeg(bytes(odd))


def _strip_segments(data: bytes, marker: int, which=None) -> bytes:
    """Drop the `which`-th (all when None) segments with this marker from the headers."""
    out, i, k = bytearray(data[:2]), 2, 0
    while True:
        m, L = data[i + 1], (data[i + 2] << 8) | data[i + 3]
        if m == 0xDA:
            return bytes(out + data[i:])
        if m != marker or (which is not None and k not in which):
            out += data[i:i + 2 + L]
        k += m == marker
        i += 2 + L


@pytest.mark.parametrize("sampling", ["420", "444", "gray"])
def test_standard_tables_stand_in_for_missing_dht(jpeg_mod, sampling):
    """Motion-JPEG frames carry no DHT: libjpeg installs the Annex K tables for ids 0 and 1
    (jdhuff.c jinit_huff_decoder -> std_huff_tables), so cv2 decodes them; so must we."""
    img = W.photo(40, 56, seed=1)
    data = W.encode(img[..., 0], 80) if sampling == "gray" else W.encode(img, 80, sampling)
    want = W.reference_decode(data)
    for which in (None, [0], [1, 3] if sampling != "gray" else [1]):
        bare = _strip_segments(data, 0xC4, which)
        assert len(bare) < len(data)
        assert np.array_equal(W.reference_decode(bare), want)
        assert np.array_equal(J.load_image_from_bytes(bare), want)
        a, b = jpeg_mod.parse_jpeg(data), jpeg_mod.parse_jpeg(bare)
        for t in range(2 if sampling != "gray" else 1):
            assert bytes(a.dc_counts[t]) == bytes(b.dc_counts[t]) and bytes(a.dc_vals[t]) == bytes(b.dc_vals[t])
            assert bytes(a.ac_counts[t]) == bytes(b.ac_counts[t]) and bytes(a.ac_vals[t]) == bytes(b.ac_vals[t])
    # an id without a default stays an error
    sos = data.index(b"\xff\xda")
    odd = bytearray(data)
    odd[sos + 6] = 0x22
    assert cv2.imdecode(np.frombuffer(bytes(odd), np.uint8), cv2.IMREAD_COLOR) is None
    with pytest.raises(ValueError):
        J.parse(bytes(odd))
    with pytest.raises(ValueError):
        jpeg_mod.parse_jpeg(bytes(odd))


def _refused_everywhere(jpeg_mod, data: bytes):
    assert cv2.imdecode(np.frombuffer(data, np.uint8), cv2.IMREAD_COLOR) is None
    with pytest.raises(ValueError):
        J.load_image_from_bytes(data)
    with pytest.raises(ValueError):
        jpeg_mod.parse_jpeg(data)


def _decoded_everywhere(jpeg_mod, data: bytes):
    want = W.reference_decode(data)
    assert np.array_equal(J.load_image_from_bytes(data), want)
    jpeg_mod.parse_jpeg(data)


def test_huffman_tables_libjpeg_refuses(jpeg_mod):
    """jdhuff.c jpeg_make_d_derived_tbl is fatal for codes that reach the all-ones code of their
    length and for DC symbols above 15 — but only for tables the scan uses."""
    data = W.encode(W.photo(40, 56, seed=1), 80, "420")
    dht = data.index(b"\xff\xc4")  # the first one: DC table 0, counts 0 1 5 1 1 1 1 1 1 0...
    assert data[dht + 4] == 0x00
    over = bytearray(data)  # two 1-bit codes, total unchanged
    over[dht + 5], over[dht + 7] = 2, over[dht + 7] - 2
    _refused_everywhere(jpeg_mod, bytes(over))
    full = bytearray(data)  # a complete tree: the last code is all ones
    full[dht + 5:dht + 21] = bytes([1] * 10 + [2] + [0] * 5)
    _refused_everywhere(jpeg_mod, bytes(full))
    sym = bytearray(data)  # DC category 16
    sym[dht + 21 + 3] = 0x10
    _refused_everywhere(jpeg_mod, bytes(sym))
    unused = bytearray(data)  # the same damage in a table no component uses: harmless
    unused[dht + 4] = 0x03
    unused[dht + 21 + 3] = 0x10
    _decoded_everywhere(jpeg_mod, bytes(unused))  # DC table 0 is now the standard one


def test_header_rules_follow_libjpeg(jpeg_mod):
    data = W.encode(W.photo(24, 40, seed=5), 80, "422", restart=2)
    app0 = (data[4] << 8) | data[5]
    dri, sof, sos = data.index(b"\xff\xdd"), data.index(b"\xff\xc0"), data.index(b"\xff\xda")
    dqt = data.index(b"\xff\xdb")

    def edit(pos, *vals):
        b = bytearray(data)
        b[pos:pos + len(vals)] = bytes(vals)
        return bytes(b)

    # OpenCV recognises JPEG by FF D8 FF
    _refused_everywhere(jpeg_mod, edit(2, 0xFB))
    # a second SOI, an unknown marker, a reserved one
    _refused_everywhere(jpeg_mod, data[:2] + b"\xff\xd8" + data[2:])
    _refused_everywhere(jpeg_mod, data[:2] + b"\xff\x02\x00\x02" + data[2:])
    _refused_everywhere(jpeg_mod, data[:2] + b"\xff\xf0\x00\x02" + data[2:])
    # FF 00 and fill bytes between segments are skipped; so is garbage
    _decoded_everywhere(jpeg_mod, data[:2] + b"\xff\x00\x12\x34\xff\xff" + data[3:])
    # parameterless markers in the headers
    _decoded_everywhere(jpeg_mod, data[:2] + b"\xff\x01\xff\xd3" + data[2:])
    # an APPn / COM length below 2 skips nothing
    _decoded_everywhere(jpeg_mod, data[:2] + b"\xff\xfe\x00\x00" + data[2:])
    # exact lengths: DRI 4, SOF 8 + 3n, SOS 6 + 2n
    _refused_everywhere(jpeg_mod, data[:dri] + b"\xff\xdd\x00\x05\x00\x02\x00" + data[dri + 6:])
    _refused_everywhere(jpeg_mod, data[:sof + 2] + b"\x00\x12" + data[sof + 4:sof + 19] + b"\x00" + data[sof + 19:])
    _refused_everywhere(jpeg_mod, data[:sos + 2] + b"\x00\x0d" + data[sos + 4:sos + 14] + b"\x00" + data[sos + 14:])
    # a second SOF, sampling factor 5, a scan component the frame does not have
    _refused_everywhere(jpeg_mod, data[:sos] + data[sof:sof + 19] + data[sos:])
    _refused_everywhere(jpeg_mod, edit(sof + 11, 0x51))
    _refused_everywhere(jpeg_mod, edit(sos + 5, 9))
    # quantisation table id 4; a component naming a table that was never defined
    _refused_everywhere(jpeg_mod, edit(dqt + 4, 0x04))
    _refused_everywhere(jpeg_mod, edit(sof + 12, 3))
    # DAC has no business in a Huffman file but libjpeg checks it
    _decoded_everywhere(jpeg_mod, data[:2] + b"\xff\xcc\x00\x04\x00\x10" + data[2:])
    _refused_everywhere(jpeg_mod, data[:2] + b"\xff\xcc\x00\x04\x00\x01" + data[2:])
    _refused_everywhere(jpeg_mod, data[:2] + b"\xff\xcc\x00\x04\x20\x00" + data[2:])
    # JFIF needs 14 bytes to count (only matters for the colour-space guess)
    assert app0 == 16


def test_header_fuzz_agrees_with_cv2():
    """scripts/jpeg_header_fuzz.py: random damage to the bytes before the scan; cv2, the oracle
    and the library's host parser must give the same verdict, and the same pixels when they
    decode (damaged tables make garbage coefficients: libjpeg-turbo's 16-bit SIMD IDCT lanes,
    its concealment of short streams and its 17-bit invalid codes are all part of that)."""
    import importlib.util
    from pathlib import Path

    spec = importlib.util.spec_from_file_location(
        "jpeg_header_fuzz", Path(__file__).resolve().parent.parent / "scripts" / "jpeg_header_fuzz.py")
    fuzz = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(fuzz)
    stats, findings = fuzz.run(iterations=500, seed=11, verbose=False)
    assert sum(stats.values()) == 500
    assert sum(v for k, v in stats.items() if k[0] == "cv2 ok") > 200
    assert not findings, [(p, k) for p, k, _ in findings]


def test_damaged_scans_are_concealed_like_libjpeg():
    """jdhuff.c decode_mcu: after a request that needed bits behind a marker the MCU in hand is
    finished from zero bits and the rest stays zero until a restart marker is consumed
    (jdmarker.c jpeg_resync_to_restart decides which).  Data that simply ends — no marker —
    makes OpenCV's source manager fail: the reference raises."""
    for sampling, kw in (("420", {}), ("444", {"restart": 2}), ("422", {"optimize": True, "restart": 3})):
        data = W.encode(W.photo(32, 48, seed=3), 85, sampling, **kw)
        scan = J.parse(data)["scan_start"]
        body = len(data) - scan
        decoded = 0
        for cut in (body // 5, body // 2, body - 20):
            for tail in (b"\xff\xd9", b"\xff\xc4\x00\x02", b"\xff\xd0\xff\xd9", b"\xff\xd5"):
                dmg = data[:scan + cut] + tail
                ref = cv2.imdecode(np.frombuffer(dmg, np.uint8), cv2.IMREAD_COLOR)
                if ref is None:  # a restart came due and the search for its marker ran off the end
                    assert kw.get("restart")
                    with pytest.raises(ValueError):
                        J.load_image_from_bytes(dmg)
                    continue
                decoded += 1
                report = {}
                info = J.parse(dmg)
                J.decode_coefficients(dmg, info, report)
                assert report["damaged"]
                assert np.array_equal(J.load_image_from_bytes(dmg), ref[..., ::-1]), (sampling, cut, tail)
            cut_off = data[:scan + cut]
            if b"\xff" in cut_off[scan:].replace(b"\xff\x00", b""):
                continue  # a restart marker is in there: libjpeg stops reading at it
            assert cv2.imdecode(np.frombuffer(cut_off, np.uint8), cv2.IMREAD_COLOR) is None
            with pytest.raises(ValueError):
                J.load_image_from_bytes(cut_off)
        assert decoded >= 3
        # a restart marker with the wrong number
        if kw.get("restart"):
            pos = [i for i in range(scan, len(data) - 1) if data[i] == 0xFF and 0xD0 <= data[i + 1] <= 0xD7]
            for delta in (1, 2, 3, 6, 7):
                dmg = bytearray(data)
                dmg[pos[1] + 1] = 0xD0 + ((dmg[pos[1] + 1] - 0xD0 + delta) & 7)
                ref = cv2.imdecode(np.frombuffer(bytes(dmg), np.uint8), cv2.IMREAD_COLOR)
                assert np.array_equal(J.load_image_from_bytes(bytes(dmg)), ref[..., ::-1]), (sampling, delta)


def test_rejects():
    with pytest.raises(J.UnsupportedJpeg):
        J.parse(W.encode(W.photo(16, 16, 1), 90, "420", progressive=True))
    with pytest.raises(ValueError):
        J.parse(b"\x89PNG\r\n\x1a\n" + b"\0" * 32)
    with pytest.raises(ValueError):
        J.load_image_from_bytes(b"")


# --------------------------------------------------------------- the library's host parser
@pytest.fixture(scope="module")
def jpeg_mod():
    from shared.processing import jpeg
    return jpeg


def test_job_struct_layout(jpeg_mod):
    from shared.processing._native import lib
    assert lib().ia_jpeg_job_size() == ctypes.sizeof(jpeg_mod.JpegJob)


@pytest.mark.parametrize("sampling", W.SAMPLINGS + ("gray",))
def test_host_parser_matches_oracle(jpeg_mod, sampling):
    img = W.photo(45, 70, seed=11)
    data = (W.encode(img[..., 0], 70) if sampling == "gray"
            else W.encode(img, 70, sampling, restart=4, optimize=sampling == "422"))
    data = W.with_exif_orientation(data, 6)
    info = J.parse(data)
    job = jpeg_mod.parse_jpeg(data)
    assert (job.width, job.height, job.ncomp) == (info["width"], info["height"], len(info["comps"]))
    assert (job.out_width, job.out_height, job.orientation) == (info["height"], info["width"], 6)
    assert job.restart_interval == info["restart"] and job.scan_offset == info["scan_start"]
    assert job.file_bytes == len(data)
    for c, comp in enumerate(info["comps"]):
        assert (job.h[c], job.v[c]) == (comp["h"], comp["v"])
        assert list(job.quant[c]) == info["qt"][comp["tq"]].tolist()
        counts, vals = info["dc"][comp["td"]]
        assert list(job.dc_counts[job.comp_dc[c]]) == counts
        assert list(job.dc_vals[job.comp_dc[c]])[: len(vals)] == vals
        counts, vals = info["ac"][comp["ta"]]
        assert list(job.ac_counts[job.comp_ac[c]]) == counts
        assert list(job.ac_vals[job.comp_ac[c]])[: len(vals)] == vals
    hmax, vmax = max(c["h"] for c in info["comps"]), max(c["v"] for c in info["comps"])
    assert (job.mcux, job.mcuy) == (-(-70 // (8 * hmax)), -(-45 // (8 * vmax)))
    bpm = sum(c["h"] * c["v"] for c in info["comps"])
    assert job.blocks_per_mcu == bpm and job.total_blocks == bpm * job.mcux * job.mcuy
    assert job.num_segments == (-(-(job.mcux * job.mcuy) // 4) if info["restart"] else 1)


def test_host_parser_errors(jpeg_mod):
    with pytest.raises(ValueError, match="Failed to decode"):
        jpeg_mod.parse_jpeg(b"")
    with pytest.raises(ValueError, match="Failed to decode"):
        jpeg_mod.parse_jpeg(b"not a jpeg at all, just text")
    with pytest.raises(ValueError, match="Failed to decode"):
        jpeg_mod.parse_jpeg(W.encode(W.photo(16, 16, 1), 90)[:60])  # truncated headers
    with pytest.raises(jpeg_mod.UnsupportedJpeg):
        jpeg_mod.parse_jpeg(W.encode(W.photo(16, 16, 1), 90, "420", progressive=True))
    assert issubclass(jpeg_mod.UnsupportedJpeg, ValueError)


def test_no_cpu_fallback(jpeg_mod):
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        jpeg_mod.decode_jpeg_batch_cuda([W.encode(W.photo(16, 16, 1), 90)])


def test_idct_range_check_flags_only_the_damaged_block():
    """oracle.jpeg.idct_left_16bit_range: no block of a valid encoding leaves the 16-bit range of the
    SIMD IDCT (there the SIMD code is plain jidctint.c arithmetic), the committed damaged file has
    exactly one that does."""
    from pathlib import Path
    import cv2
    rng = np.random.default_rng(0)
    noise = rng.integers(0, 256, (96, 128, 3), dtype=np.uint8)
    blobs = [W.encode(W.photo(120, 160, seed=q), q, s) for q in (5, 60, 100) for s in ("444", "420")]
    blobs += [cv2.imencode(".jpg", noise, [cv2.IMWRITE_JPEG_QUALITY, q])[1].tobytes() for q in (1, 100)]
    for blob in blobs:
        info = J.parse(blob)
        for c, cf in zip(info["comps"], J.decode_coefficients(blob, info)):
            assert not J.idct_left_16bit_range(cf, info["qt"][c["tq"]]).any()
    data = (Path(__file__).parent / "golden" / "jpeg_idct_overflow_case.jpg").read_bytes()
    info = J.parse(data)
    hits = [int(J.idct_left_16bit_range(cf, info["qt"][c["tq"]]).sum())
            for c, cf in zip(info["comps"], J.decode_coefficients(data, info))]
    assert hits == [1, 0, 0]
