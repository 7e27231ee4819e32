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
