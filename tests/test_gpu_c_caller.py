"""The C-ABI really is a C ABI: a plain-C program (tests/c_caller/c_caller.c — cudaMalloc
and raw pointers, no Python, no torch) drives K1 -> K2+K3 -> K4 and its outputs are compared
bit for bit with the oracle.  This is the shape of binding a non-Python host (cgo, JNI,
N-API) would make."""

import shutil
import subprocess
from pathlib import Path

import numpy as np
import pytest

from oracle import pipeline as O
from workloads import synth

pytestmark = pytest.mark.gpu
REPO = Path(__file__).resolve().parent.parent


def test_plain_c_caller(tmp_path):
    gcc = shutil.which("gcc")
    cuda = Path("/usr/local/cuda")
    if gcc is None or not (cuda / "include" / "cuda_runtime.h").exists():
        pytest.skip("gcc or the CUDA headers are not available")
    lib_dir = REPO / "inference-arena_b200" / "lib"
    exe = tmp_path / "c_caller"
    subprocess.run([gcc, "-O1", "-o", str(exe), str(REPO / "tests" / "c_caller" / "c_caller.c"),
                    "-I", str(REPO / "include"), "-I", str(cuda / "include"), "-L", str(lib_dir),
                    "-L", str(cuda / "lib64"), "-lia_b200", "-lcudart", f"-Wl,-rpath,{lib_dir}",
                    f"-Wl,-rpath,{cuda / 'lib64'}"], check=True)
    n, h, w, anchors, max_det = 3, 480, 640, 25200, 32
    frames, heads = synth.frames(n, h, w, seed=80), synth.head_v5(n, seed=81)
    frames.tofile(tmp_path / "frames.bin")
    heads.tofile(tmp_path / "heads.bin")
    out = subprocess.run([str(exe), str(tmp_path), str(n), str(h), str(w), str(anchors), str(max_det)],
                         check=True, capture_output=True, text=True).stdout
    assert "sm_100a" in out and "scale=1 " in out

    tensor = np.fromfile(tmp_path / "tensor.bin", np.float32).reshape(n, 3, 640, 640)
    dets = np.fromfile(tmp_path / "dets.bin", np.float32).reshape(n, max_det, 6)
    keep = np.fromfile(tmp_path / "keep.bin", np.int32).reshape(n, max_det)
    count = np.fromfile(tmp_path / "count.bin", np.int32)
    offs = np.fromfile(tmp_path / "offs.bin", np.int32)
    crops = np.fromfile(tmp_path / "crops.bin", np.float32).reshape(-1, 3, 224, 224)
    for i in range(n):
        want_t, want_d, want_k, _, want_c = O.run_image(frames[i], heads[i], 0.25, 0.45)
        k = int(count[i])
        assert k == len(want_k) and keep[i, :k].tolist() == want_k
        assert np.array_equal(tensor[i : i + 1], want_t) and np.array_equal(dets[i, :k], want_d)
        assert np.array_equal(crops[offs[i] : offs[i + 1]], want_c)


def _compile(tmp_path, source: str):
    gcc = shutil.which("gcc")
    cuda = Path("/usr/local/cuda")
    if gcc is None or not (cuda / "include" / "cuda_runtime.h").exists():
        pytest.skip("gcc or the CUDA headers are not available")
    lib_dir = REPO / "inference-arena_b200" / "lib"
    exe = tmp_path / Path(source).stem
    subprocess.run([gcc, "-O1", "-o", str(exe), str(REPO / "tests" / "c_caller" / source),
                    "-I", str(REPO / "include"), "-I", str(cuda / "include"), "-L", str(lib_dir),
                    "-L", str(cuda / "lib64"), "-lia_b200", "-lcudart", f"-Wl,-rpath,{lib_dir}",
                    f"-Wl,-rpath,{cuda / 'lib64'}"], check=True)
    return exe


def test_plain_c_caller_jpeg_and_topk(tmp_path, golden):
    """The N3 / N4 entries from C: a committed JPEG fixture (pixels known from the reference's
    load_image_from_bytes) and seeded logits (against the oracle)."""
    exe = _compile(tmp_path, "c_jpeg_caller.c")
    (tmp_path / "image.jpg").write_bytes(golden["jpeg7_bytes"].tobytes())
    rows, classes, k = 12, 1000, 5
    logits = synth.logits(rows, seed=90, classes=classes)
    logits.tofile(tmp_path / "logits.bin")
    out = subprocess.run([str(exe), str(tmp_path), str(rows), str(classes), str(k)], check=True,
                         capture_output=True, text=True).stdout
    want = golden["jpeg7_rgb"]
    assert f"jpeg {want.shape[1]}x{want.shape[0]} comps=3 status=0" in out
    rgb = np.fromfile(tmp_path / "rgb.bin", np.uint8).reshape(want.shape)
    assert np.array_equal(rgb, want)
    ids = np.fromfile(tmp_path / "ids.bin", np.int32).reshape(rows, k)
    probs = np.fromfile(tmp_path / "probs.bin", np.float32).reshape(rows, k)
    for r in range(rows):
        want_ids, want_probs = O.softmax_topk(logits[r], k)
        assert ids[r].tolist() == want_ids.tolist()
        np.testing.assert_allclose(probs[r], want_probs, rtol=1e-5, atol=0)
