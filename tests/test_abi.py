"""The C-ABI library must load and export every symbol include/ia_b200.h declares
(no compute calls here: this runs without a GPU)."""

import ctypes
import re
from pathlib import Path

import pytest

REPO = Path(__file__).resolve().parent.parent
HEADER = REPO / "include" / "ia_b200.h"
LIB = REPO / "inference-arena_b200" / "lib" / "libia_b200.so"


def declared_functions() -> list[str]:
    text = re.sub(r"/\*.*?\*/", "", HEADER.read_text(), flags=re.S)
    return sorted(set(re.findall(r"\b(ia_[a-z0-9_]+)\s*\(", text)))


@pytest.fixture(scope="module")
def library():
    if not LIB.exists():
        import __graft_entry__
        __graft_entry__.build()
    return ctypes.CDLL(str(LIB))


def test_header_declares_functions():
    names = declared_functions()
    assert "ia_yolo_preprocess" in names and "ia_decode_nms" in names and len(names) >= 12


def test_library_exports_every_declared_symbol(library):
    missing = [n for n in declared_functions() if not hasattr(library, n)]
    assert not missing, missing


def test_binding_table_matches_header():
    from shared.processing import _native

    assert sorted(_native.exported_symbols()) == declared_functions()


def test_host_only_entry_points(library):
    """ia_letterbox_geometry is pure host arithmetic (transforms.py:150-175)."""
    from shared.processing import _native

    library.ia_version.restype = ctypes.c_char_p
    assert b"sm_100a" in library.ia_version()
    g = _native.letterbox_geometry(1080, 1920, 640)
    assert (g.scale, g.new_w, g.new_h, g.pad_w, g.pad_h) == (0.3333333333333333, 640, 360, 0, 140)
    g = _native.letterbox_geometry(480, 640, 640)
    assert (g.scale, g.new_w, g.new_h, g.pad_w, g.pad_h) == (1.0, 640, 480, 0, 80)
    from oracle.pipeline import letterbox_geometry as ref
    for h, w, t in [(500, 375, 640), (333, 500, 640), (1281, 1279, 640), (37, 640, 320), (720, 1280, 1280)]:
        g, r = _native.letterbox_geometry(h, w, t), ref(h, w, t)
        assert (g.scale, g.new_w, g.new_h, g.pad_w, g.pad_h) == (r.scale, r.new_w, r.new_h, r.pad_w, r.pad_h)
    assert library.ia_decode_nms_workspace_bytes(0, 0, 0) == 0


def test_no_cpu_fallback_without_cuda():
    """The product path must fail loudly when it cannot reach the GPU."""
    import numpy as np
    import torch

    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    from shared.processing import YOLOPreprocessor

    with pytest.raises(RuntimeError, match="CUDA"):
        YOLOPreprocessor()(np.zeros((8, 8, 3), np.uint8))
