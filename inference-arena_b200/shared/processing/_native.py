"""ctypes binding of libia_b200.so — the C-ABI declared in ``include/ia_b200.h``.

The product path has no CPU fallback: if the library is missing, or the process has no
CUDA device, the first call raises.
"""

from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, c_char_p, c_double, c_float, c_int, c_int32, c_size_t, c_void_p
from pathlib import Path

PACKAGE_ROOT = Path(__file__).resolve().parents[2]  # .../inference-arena_b200
LIB_PATH = Path(os.environ.get("IA_B200_LIB", PACKAGE_ROOT / "lib" / "libia_b200.so"))

IA_HEAD_LAYOUT_V5 = 0
IA_HEAD_LAYOUT_V8 = 1
IA_F32_RESIZE_OPENCV = 0
IA_F32_RESIZE_IPP = 1


class LetterboxGeom(ctypes.Structure):
    """``ia_letterbox_geom`` (include/ia_b200.h)."""

    _fields_ = [("src_h", c_int32), ("src_w", c_int32), ("target", c_int32),
                ("new_h", c_int32), ("new_w", c_int32), ("pad_h", c_int32), ("pad_w", c_int32),
                ("scale", c_double)]


class FrameDesc(ctypes.Structure):
    """``ia_frame_desc`` (include/ia_b200.h): one frame of a ragged batch."""

    _fields_ = [("data_dev", c_void_p), ("h", c_int32), ("w", c_int32)]


F3 = c_float * 3
U3 = ctypes.c_uint8 * 3

_SIGNATURES = {
    "ia_version": (c_char_p, []),
    "ia_error_string": (c_char_p, [c_int]),
    "ia_letterbox_geometry": (c_int, [c_int, c_int, c_int, POINTER(LetterboxGeom)]),
    "ia_yolo_preprocess": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_float, c_void_p, c_void_p]),
    "ia_yolo_preprocess_ragged": (c_int, [c_void_p, c_int, c_int, c_float, c_void_p, c_void_p]),
    "ia_letterbox_u8": (c_int, [c_void_p, c_int, c_int, c_int, c_int, U3, c_void_p, c_void_p]),
    "ia_decode_nms_workspace_bytes": (c_size_t, [c_int, c_int, c_int]),
    "ia_decode_nms": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_float, c_float, c_int, c_int,
                              c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "ia_decode_nms_static": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_float, c_float, c_int,
                                     c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "ia_nms": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_float, c_float, c_int, c_void_p,
                       c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "ia_scale_boxes": (c_int, [c_void_p, c_int, c_int, c_float, c_int, c_int, c_int, c_int,
                               c_void_p, c_void_p]),
    "ia_scale_boxes_f64": (c_int, [c_void_p, c_int, c_int, c_double, c_int, c_int, c_int, c_int,
                                   c_void_p, c_void_p]),
    "ia_crop_resize_normalize": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_int,
                                         c_float, c_int, c_int, c_int, F3, F3, c_void_p, c_int,
                                         c_void_p, c_void_p, c_void_p]),
    "ia_crop_resize_normalize_ragged": (c_int, [c_void_p, c_int, c_int, c_void_p, c_void_p, c_int, c_int,
                                                c_int, F3, F3, c_void_p, c_int, c_void_p, c_void_p, c_void_p]),
    "ia_resize_normalize_rois": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_int, c_int, F3,
                                         F3, c_void_p, c_void_p]),
    "ia_extract_crop_rects": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    "ia_f32_scratch_bytes": (c_size_t, [c_int]),
    "ia_resize_normalize_f32_as": (c_int, [c_void_p, c_int, c_int, c_int, c_int, F3, F3, c_void_p, c_void_p, c_size_t,
                                           c_void_p]),
    "ia_resize_normalize_f32": (c_int, [c_void_p, c_int, c_int, c_int, F3, F3, c_void_p, c_void_p, c_size_t,
                                        c_void_p]),
    "ia_imagenet_normalize_f32": (c_int, [c_void_p, c_size_t, F3, F3, c_void_p, c_void_p, c_size_t, c_void_p]),
    "ia_imagenet_normalize_u8": (c_int, [c_void_p, c_size_t, F3, F3, c_void_p, c_void_p]),
    "ia_jpeg_job_size": (c_size_t, []),
    "ia_jpeg_parse": (c_int, [c_void_p, c_size_t, c_void_p]),
    "ia_jpeg_workspace_bytes": (c_size_t, [c_void_p, c_int]),
    "ia_jpeg_plan": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_size_t]),
    "ia_jpeg_decode": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    "ia_classify_topk": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                 c_void_p]),
}

_lib = None


def exported_symbols() -> list[str]:
    """Every symbol include/ia_b200.h declares (checked by tests/test_abi.py)."""
    return list(_SIGNATURES)


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise RuntimeError(
                f"libia_b200.so not found at {LIB_PATH}: build it with "
                f"`python inference-arena_b200/build.py` (there is no CPU fallback)")
        handle = ctypes.CDLL(str(LIB_PATH))
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def check(code: int, what: str) -> None:
    if code != 0:
        msg = lib().ia_error_string(code).decode()
        raise RuntimeError(f"{what} failed: {msg} (code {code})")


def letterbox_geometry(src_h: int, src_w: int, target: int) -> LetterboxGeom:
    g = LetterboxGeom()
    check(lib().ia_letterbox_geometry(int(src_h), int(src_w), int(target), ctypes.byref(g)),
          "ia_letterbox_geometry")
    return g
