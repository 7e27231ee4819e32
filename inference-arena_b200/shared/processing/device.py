"""Batched device API of the hot path ("additive" rows).

torch is plumbing only — device memory, streams, host<->device copies; every arithmetic
step is a kernel of libia_b200.so called through the C-ABI (``include/ia_b200.h``) with raw
device pointers.  Inputs and outputs are CUDA tensors, so the stages chain on the device
with no host round trip:

    tensor            = preprocess_batch_cuda(frames)                    # K1
    dets, keep, count, status = decode_nms_cuda(head, conf, iou)          # K2 + K3
    crops, offsets, rois = crop_batch_cuda(frames, dets, count, geom)    # K4
"""

from __future__ import annotations

import ctypes
from dataclasses import dataclass

import numpy as np
import torch

from . import _native
from ._native import F3, U3, check, lib


def _require_cuda(t: torch.Tensor, dtype: torch.dtype, what: str) -> torch.Tensor:
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise RuntimeError(f"{what} must be a CUDA tensor (no CPU fallback in this path)")
    if t.dtype != dtype:
        raise ValueError(f"{what} must be {dtype}, got {t.dtype}")
    return t.contiguous()


def _stream_ptr(device: torch.device) -> int:
    return torch.cuda.current_stream(device).cuda_stream


@dataclass(frozen=True)
class LetterboxGeometry:
    """Host scalars of letterbox() (reference transforms.py:150-175)."""

    src_h: int
    src_w: int
    target: int
    new_h: int
    new_w: int
    pad_h: int
    pad_w: int
    scale: float

    @property
    def padding(self) -> tuple[int, int]:
        return (self.pad_w, self.pad_h)


def letterbox_geometry(src_h: int, src_w: int, target: int) -> LetterboxGeometry:
    g = _native.letterbox_geometry(src_h, src_w, target)
    return LetterboxGeometry(g.src_h, g.src_w, g.target, g.new_h, g.new_w, g.pad_h, g.pad_w, g.scale)


# ------------------------------------------------------------------------------------- K1
def preprocess_batch_cuda(frames: torch.Tensor, target: int = 640, divisor: float = 255.0,
                          out: torch.Tensor | None = None) -> torch.Tensor:
    """uint8 [N,H,W,3] (cuda) -> float32 [N,3,target,target]; YOLOPreprocessor.preprocess
    (reference yolo_preprocess.py:128-174) for every frame of the batch, one launch."""
    frames = _require_cuda(frames, torch.uint8, "frames")
    if frames.ndim != 4 or frames.shape[3] != 3:
        raise ValueError(f"Expected [N, H, W, 3] frames, got {tuple(frames.shape)}")
    n, h, w, _ = frames.shape
    if out is None:
        out = torch.empty((n, 3, target, target), dtype=torch.float32, device=frames.device)
    else:
        _require_cuda(out, torch.float32, "out")
        if tuple(out.shape) != (n, 3, target, target) or not out.is_contiguous():
            raise ValueError("out has the wrong shape or is not contiguous")
    with torch.cuda.device(frames.device):
        check(lib().ia_yolo_preprocess(frames.data_ptr(), n, h, w, int(target), float(np.float32(divisor)),
                                       out.data_ptr(), _stream_ptr(frames.device)), "ia_yolo_preprocess")
    return out


def letterbox_batch_cuda(frames: torch.Tensor, target: int, color=(114, 114, 114)) -> torch.Tensor:
    """uint8 [N,H,W,3] -> uint8 [N,target,target,3]; letterbox() (transforms.py:118-180)."""
    frames = _require_cuda(frames, torch.uint8, "frames")
    n, h, w, _ = frames.shape
    out = torch.empty((n, target, target, 3), dtype=torch.uint8, device=frames.device)
    with torch.cuda.device(frames.device):
        check(lib().ia_letterbox_u8(frames.data_ptr(), n, h, w, int(target), U3(*[int(c) for c in color]),
                                    out.data_ptr(), _stream_ptr(frames.device)), "ia_letterbox_u8")
    return out


# -------------------------------------------------------------------------------- K2 + K3
NMS_MAX_CLASSES = 1023  # class ids 0..1022: the sort key holds a class in 10 bits (csrc kMaxClasses)
_workspaces: dict[tuple[int, int], torch.Tensor] = {}


def _workspace(device: torch.device, nbytes: int) -> torch.Tensor:
    """Default candidate-list scratch: one buffer per (device, current torch stream), so calls
    issued on different streams (or from threads with different current streams) never share
    the counters and key lists.  Calls on ONE stream reuse it in stream order."""
    index = device.index if device.index is not None else torch.cuda.current_device()
    key = (index, int(torch.cuda.current_stream(index).cuda_stream))
    ws = _workspaces.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.empty(max(nbytes, 1 << 20), dtype=torch.uint8, device=device)
        _workspaces[key] = ws
    return ws


def decode_nms_cuda(head: torch.Tensor, conf_threshold: float, iou_threshold: float,
                    layout: str = "v5", max_candidates: int | None = None, max_det: int = 300,
                    workspace: torch.Tensor | None = None):
    """Head decode + confidence filter + per-class NMS + row assembly for a batch.

    ``layout="v5"``: float32 [N, A, 5+C] (reference curator.py:289-307 decode);
    ``layout="v8"``: float32 [N, 4+C, A] (reference postprocess.py:33-41).
    Returns ``(dets [N,max_det,6], keep [N,max_det] int32, count [N] int32, status [N] int32)``;
    rows beyond ``count`` are unspecified.  ``status != 0`` flags a truncated image
    (bit0: more than ``max_candidates`` anchors passed the filter; bit1: more than
    ``max_det`` detections)."""
    head = _require_cuda(head, torch.float32, "head")
    if head.ndim != 3:
        raise ValueError(f"Expected a 3-D head tensor, got {tuple(head.shape)}")
    if layout == "v5":
        n, anchors, row = head.shape
        classes, code = row - 5, _native.IA_HEAD_LAYOUT_V5
    elif layout == "v8":
        n, row, anchors = head.shape
        classes, code = row - 4, _native.IA_HEAD_LAYOUT_V8
    else:
        raise ValueError(f"unknown head layout {layout!r}")
    if classes < 1:
        raise ValueError(f"head has no class scores: {tuple(head.shape)}")
    cap = anchors if max_candidates is None else int(max_candidates)
    dev = head.device
    dets = torch.empty((n, max_det, 6), dtype=torch.float32, device=dev)
    keep = torch.empty((n, max_det), dtype=torch.int32, device=dev)
    count = torch.empty((n,), dtype=torch.int32, device=dev)
    status = torch.empty((n,), dtype=torch.int32, device=dev)
    need = lib().ia_decode_nms_workspace_bytes(n, anchors, cap)
    ws = workspace if workspace is not None else _workspace(dev, need)
    with torch.cuda.device(dev):
        check(lib().ia_decode_nms(head.data_ptr(), code, n, anchors, classes,
                                  float(np.float32(conf_threshold)), float(np.float32(iou_threshold)),
                                  cap, int(max_det), dets.data_ptr(), keep.data_ptr(), count.data_ptr(),
                                  status.data_ptr(), ws.data_ptr(), ws.numel(), _stream_ptr(dev)),
              "ia_decode_nms")
    return dets, keep, count, status


def nms_cuda(boxes: torch.Tensor, scores: torch.Tensor, class_ids: torch.Tensor,
             conf_threshold: float, iou_threshold: float, max_det: int | None = None):
    """apply_nms (reference postprocess.py:76-160) on decoded arrays of one image.
    Returns ``(keep [max_det] int32, count [1], status [1])`` on the device."""
    boxes = _require_cuda(boxes, torch.float32, "boxes")
    scores = _require_cuda(scores, torch.float32, "scores")
    class_ids = _require_cuda(class_ids, torch.int32, "class_ids")
    m = scores.numel()
    dev = scores.device
    max_det = max(int(max_det if max_det is not None else m), 1)
    keep = torch.empty((max_det,), dtype=torch.int32, device=dev)
    count = torch.zeros((1,), dtype=torch.int32, device=dev)
    status = torch.zeros((1,), dtype=torch.int32, device=dev)
    need = lib().ia_decode_nms_workspace_bytes(1, max(m, 1), max(m, 1)) + max_det * 24
    ws = _workspace(dev, need)
    with torch.cuda.device(dev):
        check(lib().ia_nms(boxes.data_ptr(), scores.data_ptr(), class_ids.data_ptr(), m,
                           float(np.float32(conf_threshold)), float(np.float32(iou_threshold)), max_det,
                           keep.data_ptr(), count.data_ptr(), status.data_ptr(), ws.data_ptr(),
                           ws.numel(), _stream_ptr(dev)), "ia_nms")
    return keep, count, status


# ------------------------------------------------------------------------------------- K4
def scale_boxes_cuda(boxes: torch.Tensor, scale: float, padding, original_shape) -> torch.Tensor:
    """scale_boxes (reference transforms.py:183-230) on a float32 [M, >=4] CUDA tensor."""
    if isinstance(boxes, torch.Tensor) and boxes.dtype == torch.float64:
        boxes = _require_cuda(boxes, torch.float64, "boxes")
        out = torch.empty_like(boxes)
        m, cols = boxes.shape
        with torch.cuda.device(boxes.device):
            check(lib().ia_scale_boxes_f64(boxes.data_ptr(), m, cols, float(scale), int(padding[0]),
                                           int(padding[1]), int(original_shape[0]), int(original_shape[1]),
                                           out.data_ptr(), _stream_ptr(boxes.device)), "ia_scale_boxes_f64")
        return out
    boxes = _require_cuda(boxes, torch.float32, "boxes")
    out = torch.empty_like(boxes)
    m, cols = boxes.shape
    with torch.cuda.device(boxes.device):
        check(lib().ia_scale_boxes(boxes.data_ptr(), m, cols, float(np.float32(scale)), int(padding[0]),
                                   int(padding[1]), int(original_shape[0]), int(original_shape[1]),
                                   out.data_ptr(), _stream_ptr(boxes.device)), "ia_scale_boxes")
    return out


def extract_crop_rects_cuda(boxes: torch.Tensor, src_h: int, src_w: int, image_index: int = 0) -> torch.Tensor:
    """extract_crop's integer rectangles (reference mobilenet_preprocess.py:256-267) for a
    float32 [M, >=4] CUDA tensor of original-frame boxes -> int32 [M, 8] ROI records."""
    boxes = _require_cuda(boxes, torch.float32, "boxes")
    m, cols = boxes.shape
    rois = torch.empty((m, 8), dtype=torch.int32, device=boxes.device)
    with torch.cuda.device(boxes.device):
        check(lib().ia_extract_crop_rects(boxes.data_ptr(), m, cols, int(image_index), int(src_h),
                                          int(src_w), rois.data_ptr(), _stream_ptr(boxes.device)),
              "ia_extract_crop_rects")
    return rois


def crop_batch_cuda(frames: torch.Tensor, dets: torch.Tensor, count: torch.Tensor,
                    geom: LetterboxGeometry, out_size: int = 224, mean=(0.485, 0.456, 0.406),
                    std=(0.229, 0.224, 0.225), roi_capacity: int | None = None,
                    out: torch.Tensor | None = None):
    """Fan-out on the device: for every kept detection scale_boxes -> extract_crop ->
    resize -> ImageNet normalise -> CHW (reference inference.py:188-193 loop body).
    Returns ``(crops [roi_capacity,3,S,S], roi_offsets [N+1] int32, rois [roi_capacity,8] int32)``;
    ``roi_offsets[N]`` is the number of ROIs produced."""
    frames = _require_cuda(frames, torch.uint8, "frames")
    dets = _require_cuda(dets, torch.float32, "dets")
    count = _require_cuda(count, torch.int32, "count")
    n, h, w, _ = frames.shape
    max_det = dets.shape[1]
    cap = int(roi_capacity) if roi_capacity is not None else n * max_det
    dev = frames.device
    if out is None:
        out = torch.empty((cap, 3, out_size, out_size), dtype=torch.float32, device=dev)
    offsets = torch.empty((n + 1,), dtype=torch.int32, device=dev)
    rois = torch.empty((max(cap, 1), 8), dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        check(lib().ia_crop_resize_normalize(
            frames.data_ptr(), n, h, w, dets.data_ptr(), count.data_ptr(), max_det,
            float(np.float32(geom.scale)), geom.pad_w, geom.pad_h, int(out_size),
            F3(*[float(np.float32(v)) for v in mean]), F3(*[float(np.float32(v)) for v in std]),
            out.data_ptr(), cap, offsets.data_ptr(), rois.data_ptr(), _stream_ptr(dev)),
            "ia_crop_resize_normalize")
    return out, offsets, rois


def resize_normalize_rois_cuda(frames: torch.Tensor, rois: torch.Tensor, out_size: int = 224,
                               mean=(0.485, 0.456, 0.406), std=(0.229, 0.224, 0.225)) -> torch.Tensor:
    """Resize + normalise + CHW on explicit rectangles ``rois`` int32 [R,8]
    ({image, x1, y1, x2, y2, empty, 0, 0}); MobileNetPreprocessor.preprocess
    (reference mobilenet_preprocess.py:114-160) when a ROI covers a whole image."""
    frames = _require_cuda(frames, torch.uint8, "frames")
    rois = _require_cuda(rois, torch.int32, "rois")
    n, h, w, _ = frames.shape
    r = rois.shape[0]
    out = torch.empty((r, 3, out_size, out_size), dtype=torch.float32, device=frames.device)
    with torch.cuda.device(frames.device):
        check(lib().ia_resize_normalize_rois(
            frames.data_ptr(), n, h, w, rois.data_ptr(), r, int(out_size),
            F3(*[float(np.float32(v)) for v in mean]), F3(*[float(np.float32(v)) for v in std]),
            out.data_ptr(), _stream_ptr(frames.device)), "ia_resize_normalize_rois")
    return out


def imagenet_normalize_cuda(image: torch.Tensor, mean=(0.485, 0.456, 0.406),
                            std=(0.229, 0.224, 0.225)) -> torch.Tensor:
    """imagenet_normalize (reference transforms.py:238-272) on a uint8 [..., 3] CUDA tensor."""
    image = _require_cuda(image, torch.uint8, "image")
    out = torch.empty(image.shape, dtype=torch.float32, device=image.device)
    with torch.cuda.device(image.device):
        check(lib().ia_imagenet_normalize_u8(
            image.data_ptr(), image.numel() // 3, F3(*[float(np.float32(v)) for v in mean]),
            F3(*[float(np.float32(v)) for v in std]), out.data_ptr(), _stream_ptr(image.device)),
            "ia_imagenet_normalize_u8")
    return out


# ------------------------------------------------------------------- ragged batches (N2)
def _frame_descs(frames: list[torch.Tensor]):
    """ctypes array of ia_frame_desc for a list of uint8 [H,W,3] CUDA tensors."""
    if not frames:
        raise ValueError("empty frame list")
    keep = [_require_cuda(f, torch.uint8, "frame") for f in frames]
    for f in keep:
        if f.ndim != 3 or f.shape[2] != 3:
            raise ValueError(f"Expected [H, W, 3] frames, got {tuple(f.shape)}")
    arr = (_native.FrameDesc * len(keep))()
    for i, f in enumerate(keep):
        arr[i].data_dev, arr[i].h, arr[i].w = f.data_ptr(), f.shape[0], f.shape[1]
    return arr, keep


def preprocess_ragged_cuda(frames: list[torch.Tensor], target: int = 640, divisor: float = 255.0) -> torch.Tensor:
    """YOLOPreprocessor.preprocess (reference yolo_preprocess.py:128-174) for frames of
    different sizes: list of uint8 [H_i,W_i,3] CUDA tensors -> float32 [N,3,target,target]."""
    descs, keep = _frame_descs(frames)
    dev = keep[0].device
    out = torch.empty((len(keep), 3, target, target), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        check(lib().ia_yolo_preprocess_ragged(ctypes.addressof(descs), len(keep), int(target),
                                              float(np.float32(divisor)), out.data_ptr(), _stream_ptr(dev)),
              "ia_yolo_preprocess_ragged")
    return out


def crop_ragged_cuda(frames: list[torch.Tensor], dets: torch.Tensor, count: torch.Tensor, target: int = 640,
                     out_size: int = 224, mean=(0.485, 0.456, 0.406), std=(0.229, 0.224, 0.225),
                     roi_capacity: int | None = None):
    """The fan-out (scale_boxes -> extract_crop -> resize -> normalise -> CHW) over frames of
    different sizes; every image's letterbox geometry is derived on the device.
    Returns ``(crops, roi_offsets, rois)`` like ``crop_batch_cuda``."""
    descs, keep = _frame_descs(frames)
    dev = keep[0].device
    dets = _require_cuda(dets, torch.float32, "dets")
    count = _require_cuda(count, torch.int32, "count")
    n, max_det = len(keep), dets.shape[1]
    cap = int(roi_capacity) if roi_capacity is not None else n * max_det
    # the descriptor table lives on the device for this entry point (ROI -> image -> frame)
    table = torch.frombuffer(bytearray(bytes(descs)), dtype=torch.uint8).to(dev)
    out = torch.empty((cap, 3, out_size, out_size), dtype=torch.float32, device=dev)
    offsets = torch.empty((n + 1,), dtype=torch.int32, device=dev)
    rois = torch.empty((max(cap, 1), 8), dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        check(lib().ia_crop_resize_normalize_ragged(
            table.data_ptr(), n, max(f.shape[1] for f in keep), dets.data_ptr(), count.data_ptr(), max_det,
            int(target), int(out_size), F3(*[float(np.float32(v)) for v in mean]),
            F3(*[float(np.float32(v)) for v in std]), out.data_ptr(), cap, offsets.data_ptr(), rois.data_ptr(),
            _stream_ptr(dev)), "ia_crop_resize_normalize_ragged")
    return out, offsets, rois


# --------------------------------------------------------------- float32-input branch (N2)
def float_resize_flavour() -> str:
    """Which cv2 build the float32-crop resize reproduces bit for bit: ``"ipp"`` (default — the
    opencv-python wheels the reference pins route float resizes through Intel IPP) or ``"opencv"``
    (OpenCV's own generic path: cv2 built without IPP or with ``cv2.ipp.setUseIPP(False)``).
    Chosen with the environment variable ``IA_F32_RESIZE``."""
    import os
    flavour = os.environ.get("IA_F32_RESIZE", "ipp").lower()
    if flavour not in ("ipp", "opencv"):
        raise ValueError(f"IA_F32_RESIZE must be 'ipp' or 'opencv', got {flavour!r}")
    return flavour


def resize_normalize_f32_cuda(crop: torch.Tensor, out_size: int = 224, mean=(0.485, 0.456, 0.406),
                              std=(0.229, 0.224, 0.225), flavour: str | None = None) -> torch.Tensor:
    """MobileNetPreprocessor.preprocess on a float32 [H,W,3] CUDA crop (reference
    mobilenet_preprocess.py:114-160, float branch) -> float32 [1,3,S,S].  ``flavour``: see
    :func:`float_resize_flavour`."""
    flavour = flavour or float_resize_flavour()
    code = _native.IA_F32_RESIZE_IPP if flavour == "ipp" else _native.IA_F32_RESIZE_OPENCV
    crop = _require_cuda(crop, torch.float32, "crop")
    h, w, _ = crop.shape
    dev = crop.device
    out = torch.empty((1, 3, out_size, out_size), dtype=torch.float32, device=dev)
    need = lib().ia_f32_scratch_bytes(int(out_size))
    scratch = torch.empty(need, dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        check(lib().ia_resize_normalize_f32_as(crop.data_ptr(), h, w, int(out_size), code,
                                               F3(*[float(np.float32(v)) for v in mean]),
                                               F3(*[float(np.float32(v)) for v in std]), out.data_ptr(),
                                               scratch.data_ptr(), need, _stream_ptr(dev)),
              "ia_resize_normalize_f32_as")
    return out


def imagenet_normalize_f32_cuda(image: torch.Tensor, mean=(0.485, 0.456, 0.406),
                                std=(0.229, 0.224, 0.225)) -> torch.Tensor:
    """imagenet_normalize on float32 [..., 3] (reference transforms.py:264-270): divided by 255
    only when the image's maximum exceeds 1."""
    image = _require_cuda(image, torch.float32, "image")
    out = torch.empty_like(image)
    scratch = torch.empty(16, dtype=torch.uint8, device=image.device)
    with torch.cuda.device(image.device):
        check(lib().ia_imagenet_normalize_f32(image.data_ptr(), image.numel() // 3,
                                              F3(*[float(np.float32(v)) for v in mean]),
                                              F3(*[float(np.float32(v)) for v in std]), out.data_ptr(),
                                              scratch.data_ptr(), 16, _stream_ptr(image.device)),
              "ia_imagenet_normalize_f32")
    return out


# ------------------------------------------------------------------ classifier logits (N4)
def classify_topk_cuda(logits: torch.Tensor, k: int = 5, softmax: bool = True,
                       low_index_first: bool = False) -> tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """float32 [R, C] CUDA logits -> (ids int32 [R,k], values float32 [R,k], status int32 [R]).

    ``softmax=True``: the probabilities and order of ClassificationInference.classify (reference
    architectures/microservices/classification/app/inference.py:112-118).  ``softmax=False, k=1,
    low_index_first=True``: np.argmax and the raw value (monolithic/app/inference.py:201-202).
    ``status`` bit 0: the row holds a NaN."""
    logits = _require_cuda(logits, torch.float32, "logits")
    if logits.ndim != 2:
        raise ValueError(f"logits must be [rows, classes], got {tuple(logits.shape)}")
    rows, classes = logits.shape
    dev = logits.device
    ids = torch.empty((rows, k), dtype=torch.int32, device=dev)
    values = torch.empty((rows, k), dtype=torch.float32, device=dev)
    status = torch.empty((rows,), dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        check(lib().ia_classify_topk(logits.data_ptr(), rows, classes, int(k), int(bool(softmax)),
                                     int(bool(low_index_first)), ids.data_ptr(), values.data_ptr(),
                                     status.data_ptr(), _stream_ptr(dev)), "ia_classify_topk")
    return ids, values, status
