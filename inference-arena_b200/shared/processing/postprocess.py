"""YOLO head post-processing — same interface as the reference's
``architectures/*/app/postprocess.py`` (parse_yolo_output :12-73, apply_nms :76-160), plus
the v5 ``(N, 25200, 85)`` decode of ``src/shared/data/curator.py:289-307`` behind the same
``parse_yolo_output`` name, dispatched on the head's shape the way ``curator.py:251-261``
does.  Decode, filter, NMS and row assembly run on the GPU (K2 + K3).
"""

from __future__ import annotations

import numpy as np
import torch

from shared.processing import _hostio, device as _dev

_FIRST_MAX_DET = 1024


def _run(head: torch.Tensor, layout: str, conf: float, iou: float):
    anchors = head.shape[1] if layout == "v5" else head.shape[2]
    max_det = min(_FIRST_MAX_DET, anchors)
    while True:
        dets, keep, count, status = _dev.decode_nms_cuda(head, conf, iou, layout=layout, max_det=max_det)
        if int(status.max().item()) == 0 or max_det >= anchors:
            return dets, keep, count
        max_det = anchors  # rare: more detections than rows; redo with room for every anchor


def parse_yolo_output(raw_output: np.ndarray, confidence_threshold: float,
                      iou_threshold: float) -> np.ndarray:
    """``[1, 84, A]`` (v8 layout) or ``[1, A, 85]`` (v5 layout) head -> float32 ``[K, 6]`` rows
    ``[x1, y1, x2, y2, confidence, class_id]``; ``zeros((0, 6))`` when nothing is kept."""
    if raw_output.ndim != 3:
        raise ValueError(f"Expected a 3D head tensor, got {raw_output.ndim}D")
    layout = "v5" if raw_output.shape[2] == 85 and raw_output.shape[1] != 84 else "v8"
    head = _hostio.to_device(np.asarray(raw_output[:1], dtype=np.float32))
    dets, _, count = _run(head, layout, confidence_threshold, iou_threshold)
    k = int(count[0].item())
    if k == 0:
        return np.zeros((0, 6), dtype=np.float32)
    return _hostio.to_host(dets[0, :k])


def apply_nms(boxes: np.ndarray, scores: np.ndarray, class_ids: np.ndarray, conf_threshold: float,
              iou_threshold: float) -> list[int]:
    """Class-aware NMS on ``[N,4]`` cx,cy,w,h boxes; returns the indices to keep
    (class-ascending, score-descending).

    The reference groups with ``np.unique(class_ids)`` (postprocess.py:116), so any integer ids
    work there; the device key holds a class in 10 bits.  The ids are therefore renumbered
    densely in ascending order first (order-preserving, so the output order is unchanged)."""
    m = int(len(scores))
    if m == 0:
        return []
    _, dense = np.unique(np.asarray(class_ids), return_inverse=True)
    dense = dense.reshape(-1)
    if int(dense.max()) >= _dev.NMS_MAX_CLASSES:
        raise ValueError(f"apply_nms: more than {_dev.NMS_MAX_CLASSES} distinct class ids")
    keep, count, _ = _dev.nms_cuda(_hostio.to_device(np.asarray(boxes, dtype=np.float32)),
                                   _hostio.to_device(np.asarray(scores, dtype=np.float32)),
                                   _hostio.to_device(dense.astype(np.int32)),
                                   conf_threshold, iou_threshold, max_det=m)
    k = int(count[0].item())
    return [int(v) for v in _hostio.to_host(keep[:k])]
