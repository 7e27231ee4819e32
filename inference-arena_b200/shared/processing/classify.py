"""What the reference's services do with MobileNetV2's logits, on the GPU.

``softmax_top_k`` is the arithmetic inside ``ClassificationInference.classify``
(architectures/microservices/classification/app/inference.py:112-132); ``top1`` is the
arg-max + raw value of the monolithic and Triton-gateway pipelines
(architectures/monolithic/app/inference.py:201-202, triton/gateway/app/pipeline.py:181-182).
The batched device call is ``device.classify_topk_cuda``; these wrap it for one numpy row.
"""

from __future__ import annotations

import numpy as np

from shared.processing import _hostio, device as _dev


def _row(scores: np.ndarray) -> np.ndarray:
    scores = np.asarray(scores)
    if scores.ndim != 1 or scores.size == 0:
        raise ValueError(f"Expected a non-empty 1D score vector, got shape {scores.shape}")
    return np.ascontiguousarray(scores, dtype=np.float32)[None]


def softmax_top_k(scores: np.ndarray, top_k: int = 5) -> list[tuple[int, float]]:
    """``[(class_id, probability), ...]`` for the ``top_k`` most probable classes, most
    probable first (``top_k`` larger than the class count is cut to it, like the slice)."""
    row = _row(scores)
    k = min(int(top_k), row.shape[1])
    if k <= 0:
        return []
    ids, probs, _ = _dev.classify_topk_cuda(_hostio.to_device(row), k, softmax=True)
    ids, probs = _hostio.to_host(ids)[0], _hostio.to_host(probs)[0]
    return [(int(i), float(p)) for i, p in zip(ids, probs)]


def top1(cls_output: np.ndarray) -> tuple[int, float]:
    """``(int(np.argmax(cls_output[0])), float(cls_output[0, class_id]))``."""
    cls_output = np.asarray(cls_output)
    if cls_output.ndim != 2 or cls_output.shape[1] == 0:
        raise ValueError(f"Expected a [1, classes] array, got shape {cls_output.shape}")
    ids, vals, _ = _dev.classify_topk_cuda(_hostio.to_device(_row(cls_output[0])), 1, softmax=False,
                                           low_index_first=True)
    return int(_hostio.to_host(ids)[0, 0]), float(_hostio.to_host(vals)[0, 0])
