"""numpy restatement of the reference's hot-path functions.

TEST INFRASTRUCTURE (see ``oracle/__init__.py``).  Every function cites the
reference lines it follows (paths relative to ``/root/reference``).  Arithmetic
types and operation order follow the reference exactly (float32 element maths,
float64 host scalars); only the code structure is ours.

``resize`` defaults to the dependency-free restatement in ``cv_resize``; pass
``resize=cv2_resize`` to call OpenCV the way the reference does (used for the
timed CPU arm, where OpenCV's own speed is what the reference would get).
"""

from __future__ import annotations

from dataclasses import dataclass

import numpy as np

from .cv_resize import resize_linear_f32, resize_linear_f32_ipp, resize_linear_u8

# experiment.yaml:232-245 (controlled_variables.preprocessing) and transforms.py:37
YOLO_SIZE = 640
YOLO_DIVISOR = 255.0
MOBILENET_SIZE = 224
MEAN = np.array([0.485, 0.456, 0.406], dtype=np.float32)
STD = np.array([0.229, 0.224, 0.225], dtype=np.float32)
PAD_VALUE = 114
NUM_CLASSES = 80


def cv2_resize(img: np.ndarray, dst_w: int, dst_h: int) -> np.ndarray:
    import cv2

    return cv2.resize(img, (dst_w, dst_h), interpolation=cv2.INTER_LINEAR)


# --------------------------------------------------------------------------- a1
@dataclass(frozen=True)
class LetterboxGeometry:
    src_h: int
    src_w: int
    target: int
    scale: float
    new_w: int
    new_h: int
    pad_w: int
    pad_h: int


def letterbox_geometry(src_h: int, src_w: int, target: int) -> LetterboxGeometry:
    """Host scalars of ``letterbox`` (transforms.py:150-157, 174-175): float64 scale,
    truncating ``int()``, floor-halved padding."""
    scale = min(target / src_h, target / src_w)
    new_w = int(src_w * scale)
    new_h = int(src_h * scale)
    return LetterboxGeometry(
        src_h, src_w, target, scale, new_w, new_h, (target - new_w) // 2, (target - new_h) // 2
    )


def letterbox(image: np.ndarray, target: int, color=(PAD_VALUE,) * 3, resize=resize_linear_u8):
    """transforms.py:118-180 -> (uint8[T,T,3], scale, (pad_w, pad_h))."""
    g = letterbox_geometry(image.shape[0], image.shape[1], target)
    canvas = np.empty((target, target, 3), dtype=np.uint8)
    canvas[...] = np.asarray(color, dtype=np.uint8)
    canvas[g.pad_h : g.pad_h + g.new_h, g.pad_w : g.pad_w + g.new_w] = resize(image, g.new_w, g.new_h)
    return canvas, g.scale, (g.pad_w, g.pad_h)


# --------------------------------------------------------------------------- a2
def _check_image(image, what: str, dtypes) -> None:
    """yolo_preprocess.py:176-195 / mobilenet_preprocess.py:186-210 (same messages)."""
    if not isinstance(image, np.ndarray):
        raise ValueError(f"Expected numpy array, got {type(image)}")
    if image.ndim != 3:
        raise ValueError(f"Expected 3D array [H, W, C], got {image.ndim}D")
    if image.shape[2] != 3:
        raise ValueError(f"Expected 3 channels, got {image.shape[2]}")
    if image.dtype not in dtypes:
        names = " or ".join(np.dtype(d).name for d in dtypes)
        raise ValueError(f"Expected {names} dtype, got {image.dtype}")
    if what == "crop" and (image.shape[0] < 1 or image.shape[1] < 1):
        raise ValueError(f"Invalid crop dimensions: {image.shape[:2]}")


def yolo_preprocess(image: np.ndarray, target: int = YOLO_SIZE, divisor: float = YOLO_DIVISOR,
                    resize=resize_linear_u8):
    """yolo_preprocess.py:128-174 -> (float32[1,3,T,T], scale, (pad_w,pad_h), (h,w)).

    float32 true division by ``divisor`` (:158), HWC->CHW (:161), batch dim (:164),
    C-contiguous (:167)."""
    _check_image(image, "image", (np.uint8,))
    boxed, scale, pad = letterbox(image, target, resize=resize)
    chw = (boxed.astype(np.float32) / np.float32(divisor)).transpose(2, 0, 1)
    return np.ascontiguousarray(chw[None]), scale, pad, (image.shape[0], image.shape[1])


# --------------------------------------------------------------------------- a3
def decode_v5(head_image: np.ndarray):
    """curator.py:295-304 for one image ``[A, 5+C]``: conf = obj * max(cls) in float32,
    class id = first arg-max.  Returns (boxes_cxcywh[A,4], conf[A], cid[A])."""
    cls = head_image[:, 5:]
    conf = head_image[:, 4] * cls.max(axis=1)
    return head_image[:, :4], conf, cls.argmax(axis=1)


def decode_v8(raw_image: np.ndarray):
    """postprocess.py:33-41 for one image ``[4+C, A]`` (channel-major, no objectness)."""
    rows = raw_image.T
    cls = rows[:, 4:]
    return rows[:, :4], cls.max(axis=1), cls.argmax(axis=1)


# --------------------------------------------------------------------------- a4
def nms(boxes: np.ndarray, scores: np.ndarray, class_ids: np.ndarray,
        conf_thr: float, iou_thr: float) -> list[int]:
    """postprocess.py:76-160 (== curator.py:323-400): confidence filter, cxcywh->xyxy,
    per-class greedy NMS; returns ORIGINAL indices, class-ascending then
    score-descending.

    Score ties: the reference's ``argsort()[::-1]`` (postprocess.py:129) leaves the
    order of equal scores to numpy's unstable sort.  The oracle fixes it as "larger
    original index first" (a stable argsort, reversed), which is what numpy does for
    short arrays; workloads used for parity are asserted tie-free."""
    scores = np.asarray(scores)
    cand = np.flatnonzero(scores >= conf_thr)  # :99, :107
    if cand.size == 0:
        return []
    b = boxes[cand]
    s = scores[cand]
    c = np.asarray(class_ids)[cand]
    half_w = b[:, 2] / 2
    half_h = b[:, 3] / 2
    x1, y1 = b[:, 0] - half_w, b[:, 1] - half_h  # :110-113
    x2, y2 = b[:, 0] + half_w, b[:, 1] + half_h
    area = (x2 - x1) * (y2 - y1)  # :148-151
    eps = 1e-6

    kept: list[int] = []
    for cls in np.unique(c):  # :116-119, ascending
        members = np.flatnonzero(c == cls)
        order = members[np.argsort(s[members], kind="stable")[::-1]]  # :129
        while order.size:
            head, rest = order[0], order[1:]
            kept.append(int(cand[head]))  # :133
            if rest.size == 0:
                break
            iw = np.maximum(0, np.minimum(x2[head], x2[rest]) - np.maximum(x1[head], x1[rest]))
            ih = np.maximum(0, np.minimum(y2[head], y2[rest]) - np.maximum(y1[head], y1[rest]))
            inter = iw * ih  # :144-146
            union = area[head] + area[rest] - inter  # :152
            iou = inter / (union + eps)  # :154
            order = rest[iou <= iou_thr]  # :157-158
    return kept


# --------------------------------------------------------------------------- a5
def assemble_detections(boxes: np.ndarray, conf: np.ndarray, cid: np.ndarray, keep) -> np.ndarray:
    """postprocess.py:52-73 -> float32[K,6] rows [x1,y1,x2,y2,conf,class]."""
    if len(keep) == 0:
        return np.zeros((0, 6), dtype=np.float32)
    k = np.asarray(keep, dtype=np.int64)
    b = boxes[k]
    half_w, half_h = b[:, 2] / 2, b[:, 3] / 2
    cols = [b[:, 0] - half_w, b[:, 1] - half_h, b[:, 0] + half_w, b[:, 1] + half_h, conf[k], cid[k]]
    return np.column_stack(cols).astype(np.float32)


def detect_v5(head_image: np.ndarray, conf_thr: float, iou_thr: float):
    """a3+a4+a5 for one image of a ``(N, A, 85)`` head (SURVEY §8c composition).
    Returns (dets float32[K,6], keep list[int])."""
    boxes, conf, cid = decode_v5(head_image)
    keep = nms(boxes, conf, cid, conf_thr, iou_thr)
    return assemble_detections(boxes, conf, cid, keep), keep


def detect_v8(raw_image: np.ndarray, conf_thr: float, iou_thr: float):
    """postprocess.py:12-73 for one image of a ``(N, 84, A)`` head."""
    boxes, conf, cid = decode_v8(raw_image)
    keep = nms(boxes, conf, cid, conf_thr, iou_thr)
    return assemble_detections(boxes, conf, cid, keep), keep


# --------------------------------------------------------------------------- a6
def scale_boxes(boxes: np.ndarray, scale: float, padding, original_shape) -> np.ndarray:
    """transforms.py:183-230: un-pad, divide by scale (float32 array / python float ->
    float32 division by float32(scale)), clip to the original frame."""
    out = np.array(boxes, copy=True)
    pad_w, pad_h = padding
    h, w = original_shape
    out[:, 0] -= pad_w
    out[:, 1] -= pad_h
    out[:, 2] -= pad_w
    out[:, 3] -= pad_h
    out[:, :4] /= scale
    out[:, 0] = np.clip(out[:, 0], 0, w)
    out[:, 1] = np.clip(out[:, 1], 0, h)
    out[:, 2] = np.clip(out[:, 2], 0, w)
    out[:, 3] = np.clip(out[:, 3], 0, h)
    return out


# --------------------------------------------------------------------------- a7
def crop_rect(box, height: int, width: int):
    """mobilenet_preprocess.py:256-267: truncating int(), clamp; None when empty."""
    x1, y1, x2, y2 = (int(v) for v in box[:4])
    x1, y1 = max(0, x1), max(0, y1)
    x2, y2 = min(width, x2), min(height, y2)
    if x2 <= x1 or y2 <= y1:
        return None
    return x1, y1, x2, y2


def extract_crop(image: np.ndarray, box) -> np.ndarray:
    """mobilenet_preprocess.py:236-269 (empty rectangle -> zeros((1,1,3)))."""
    r = crop_rect(box, image.shape[0], image.shape[1])
    if r is None:
        return np.zeros((1, 1, 3), dtype=np.uint8)
    x1, y1, x2, y2 = r
    return image[y1:y2, x1:x2].copy()


# --------------------------------------------------------------------------- a9
def imagenet_normalize(image: np.ndarray, mean=MEAN, std=STD) -> np.ndarray:
    """transforms.py:238-272: /255 (uint8 always; float input only when max > 1),
    then (x - mean) / std, all float32."""
    if image.dtype == np.uint8:
        x = image.astype(np.float32) / 255.0
    else:
        x = image.astype(np.float32)
        if x.max() > 1.0:
            x /= 255.0
    return (x - mean) / std


# --------------------------------------------------------------------------- a8
def mobilenet_preprocess(crop: np.ndarray, size: int = MOBILENET_SIZE, resize=resize_linear_u8):
    """mobilenet_preprocess.py:114-160 -> float32[1,3,S,S] (uint8 or float32 crops)."""
    _check_image(crop, "crop", (np.uint8, np.float32))
    if crop.dtype == np.float32 and resize is resize_linear_u8:
        # :204-206 float crops go through cv2's float path: Intel IPP in the opencv-python wheels the
        # reference pins (IA_F32_RESIZE=opencv selects OpenCV's own generic path instead)
        import os
        resize = resize_linear_f32 if os.environ.get("IA_F32_RESIZE", "ipp").lower() == "opencv" else resize_linear_f32_ipp
    chw = imagenet_normalize(resize(crop, size, size)).transpose(2, 0, 1)
    return np.ascontiguousarray(chw[None])


# ------------------------------------------------------- classifier logits
def softmax_topk(scores: np.ndarray, top_k: int = 5):
    """``ClassificationInference.classify``'s arithmetic,
    architectures/microservices/classification/app/inference.py:112-118, on one score vector:
    returns (indices, probabilities of those indices).  The reference's ``argsort`` is an
    unstable sort, so equal probabilities have no defined order there; a stable sort is used
    here (reversed: the larger index first) and the synthetic logits are tie-free at the top."""
    exp_scores = np.exp(scores - np.max(scores))
    probs = exp_scores / exp_scores.sum()
    idx = np.argsort(probs, kind="stable")[::-1][:top_k]
    return idx.astype(np.int64), probs[idx]


def top1(cls_output: np.ndarray):
    """``architectures/monolithic/app/inference.py:201-202``: arg-max and the raw value."""
    class_id = int(np.argmax(cls_output[0]))
    return class_id, float(cls_output[0, class_id])


# --------------------------------------------------------------------- the chain
def run_image(frame: np.ndarray, head_image: np.ndarray, conf_thr: float, iou_thr: float,
              resize=resize_linear_u8):
    """One request through the path the way ``monolithic/app/inference.py:161-193``
    drives it: a2 -> a3/a4/a5 -> a6 -> per detection (a7 -> a8)."""
    tensor, scale, pad, shape = yolo_preprocess(frame, resize=resize)
    dets, keep = detect_v5(head_image, conf_thr, iou_thr)
    dets_orig = scale_boxes(dets, scale, pad, shape) if len(dets) else dets
    crops = [mobilenet_preprocess(extract_crop(frame, d), resize=resize) for d in dets_orig]
    crops = (np.concatenate(crops, axis=0) if crops
             else np.zeros((0, 3, MOBILENET_SIZE, MOBILENET_SIZE), np.float32))
    return tensor, dets, keep, dets_orig, crops
