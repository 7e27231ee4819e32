"""Seeded synthetic workloads for the hot path.

TEST INFRASTRUCTURE (see ``oracle/__init__.py``).  There is no network, dataset or
model in the loop: frames are uniform uint8 noise of COCO-like shape and heads are
synthetic ``(N, 25200, 85)`` tensors with a planted number of objects per image that
follows the reference's curated fan-out ({3: 25 %, 4: 50 %, 5: 25 %},
``data/thesis_test_set/manifest.json``).
"""

from __future__ import annotations

import numpy as np

ANCHORS = 25200
CLASSES = 80
FANOUT_VALUES = np.array([3, 4, 5])
FANOUT_PROBS = np.array([0.25, 0.5, 0.25])


def frames(n: int, h: int, w: int, seed: int) -> np.ndarray:
    """uint8[n,h,w,3] uniform noise (the reference's own fixtures use the same,
    tests/conftest.py:24-100)."""
    return np.random.default_rng(seed).integers(0, 256, (n, h, w, 3), dtype=np.uint8)


def head_v5(n: int, seed: int, anchors: int = ANCHORS, classes: int = CLASSES,
            objects=None, dup: int = 30, size: float = 640.0) -> np.ndarray:
    """float32[n,anchors,5+classes]: low-confidence background plus, per image, k planted
    objects each covered by ``dup`` jittered anchors of one class (~k*dup candidates at
    conf 0.25, ~k survivors of NMS at IoU 0.45)."""
    rng = np.random.default_rng(seed)
    head = np.empty((n, anchors, 5 + classes), dtype=np.float32)
    for i in range(n):
        img = head[i]
        img[:, 0:2] = rng.random((anchors, 2), dtype=np.float32) * np.float32(size)
        img[:, 2:4] = np.float32(8) + rng.random((anchors, 2), dtype=np.float32) * np.float32(192)
        img[:, 4] = rng.random(anchors, dtype=np.float32) * np.float32(0.05)
        img[:, 5:] = rng.random((anchors, classes), dtype=np.float32) * np.float32(0.3)
        k = int(rng.choice(FANOUT_VALUES, p=FANOUT_PROBS)) if objects is None else int(objects)
        if k * dup > anchors:
            raise ValueError("too many planted anchors")
        slots = rng.choice(anchors, size=k * dup, replace=False)
        for o in range(k):
            rows = slots[o * dup : (o + 1) * dup]
            centre = rng.uniform(80, size - 80, size=2)
            wh = rng.uniform(60, 220, size=2)
            cls = int(rng.integers(0, classes))
            img[rows, 0:2] = (centre + rng.normal(0, 3, size=(dup, 2))).astype(np.float32)
            img[rows, 2:4] = (wh + rng.normal(0, 4, size=(dup, 2))).astype(np.float32)
            img[rows, 4] = rng.uniform(0.5, 0.99, size=dup).astype(np.float32)
            img[rows, 5 + cls] = rng.uniform(0.6, 0.99, size=dup).astype(np.float32)
    return head


def head_v5_dense(n: int, seed: int, candidates: int, anchors: int = ANCHORS,
                  classes: int = CLASSES, used_classes: int = 80) -> np.ndarray:
    """Stress variant: ``candidates`` random anchors per image pass the threshold with
    random boxes/classes (many survive NMS); ``candidates == anchors`` is the all-pass
    worst case."""
    rng = np.random.default_rng(seed)
    head = np.empty((n, anchors, 5 + classes), dtype=np.float32)
    head[..., 0:2] = rng.random((n, anchors, 2), dtype=np.float32) * np.float32(640)
    head[..., 2:4] = np.float32(8) + rng.random((n, anchors, 2), dtype=np.float32) * np.float32(120)
    head[..., 4] = rng.random((n, anchors), dtype=np.float32) * np.float32(0.05)
    head[..., 5:] = rng.random((n, anchors, classes), dtype=np.float32) * np.float32(0.3)
    for i in range(n):
        rows = rng.choice(anchors, size=candidates, replace=False)
        head[i, rows, 4] = rng.uniform(0.6, 0.99, size=candidates).astype(np.float32)
        cls = rng.integers(0, used_classes, size=candidates)
        head[i, rows, 5 + cls] = rng.uniform(0.6, 0.99, size=candidates).astype(np.float32)
        for _ in range(64):  # re-draw objectness of rows whose (class, conf) pair collides
            conf = head[i, rows, 4] * head[i, rows, 5 + cls]
            key = cls.astype(np.int64) << 32 | conf.view(np.uint32).astype(np.int64)
            _, first = np.unique(key, return_index=True)
            dup_rows = np.setdiff1d(np.arange(candidates), first)
            if dup_rows.size == 0:
                break
            head[i, rows[dup_rows], 4] = rng.uniform(0.6, 0.99, size=dup_rows.size).astype(np.float32)
    return head


def head_v8_from_v5(head5: np.ndarray) -> np.ndarray:
    """Channel-major ``(n, 4+C, A)`` head whose class scores are obj*cls of a v5 head, so
    the same objects come out (postprocess.py:12-41 layout)."""
    rows = np.concatenate([head5[..., :4], head5[..., 4:5] * head5[..., 5:]], axis=-1)
    return np.ascontiguousarray(rows.transpose(0, 2, 1))


def assert_tie_free(head: np.ndarray, conf_thr: float) -> None:
    """Post-condition of SURVEY §8d/C3: no two candidates of one class in an image share a
    float32 confidence (the reference's order would then be implementation-defined)."""
    for i in range(head.shape[0]):
        cls = head[i, :, 5:]
        conf = head[i, :, 4] * cls.max(axis=1)
        cid = cls.argmax(axis=1)
        m = conf >= conf_thr
        key = np.stack([cid[m].astype(np.int64), conf[m].view(np.uint32).astype(np.int64)], axis=1)
        if len(np.unique(key, axis=0)) != len(key):
            raise AssertionError(f"confidence tie within a class in image {i}")


def detections(n: int, seed: int, kmax: int = 8, size: float = 640.0):
    """C4 input: float32[n,kmax,6] letterbox-space rows [x1,y1,x2,y2,conf,cls] and
    int32[n] counts with the {3,4,5} fan-out; unused rows are zero."""
    rng = np.random.default_rng(seed)
    dets = np.zeros((n, kmax, 6), dtype=np.float32)
    count = rng.choice(FANOUT_VALUES, size=n, p=FANOUT_PROBS).astype(np.int32)
    for i in range(n):
        k = int(count[i])
        c = rng.uniform(0, size, size=(k, 2))
        wh = rng.uniform(40, 320, size=(k, 2))
        dets[i, :k, 0:2] = (c - wh / 2).astype(np.float32)
        dets[i, :k, 2:4] = (c + wh / 2).astype(np.float32)
        dets[i, :k, 4] = rng.uniform(0.25, 0.99, size=k).astype(np.float32)
        dets[i, :k, 5] = rng.integers(0, CLASSES, size=k).astype(np.float32)
    return dets, count


def logits(rows: int, seed: int, classes: int = 1000, spread: float = 4.0) -> np.ndarray:
    """MobileNetV2-like logits: N(0, spread) with, per row, the five largest values pushed at
    least 0.05 apart so the top-5 order survives a 1-ulp difference in exp()."""
    rng = np.random.default_rng(seed)
    x = rng.normal(0.0, spread, (rows, classes)).astype(np.float32)
    take = min(6, classes)
    top = np.argsort(x, axis=1)[:, ::-1][:, :take]
    for r in range(rows):
        base = x[r, top[r, take - 1]]
        for j in range(take - 1):
            x[r, top[r, j]] = np.float32(base + 0.05 * (take - 1 - j) + abs(x[r, top[r, j]] - base))
        order = np.argsort(x[r])[::-1][:take]
        assert np.all(np.diff(x[r, order]) <= -0.04), "logit generator produced a near tie"
    return x
