#!/usr/bin/env python
"""Generate the golden vectors in this directory FROM THE UNMODIFIED REFERENCE.

Run in the build container (the only place ``/root/reference`` exists):

    python tests/golden/make_golden.py

It imports the reference's own modules (``src/shared/processing/*``,
``src/shared/data/curator.py``, ``architectures/monolithic/app/postprocess.py``,
``architectures/microservices/classification/app/inference.py``; for JPEG decode
``load_image_from_bytes`` on files encoded by ``workloads/jpeg_cases.py``),
feeds them the seeded inputs of ``workloads/synth.py`` and stores what they return in
``golden.npz`` (small arrays verbatim; large tensors as SHA-256 of their bytes plus a
strided sample).  Nothing at test/bench time reads ``/root/reference``; the tests read
only the committed ``golden.npz``.
"""

from __future__ import annotations

import hashlib
import importlib.util
import json
import sys
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
REPO = HERE.parent.parent
REF = Path("/root/reference")
sys.path.insert(0, str(REPO))
sys.path.insert(0, str(REF / "src"))

from workloads import synth  # noqa: E402

LETTERBOX_SHAPES = [(480, 640), (1080, 1920), (1920, 1080), (500, 375), (333, 500), (427, 640),
                    (640, 640), (100, 50), (1281, 1279), (720, 1280), (1280, 1280), (37, 640)]
CROP_SHAPES = [(100, 150), (10, 10), (500, 800), (224, 224), (448, 448), (1, 1), (300, 17),
               (223, 225), (1, 300), (449, 447), (112, 112), (64, 1000)]
BOX_CASES = [  # (frame h, w) and boxes exercising clip / truncation / empty / negative
    ((480, 640), [[100.7, 100.2, 300.9, 400.5, 0.9, 0], [-20.5, -3.2, 50.5, 60.9, 0.8, 1],
                  [600.0, 400.0, 900.0, 700.0, 0.7, 2], [50.0, 50.0, 50.9, 300.0, 0.6, 3],
                  [10.0, 20.0, 5.0, 10.0, 0.5, 4], [0.0, 0.0, 640.0, 480.0, 0.4, 5],
                  [639.2, 479.1, 640.0, 480.0, 0.3, 6], [-0.9, -0.9, 0.9, 0.9, 0.3, 7]]),
]


def sha(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def sample(a: np.ndarray, n: int = 257) -> np.ndarray:
    flat = np.ascontiguousarray(a).reshape(-1)
    idx = np.linspace(0, flat.size - 1, num=min(n, flat.size)).astype(np.int64)
    return flat[idx].copy()


def load_postprocess():
    spec = importlib.util.spec_from_file_location(
        "ref_postprocess", REF / "architectures/monolithic/app/postprocess.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def main() -> None:
    from shared.data.curator import DetectionCounter
    from shared.processing.mobilenet_preprocess import MobileNetPreprocessor, extract_crop
    from shared.processing.transforms import imagenet_normalize, letterbox, scale_boxes
    from shared.processing.yolo_preprocess import YOLOPreprocessor

    post = load_postprocess()
    out: dict[str, np.ndarray] = {}
    meta: dict = {"letterbox": [], "crops": [], "heads": [], "versions": {}}
    import cv2
    meta["versions"] = {"numpy": np.__version__, "cv2": cv2.__version__}

    # ---- a1/a2: letterbox + YOLOPreprocessor --------------------------------
    yolo = YOLOPreprocessor()
    for i, (h, w) in enumerate(LETTERBOX_SHAPES):
        img = synth.frames(1, h, w, seed=100 + i)[0]
        boxed, scale, pad = letterbox(img, 640)
        res = yolo(img)
        meta["letterbox"].append({"h": h, "w": w, "seed": 100 + i, "scale": scale,
                                  "pad": list(pad), "u8_sha": sha(boxed),
                                  "tensor_sha": sha(res.tensor)})
        out[f"lb{i}_u8_sample"] = sample(boxed)
        out[f"lb{i}_tensor_sample"] = sample(res.tensor)
    # other target sizes (tests/shared/test_processing.py:115-123 uses 320..1280)
    for T in (320, 416, 1280):
        img = synth.frames(1, 480, 640, seed=150)[0]
        boxed, scale, pad = letterbox(img, T)
        meta["letterbox"].append({"h": 480, "w": 640, "seed": 150, "target": T, "scale": scale,
                                  "pad": list(pad), "u8_sha": sha(boxed)})

    # ---- a8/a9: MobileNetPreprocessor ---------------------------------------
    mob = MobileNetPreprocessor()
    for i, (h, w) in enumerate(CROP_SHAPES):
        crop = synth.frames(1, h, w, seed=200 + i)[0]
        t = mob(crop).tensor
        meta["crops"].append({"h": h, "w": w, "seed": 200 + i, "tensor_sha": sha(t)})
        out[f"crop{i}_tensor_sample"] = sample(t)
    norm_in = synth.frames(1, 8, 8, seed=250)[0]
    out["normalize_in"] = norm_in
    out["normalize_out"] = imagenet_normalize(norm_in)

    # ---- a6/a7: scale_boxes + extract_crop ----------------------------------
    for i, ((h, w), boxes) in enumerate(BOX_CASES):
        frame = synth.frames(1, h, w, seed=300 + i)[0]
        b = np.array(boxes, dtype=np.float32)
        out[f"box{i}_in"] = b
        for j, (sc, pad, shape) in enumerate([(1.0, (0, 80), (480, 640)),
                                              (0.3333333333333333, (0, 140), (1080, 1920)),
                                              (0.5, (10, 0), (1280, 1240))]):
            out[f"box{i}_scaled{j}"] = scale_boxes(b, sc, pad, shape)
        shapes, hashes = [], []
        for row in b:
            c = extract_crop(frame, row)
            shapes.append(c.shape)
            hashes.append(sha(c))
        out[f"box{i}_crop_shapes"] = np.array(shapes, dtype=np.int32)
        meta[f"box{i}_crop_sha"] = hashes

    # ---- a3/a4/a5: v5 decode + apply_nms + assembly --------------------------
    counter = DetectionCounter(models_dir=Path("/nonexistent"), confidence_threshold=0.25,
                               iou_threshold=0.45)
    head_cases = {
        "std": synth.head_v5(4, seed=2),
        "small": synth.head_v5(3, seed=5, anchors=1000, dup=12),
        "dense": synth.head_v5_dense(1, seed=6, candidates=2000),
        "allpass": synth.head_v5_dense(1, seed=7, candidates=3000, anchors=3000),
        "oneclass": synth.head_v5_dense(1, seed=8, candidates=1500, anchors=4000, used_classes=1),
        "empty": synth.head_v5(2, seed=9, objects=0),
    }
    for name, head in head_cases.items():
        synth.assert_tie_free(head, 0.25)
        for n in range(head.shape[0]):
            d = head[n]
            conf = d[:, 4] * d[:, 5:].max(axis=1)  # curator.py:302-303
            cid = d[:, 5:].argmax(axis=1)  # curator.py:304
            keep = post.apply_nms(d[:, :4], conf, cid, 0.25, 0.45)
            count = counter._parse_yolov5_raw_output(head[n:n + 1])
            assert count == len(keep), (name, n, count, len(keep))
            if keep:
                b = d[:, :4][keep]
                rows = np.column_stack([b[:, 0] - b[:, 2] / 2, b[:, 1] - b[:, 3] / 2,
                                        b[:, 0] + b[:, 2] / 2, b[:, 1] + b[:, 3] / 2,
                                        conf[keep], cid[keep]]).astype(np.float32)  # postprocess.py:56-73
            else:
                rows = np.zeros((0, 6), dtype=np.float32)
            out[f"head_{name}{n}_keep"] = np.array(keep, dtype=np.int64)
            out[f"head_{name}{n}_dets"] = rows
        meta["heads"].append({"name": name, "n": int(head.shape[0]),
                              "candidates": int(((head[..., 4] * head[..., 5:].max(-1)) >= 0.25).sum())})

    # ---- N1: v8 layout through the reference's own parse_yolo_output ---------
    h8 = synth.head_v8_from_v5(synth.head_v5(2, seed=11, anchors=8400))
    for n in range(2):
        out[f"v8_{n}_dets"] = post.parse_yolo_output(h8[n:n + 1], 0.5, 0.45)

    # ---- C1: the whole chain on one 640x480 frame ----------------------------
    frame = synth.frames(1, 480, 640, seed=0)[0]
    head = synth.head_v5(1, seed=1, objects=4)
    res = yolo(frame)
    d = head[0]
    conf = d[:, 4] * d[:, 5:].max(axis=1)
    cid = d[:, 5:].argmax(axis=1)
    keep = post.apply_nms(d[:, :4], conf, cid, 0.25, 0.45)
    b = d[:, :4][keep]
    dets = np.column_stack([b[:, 0] - b[:, 2] / 2, b[:, 1] - b[:, 3] / 2, b[:, 0] + b[:, 2] / 2,
                            b[:, 1] + b[:, 3] / 2, conf[keep], cid[keep]]).astype(np.float32)
    dets_orig = res.scale_boxes_to_original(dets)
    crops = [mob(extract_crop(frame, r)).tensor for r in dets_orig]
    out["c1_keep"] = np.array(keep, dtype=np.int64)
    out["c1_dets"] = dets
    out["c1_dets_orig"] = dets_orig
    meta["c1"] = {"tensor_sha": sha(res.tensor), "crops_sha": [sha(c) for c in crops]}
    out["c1_crop_samples"] = np.stack([sample(c) for c in crops])

    # ---- N4: softmax/top-k and arg-max through the reference's own classify() ----------
    # ClassificationInference is built without __init__ (that needs onnxruntime and a model
    # file); the session is a stub that returns the seeded logits, so every line of
    # classify() after session.run — inference.py:112-134 — runs unmodified.
    sys.path.insert(0, str(REF / "architectures/microservices/classification"))
    from app.inference import ClassificationInference

    class _Session:
        def __init__(self, row):
            self.row = row

        def run(self, _names, _feed):
            return [self.row[None]]

    class _Pre:
        tensor = None

    logits = synth.logits(16, seed=400)
    small = synth.logits(4, seed=401, classes=10)
    for name, rows, k in (("cls", logits, 5), ("cls_small", small, 3), ("cls_k1", logits[:4], 1)):
        ids, probs = [], []
        for row in rows:
            eng = object.__new__(ClassificationInference)
            eng.preprocessor = lambda crop: _Pre
            eng.mobilenet_session = _Session(row)
            eng.labels = [str(i) for i in range(rows.shape[1])]
            top_id, _, top_conf, top = eng.classify(None, top_k=k)
            assert top[0][0] == top_id and top[0][2] == top_conf
            ids.append([t[0] for t in top])
            probs.append([t[2] for t in top])
        out[f"{name}_ids"] = np.array(ids, dtype=np.int64)
        out[f"{name}_probs"] = np.array(probs, dtype=np.float64)  # float(probs[idx]) of f32 values

    # ---- N3: baseline JPEGs through the reference's own load_image_from_bytes --------------
    from shared.processing.transforms import load_image_from_bytes
    from workloads import jpeg_cases

    meta["jpeg"] = []
    cases = [((40, 56), "420", 85, {}), ((33, 17), "444", 60, {"restart": 2}), ((24, 40), "422", 90, {"optimize": True}),
             ((17, 33), "440", 30, {}), ((20, 30), "411", 75, {}), ((19, 23), "gray", 80, {}),
             ((24, 40), "420", 85, {"orientation": 6}), ((96, 128), "420", 95, {})]
    for i, (shape, sampling, q, kw) in enumerate(cases):
        kw = dict(kw)
        orientation = kw.pop("orientation", None)
        img = jpeg_cases.photo(*shape, seed=700 + i)
        data = (jpeg_cases.encode(img[..., 0], q, **kw) if sampling == "gray"
                else jpeg_cases.encode(img, q, sampling, **kw))
        if orientation:
            data = jpeg_cases.with_exif_orientation(data, orientation)
        rgb = load_image_from_bytes(data)
        out[f"jpeg{i}_bytes"] = np.frombuffer(data, dtype=np.uint8)
        out[f"jpeg{i}_rgb"] = rgb
        meta["jpeg"].append({"shape": list(rgb.shape), "sampling": sampling, "quality": q, "sha": sha(rgb)})

    out["meta_json"] = np.frombuffer(json.dumps(meta).encode(), dtype=np.uint8)
    np.savez_compressed(HERE / "golden.npz", **out)
    print("wrote", HERE / "golden.npz", (HERE / "golden.npz").stat().st_size, "bytes;",
          {h["name"]: h["candidates"] for h in meta["heads"]})


if __name__ == "__main__":
    main()
