#!/usr/bin/env python
"""The reference's monolithic request flow (architectures/monolithic/app/inference.py:127-227)
written against this package's drop-in names — the only change a service needs is where
`shared.processing` and `postprocess` are imported from.  The two ONNX sessions are stand-ins
(no model files or onnxruntime here): a synthetic YOLOv5n head and random logits.

    python examples/monolithic_predict.py            # needs a CUDA device
"""

import sys
import time
from pathlib import Path

import numpy as np

REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO / "inference-arena_b200"))
sys.path.insert(0, str(REPO))

from shared.processing import MobileNetPreprocessor, YOLOPreprocessor  # noqa: E402
from shared.processing.mobilenet_preprocess import extract_crop  # noqa: E402
from shared.processing.postprocess import parse_yolo_output  # noqa: E402


class FakeYolo:
    """Stands in for ModelRegistry.get_session("yolov5n").run (src/shared/model/registry.py)."""

    def run(self, _names, feeds):
        from workloads import synth  # seeded synthetic head generator

        assert feeds["images"].shape == (1, 3, 640, 640) and feeds["images"].dtype == np.float32
        return [synth.head_v5(1, seed=1, objects=4)]


class FakeMobileNet:
    def run(self, _names, feeds):
        assert feeds["input"].shape == (1, 3, 224, 224) and feeds["input"].dtype == np.float32
        return [np.random.default_rng(0).normal(size=(1, 1000)).astype(np.float32)]


def predict(image: np.ndarray, conf: float = 0.5, iou: float = 0.45):
    yolo_pre, mob_pre, yolo, mobilenet = YOLOPreprocessor(), MobileNetPreprocessor(), FakeYolo(), FakeMobileNet()
    t0 = time.perf_counter()
    yolo_result = yolo_pre(image)                                              # inference.py:161
    head = yolo.run(None, {"images": yolo_result.tensor})[0]                   # :164
    dets = parse_yolo_output(head, conf, iou)                                  # :167
    dets_orig = yolo_result.scale_boxes_to_original(dets) if len(dets) else dets  # :174-177
    results = []
    for det in dets_orig:                                                      # :188
        crop = extract_crop(image, det)                                        # :190
        tensor = mob_pre(crop).tensor                                          # :193
        logits = mobilenet.run(None, {"input": tensor})[0]                     # :196-198
        results.append({"box": [float(v) for v in det[:4]], "confidence": float(det[4]),
                        "class_id": int(det[5]), "top1": int(np.argmax(logits[0]))})
    return results, (time.perf_counter() - t0) * 1e3


if __name__ == "__main__":
    frame = np.random.default_rng(0).integers(0, 256, (480, 640, 3), dtype=np.uint8)
    predict(frame)  # warm-up (library load, CUDA context)
    out, ms = predict(frame)
    print(f"{len(out)} detections in {ms:.2f} ms (pre/post-processing on the GPU through the drop-in API)")
    for r in out:
        print(r)
