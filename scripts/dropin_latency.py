"""Ad-hoc: per-call latency of the numpy drop-in API (host arrays in/out) on one request."""
import sys, time
from pathlib import Path
R = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(R)); sys.path.insert(0, str(R / "inference-arena_b200"))
import numpy as np
from workloads import synth
from shared.processing import YOLOPreprocessor, MobileNetPreprocessor
from shared.processing.mobilenet_preprocess import extract_crop
from shared.processing.postprocess import parse_yolo_output
def t(fn, n=50):
    for _ in range(5): fn()
    t0 = time.perf_counter()
    for _ in range(n): r = fn()
    return (time.perf_counter() - t0) / n * 1e3, r
for (h, w) in [(480, 640), (1080, 1920)]:
    img = synth.frames(1, h, w, seed=0)[0]; head = synth.head_v5(1, seed=1, objects=4)
    yp, mp = YOLOPreprocessor(), MobileNetPreprocessor()
    a, res = t(lambda: yp(img)); b, dets = t(lambda: parse_yolo_output(head, 0.25, 0.45))
    orig = res.scale_boxes_to_original(dets)
    c, _ = t(lambda: res.scale_boxes_to_original(dets))
    d, crops = t(lambda: [extract_crop(img, x) for x in orig]); e, _ = t(lambda: [mp(cr) for cr in crops])
    print(f"{h}x{w}: YOLOPreprocessor {a:.2f} ms, parse_yolo_output {b:.2f} ms, scale_boxes {c:.2f} ms, "
          f"{len(orig)} x extract_crop {d:.2f} ms, {len(orig)} x MobileNetPreprocessor {e:.2f} ms -> total {a+b+c+d+e:.2f} ms")
