"""TEST / MEASUREMENT INFRASTRUCTURE — never imported by the product path.

Runs the UNMODIFIED reference modules (matthewhoung/inference-arena) for one request of the hot
path, the way ``architectures/monolithic/app/inference.py:161-193`` drives them.  The files are
looked for under ``$IA_REFERENCE``, ``baseline/_ref`` (staged byte for byte by
``scripts/stage_reference.py``; it travels to the GPU box) and ``/root/reference``.  Used by
``bench.py``'s CPU arm (``kind: "reference"``) and by the tests that pin ``oracle/pipeline.py``
to it.  The one step the reference has no index-returning function for is the v5 head: its
``curator.py:295-304`` decode feeds a count-only NMS, so — as SURVEY §8c prescribes — that
decode is restated here (three numpy lines) in front of the reference's own ``apply_nms``
(``postprocess.py:76-160``) and its row assembly (``postprocess.py:56-73``, restated).
"""

from __future__ import annotations

import importlib.util
import os
import sys
from pathlib import Path

import numpy as np

REPO = Path(__file__).resolve().parent.parent


def find_reference() -> Path | None:
    cands = [os.environ.get("IA_REFERENCE"), REPO / "baseline" / "_ref", "/root/reference"]
    for c in cands:
        if c and (Path(c) / "src" / "shared" / "processing" / "transforms.py").exists():
            return Path(c)
    return None


class Reference:
    """The reference's functions, imported from its own files (module names are kept private
    to this object so that the drop-in ``shared`` package of this repository is not shadowed)."""

    def __init__(self, root: Path | None = None) -> None:
        root = root or find_reference()
        if root is None:
            raise FileNotFoundError("reference files not found (scripts/stage_reference.py stages them)")
        self.root = root
        saved_path, saved_mods = list(sys.path), {k: v for k, v in sys.modules.items()
                                                  if k == "shared" or k.startswith("shared.")}
        for k in saved_mods:
            del sys.modules[k]
        sys.path.insert(0, str(root / "src"))
        cwd = os.getcwd()
        try:
            os.chdir(root)  # config.py also looks for experiment.yaml in the working directory
            from shared.processing.mobilenet_preprocess import MobileNetPreprocessor, extract_crop
            from shared.processing.yolo_preprocess import YOLOPreprocessor
            self.yolo = YOLOPreprocessor()
            self.mobilenet = MobileNetPreprocessor()
            self.extract_crop = extract_crop
            spec = importlib.util.spec_from_file_location(
                "ia_ref_postprocess", root / "architectures" / "monolithic" / "app" / "postprocess.py")
            post = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(post)
            self.apply_nms = post.apply_nms
            self.shared_file = sys.modules["shared"].__file__
        finally:
            os.chdir(cwd)
            sys.path[:] = saved_path
            for k in [k for k in sys.modules if k == "shared" or k.startswith("shared.")]:
                del sys.modules[k]
            sys.modules.update(saved_mods)

    def run_image(self, frame: np.ndarray, head_image: np.ndarray, conf_thr: float, iou_thr: float):
        """(tensor, dets, keep, dets_orig, crops) — same tuple as oracle.pipeline.run_image."""
        res = self.yolo(frame)
        boxes, obj, cls = head_image[:, :4], head_image[:, 4], head_image[:, 5:]
        conf = obj * cls.max(axis=1)  # curator.py:302-303
        cid = cls.argmax(axis=1)      # curator.py:304
        keep = self.apply_nms(boxes, conf, cid, conf_thr, iou_thr)
        if len(keep) == 0:
            dets = np.zeros((0, 6), dtype=np.float32)
        else:
            b = boxes[keep]
            dets = np.column_stack([b[:, 0] - b[:, 2] / 2, b[:, 1] - b[:, 3] / 2, b[:, 0] + b[:, 2] / 2,
                                    b[:, 1] + b[:, 3] / 2, conf[keep], cid[keep]]).astype(np.float32)
        dets_orig = res.scale_boxes_to_original(dets) if len(dets) else dets
        crops = [self.mobilenet(self.extract_crop(frame, d)).tensor for d in dets_orig]
        crops = np.concatenate(crops, axis=0) if crops else np.zeros((0, 3, 224, 224), np.float32)
        return res.tensor, dets, list(keep), dets_orig, crops
