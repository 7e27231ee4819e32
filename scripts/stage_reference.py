#!/usr/bin/env python
"""Stage the UNMODIFIED reference files of the hot path under ``baseline/_ref`` (git-ignored;
it travels to the GPU box with the working tree, where ``/root/reference`` does not exist).

    python scripts/stage_reference.py [--reference /root/reference]

Two users, both test / measurement infrastructure, never the product path:
  * ``bench.py --impl reference`` (and the ``cpu_baseline`` leg) times the reference's own
    functions from here (``kind: "reference"``);
  * ``tests/test_gpu_reference_suite.py`` runs the reference's own acceptance suite
    (``tests/shared/test_processing.py``, 71 tests) against the drop-in package.
Files are copied byte for byte; nothing here is committed.
"""
from __future__ import annotations

import argparse
import shutil
import sys
from pathlib import Path

REPO = Path(__file__).resolve().parent.parent
DEST = REPO / "baseline" / "_ref"
FILES = [
    "experiment.yaml",
    "src/shared/__init__.py",
    "src/shared/config.py",
    "src/shared/processing/__init__.py",
    "src/shared/processing/transforms.py",
    "src/shared/processing/yolo_preprocess.py",
    "src/shared/processing/mobilenet_preprocess.py",
    "src/shared/data/__init__.py",
    "src/shared/data/coco_dataset.py",
    "src/shared/data/curator.py",
    "architectures/monolithic/app/postprocess.py",
    "tests/__init__.py",
    "tests/conftest.py",
    "tests/shared/__init__.py",
    "tests/shared/test_processing.py",
]


def stage(reference: Path) -> int:
    if not reference.is_dir():
        return 0
    n = 0
    for rel in FILES:
        src = reference / rel
        if not src.exists():
            continue
        dst = DEST / rel
        dst.parent.mkdir(parents=True, exist_ok=True)
        shutil.copyfile(src, dst)
        n += 1
    return n


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--reference", default="/root/reference")
    n = stage(Path(ap.parse_args().reference))
    print(f"staged {n} reference files under {DEST}" if n else "reference tree not found: nothing staged")
    sys.exit(0)
