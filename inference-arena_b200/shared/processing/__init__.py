"""`shared.processing` — B200 drop-in for the reference package of the same name
(reference src/shared/processing/__init__.py:14-34).  Same public names and call
signatures; the arithmetic runs in sm_100a CUDA kernels (libia_b200.so) reached through a
C-ABI.  There is no CPU fallback: without the library or a CUDA device the calls raise.

Batched, device-resident entry points live in ``shared.processing.device``; the YOLO head
post-processing (reference architectures/*/app/postprocess.py) is in
``shared.processing.postprocess``.
"""

from shared.processing.mobilenet_preprocess import MobileNetPreprocessor
from shared.processing.transforms import (
    imagenet_normalize,
    letterbox,
    load_image,
    load_image_from_bytes,
    scale_boxes,
)
from shared.processing.yolo_preprocess import YOLOPreprocessor

__all__ = [
    "letterbox",
    "imagenet_normalize",
    "scale_boxes",
    "load_image",
    "load_image_from_bytes",
    "YOLOPreprocessor",
    "MobileNetPreprocessor",
]
