"""Drop-in `shared` package: the B200 implementation of Inference Arena's shared
image-processing path.  Same import surface as the reference's ``src/shared/__init__.py``
(:15-29) for the processing names; every arithmetic step runs in libia_b200.so on the GPU.
"""

from shared.processing import (
    MobileNetPreprocessor,
    YOLOPreprocessor,
    imagenet_normalize,
    letterbox,
    scale_boxes,
)

__all__ = [
    "YOLOPreprocessor",
    "MobileNetPreprocessor",
    "letterbox",
    "imagenet_normalize",
    "scale_boxes",
]

__version__ = "0.1.0"
