"""Drop-in `shared` package: the B200 implementation of Inference Arena's shared
image-processing path.  Re-exports the processing names the reference's
``src/shared/__init__.py`` (:15-29) exposes; every arithmetic step runs in libia_b200.so on
the GPU.
"""

import shared.processing as _processing

__version__ = "0.1.0"
__all__ = ["YOLOPreprocessor", "MobileNetPreprocessor", "letterbox", "imagenet_normalize", "scale_boxes"]
globals().update({name: getattr(_processing, name) for name in __all__})
