"""`shared.processing` — B200 drop-in for the reference package of the same name
(reference src/shared/processing/__init__.py:14-34).  Same public names and call
signatures; the arithmetic runs in sm_100a CUDA kernels (libia_b200.so) reached through a
C-ABI.  There is no CPU fallback: without the library or a CUDA device the calls raise.

Batched, device-resident entry points live in ``shared.processing.device`` and
``shared.processing.pipeline``; the YOLO head post-processing (reference
architectures/*/app/postprocess.py) is in ``shared.processing.postprocess``.
"""

from shared.processing import mobilenet_preprocess as _mobilenet
from shared.processing import transforms as _transforms
from shared.processing import yolo_preprocess as _yolo

# the names the reference package exports, bound to the GPU-backed implementations
_EXPORTS = {
    "letterbox": _transforms, "imagenet_normalize": _transforms, "scale_boxes": _transforms,
    "load_image": _transforms, "load_image_from_bytes": _transforms,
    "YOLOPreprocessor": _yolo, "MobileNetPreprocessor": _mobilenet,
}
globals().update({name: getattr(module, name) for name, module in _EXPORTS.items()})
__all__ = list(_EXPORTS)
