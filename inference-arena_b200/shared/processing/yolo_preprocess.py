"""YOLOv5n input preparation — same interface as the reference's
``src/shared/processing/yolo_preprocess.py`` (YOLOPreprocessResult :44-69,
YOLOPreprocessor :77-213).  ``preprocess`` runs the fused K1 kernel (letterbox + /255 +
HWC->CHW) on the GPU; ``preprocess_batch`` is the additive batched form.
"""

from __future__ import annotations

from dataclasses import dataclass

import numpy as np

from shared.config import get_controlled_variable
from shared.processing import _hostio, device as _dev
from shared.processing.transforms import scale_boxes

_yolo_config = get_controlled_variable("preprocessing", "yolo")
YOLO_INPUT_SIZE: int = _yolo_config["target_size"]
YOLO_NORMALIZATION_SCALE: float = _yolo_config["normalization_scale"]


@dataclass
class YOLOPreprocessResult:
    """tensor [1,3,S,S] float32 in [0,1]; scale / padding / original_shape for mapping
    detections back (reference yolo_preprocess.py:44-69)."""

    tensor: np.ndarray
    scale: float
    padding: tuple[int, int]
    original_shape: tuple[int, int]

    def scale_boxes_to_original(self, boxes: np.ndarray) -> np.ndarray:
        return scale_boxes(boxes, self.scale, self.padding, self.original_shape)


class YOLOPreprocessor:
    """Letterbox to ``input_size``, divide by ``normalization_scale``, CHW, batch dim."""

    def __init__(self, input_size: int = YOLO_INPUT_SIZE,
                 normalization_scale: float = YOLO_NORMALIZATION_SCALE) -> None:
        self.input_size = input_size
        self.normalization_scale = normalization_scale

    def __call__(self, image: np.ndarray) -> YOLOPreprocessResult:
        return self.preprocess(image)

    def preprocess(self, image: np.ndarray) -> YOLOPreprocessResult:
        self._validate_input(image)
        height, width = image.shape[0], image.shape[1]
        geom = _dev.letterbox_geometry(height, width, self.input_size)
        frames = _hostio.to_device(image)[None]
        tensor = _dev.preprocess_batch_cuda(frames, self.input_size, self.normalization_scale)
        return YOLOPreprocessResult(tensor=_hostio.to_host(tensor), scale=geom.scale,
                                    padding=geom.padding, original_shape=(height, width))

    def preprocess_batch(self, frames: np.ndarray) -> np.ndarray:
        """uint8 [N,H,W,3] -> float32 [N,3,S,S] in one launch (host arrays in and out)."""
        if not isinstance(frames, np.ndarray) or frames.ndim != 4:
            raise ValueError(f"Expected 4D array [N, H, W, C], got {getattr(frames, 'ndim', '?')}D")
        for frame in frames[:1]:
            self._validate_input(frame)
        out = _dev.preprocess_batch_cuda(_hostio.to_device(frames), self.input_size,
                                         self.normalization_scale)
        return _hostio.to_host(out)

    def _validate_input(self, image: np.ndarray) -> None:
        if not isinstance(image, np.ndarray):
            raise ValueError(f"Expected numpy array, got {type(image)}")
        if image.ndim != 3:
            raise ValueError(f"Expected 3D array [H, W, C], got {image.ndim}D")
        if image.shape[2] != 3:
            raise ValueError(f"Expected 3 channels, got {image.shape[2]}")
        if image.dtype != np.uint8:
            raise ValueError(f"Expected uint8 dtype, got {image.dtype}")

    @staticmethod
    def get_input_shape() -> tuple[int, int, int, int]:
        return (1, 3, YOLO_INPUT_SIZE, YOLO_INPUT_SIZE)

    @staticmethod
    def get_input_dtype() -> np.dtype:
        return np.dtype(np.float32)
