"""MobileNetV2 input preparation — same interface as the reference's
``src/shared/processing/mobilenet_preprocess.py`` (MobileNetPreprocessResult :40-50,
MobileNetPreprocessor :58-233, extract_crop :236-269).  Resize + ImageNet normalise + CHW
run in the K4 kernel with the crop as a single whole-image ROI.
"""

from __future__ import annotations

import functools
from dataclasses import dataclass

import numpy as np
import torch

from shared.config import get_controlled_variable
from shared.processing import _hostio, device as _dev
from shared.processing.transforms import IMAGENET_MEAN, IMAGENET_STD

_mobilenet_config = get_controlled_variable("preprocessing", "mobilenet")
MOBILENET_INPUT_SIZE: int = _mobilenet_config["target_size"]


@dataclass
class MobileNetPreprocessResult:
    """tensor [1,3,S,S] float32 (ImageNet-normalised) and the crop's (height, width)."""

    tensor: np.ndarray
    original_shape: tuple[int, int]


@functools.lru_cache(maxsize=1024)
def _whole_image_roi_on(device_index: int, height: int, width: int) -> torch.Tensor:
    """The one-rectangle table {image 0, x1 0, y1 0, x2 w, y2 h, not empty} of a (height, width)
    crop, kept on the device: the kernel only reads it, and a service sees the same few hundred
    crop shapes again and again."""
    row = torch.tensor([[0, 0, 0, width, height, 0, 0, 0]], dtype=torch.int32)
    return row.to(torch.device("cuda", device_index))


class MobileNetPreprocessor:
    """Resize a crop to ``input_size`` (bilinear), normalise with the ImageNet statistics,
    CHW, batch dim."""

    def __init__(self, input_size: int = MOBILENET_INPUT_SIZE, mean: np.ndarray = IMAGENET_MEAN,
                 std: np.ndarray = IMAGENET_STD) -> None:
        self.input_size = input_size
        self.mean = mean
        self.std = std

    def __call__(self, crop: np.ndarray) -> MobileNetPreprocessResult:
        return self.preprocess(crop)

    def preprocess(self, crop: np.ndarray) -> MobileNetPreprocessResult:
        self._validate_input(crop)
        height, width = crop.shape[0], crop.shape[1]
        if crop.dtype == np.float32:  # OpenCV's float resize + the float branch of imagenet_normalize
            tensor = _dev.resize_normalize_f32_cuda(_hostio.to_device(crop), self.input_size, IMAGENET_MEAN,
                                                    IMAGENET_STD)
            return MobileNetPreprocessResult(tensor=_hostio.to_host(tensor), original_shape=(height, width))
        frames = _hostio.to_device(crop)[None]
        rois = _whole_image_roi_on(frames.device.index, height, width)
        # like the reference (mobilenet_preprocess.py:146 -> transforms.py:270) the module-level
        # ImageNet statistics are applied, not self.mean / self.std
        tensor = _dev.resize_normalize_rois_cuda(frames, rois, self.input_size, IMAGENET_MEAN, IMAGENET_STD)
        return MobileNetPreprocessResult(tensor=_hostio.to_host(tensor), original_shape=(height, width))

    def preprocess_batch(self, crops: list[np.ndarray]) -> np.ndarray:
        if not crops:
            return np.zeros((0, 3, self.input_size, self.input_size), dtype=np.float32)
        return np.concatenate([self.preprocess(c).tensor for c in crops], axis=0)

    def _validate_input(self, crop: np.ndarray) -> None:
        if not isinstance(crop, np.ndarray):
            raise ValueError(f"Expected numpy array, got {type(crop)}")
        if crop.ndim != 3:
            raise ValueError(f"Expected 3D array [H, W, C], got {crop.ndim}D")
        if crop.shape[2] != 3:
            raise ValueError(f"Expected 3 channels, got {crop.shape[2]}")
        if crop.dtype not in (np.uint8, np.float32):
            raise ValueError(f"Expected uint8 or float32 dtype, got {crop.dtype}")
        if crop.shape[0] < 1 or crop.shape[1] < 1:
            raise ValueError(f"Invalid crop dimensions: {crop.shape[:2]}")

    @staticmethod
    def get_input_shape() -> tuple[int, int, int, int]:
        return (1, 3, MOBILENET_INPUT_SIZE, MOBILENET_INPUT_SIZE)

    @staticmethod
    def get_input_dtype() -> np.dtype:
        return np.dtype(np.float32)


def extract_crop(image: np.ndarray, box: np.ndarray) -> np.ndarray:
    """Copy the region ``box[:4] = [x1, y1, x2, y2]`` (truncated toward zero, clamped to the
    frame) out of ``image``; an empty region yields ``zeros((1, 1, 3))``.

    The four coordinates are truncated on the host in the caller's own dtype, exactly as the
    reference's ``map(int, box[:4])`` does (mobilenet_preprocess.py:256): a float64 99.99999999
    stays 99, integer boxes pass through, NaN / inf raise what ``int()`` raises.  (Device-resident
    float32 boxes take ``extract_crop_rects_cuda`` / the fused K4 kernel instead.)  The pixels
    of a host image are then a plain host-side slice copy — moving the whole frame to the GPU
    and back just to copy a window of it would only add PCIe traffic."""
    height, width = image.shape[:2]
    left, top, right, bottom = (int(v) for v in box[:4])
    left, top = max(left, 0), max(top, 0)
    right, bottom = min(right, width), min(bottom, height)
    if right <= left or bottom <= top:
        return np.zeros((1, 1, 3), dtype=np.uint8)
    return image[top:bottom, left:right].copy()
