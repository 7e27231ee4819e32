"""numpy <-> device plumbing for the single-call drop-in functions (torch as the copier)."""

from __future__ import annotations

import numpy as np
import torch


def device() -> torch.device:
    if not torch.cuda.is_available():
        raise RuntimeError("shared.processing needs a CUDA device: the hot path runs in "
                           "libia_b200.so kernels and has no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device())


def to_device(a: np.ndarray) -> torch.Tensor:
    return torch.from_numpy(np.ascontiguousarray(a)).to(device(), non_blocking=False)


def to_host(t: torch.Tensor) -> np.ndarray:
    return np.ascontiguousarray(t.cpu().numpy())
