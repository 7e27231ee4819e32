"""numpy <-> device plumbing for the single-call drop-in functions (torch as the copier).

Results come back in page-locked memory: a pageable ``tensor.cpu()`` of the 4.9 MB model input
costs ~0.4 ms (the driver stages it through its own bounce buffer at ~12 GB/s), the same copy into
a pinned block ~0.1 ms.  The blocks come from torch's caching host allocator and the returned
array keeps its block alive, so the block is recycled — not freed — when the caller drops the
array.  A caller that hoards results (a curator caching thousands of tensors) would pin that
much RAM; above ``IA_PINNED_RESULT_BYTES`` (default 1 GiB) of results still alive, further ones
are ordinary pageable arrays.
"""

from __future__ import annotations

import os
import threading
import weakref

import numpy as np
import torch

_PINNED_CAP = int(os.environ.get("IA_PINNED_RESULT_BYTES", str(1 << 30)))
_PINNED_MIN = 64 << 10  # below this a pageable copy is as fast
_lock = threading.Lock()
_pinned_alive = 0


def device() -> torch.device:
    if not torch.cuda.is_available():
        raise RuntimeError("shared.processing needs a CUDA device: the hot path runs in "
                           "libia_b200.so kernels and has no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device())


def to_device(a: np.ndarray) -> torch.Tensor:
    return torch.from_numpy(np.ascontiguousarray(a)).to(device(), non_blocking=False)


def _release(nbytes: int) -> None:
    global _pinned_alive
    with _lock:
        _pinned_alive -= nbytes


def pinned_result_bytes() -> int:
    """Bytes of results handed out in page-locked memory and still referenced (for tests)."""
    with _lock:
        return _pinned_alive


def to_host(t: torch.Tensor) -> np.ndarray:
    global _pinned_alive
    nbytes = t.numel() * t.element_size()
    if nbytes >= _PINNED_MIN and t.is_cuda:
        with _lock:
            take = _pinned_alive + nbytes <= _PINNED_CAP
            if take:
                _pinned_alive += nbytes
        if take:
            try:
                out = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
                out.copy_(t, non_blocking=True)
                torch.cuda.current_stream(t.device).synchronize()
                arr = out.numpy()  # shares the block; `arr.base` keeps `out` alive
                # (a view taken by the caller may outlive `arr` itself: the block stays alive through
                # the view's base, only this count lets go early)
                weakref.finalize(arr, _release, nbytes)
                return arr
            except BaseException:
                _release(nbytes)
                raise
    return np.ascontiguousarray(t.cpu().numpy())
