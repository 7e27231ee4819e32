"""pytest configuration: the ``gpu`` marker and shared helpers.

``-m "not gpu"`` runs everywhere (oracle vs golden vectors, host logic, C-ABI symbol
export); ``-m gpu`` needs a B200 and calls the CUDA path through the C-ABI.
"""

from __future__ import annotations

import hashlib
import json
import sys
from pathlib import Path

import numpy as np
import pytest

REPO = Path(__file__).resolve().parent.parent
if str(REPO) not in sys.path:
    sys.path.insert(0, str(REPO))
PKG = REPO / "inference-arena_b200"
if str(PKG) not in sys.path:
    sys.path.insert(0, str(PKG))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def sha(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def strided_sample(a: np.ndarray, n: int = 257) -> np.ndarray:
    flat = np.ascontiguousarray(a).reshape(-1)
    idx = np.linspace(0, flat.size - 1, num=min(n, flat.size)).astype(np.int64)
    return flat[idx].copy()


class Golden:
    """Reference-generated vectors (tests/golden/make_golden.py)."""

    def __init__(self) -> None:
        self.arrays = np.load(REPO / "tests" / "golden" / "golden.npz")
        self.meta = json.loads(bytes(self.arrays["meta_json"]).decode())

    def __getitem__(self, key: str) -> np.ndarray:
        return self.arrays[key]


@pytest.fixture(scope="session")
def golden() -> Golden:
    return Golden()
