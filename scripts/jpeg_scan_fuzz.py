"""Differential fuzz of the entropy-coded part of JPEG files, device decoder vs cv2.imdecode
(what the reference's load_image_from_bytes calls).  Needs a GPU.

Damage classes: byte flips and bit flips in the data, bytes forced to FF, bytes deleted or
inserted, cuts (with and without a closing marker), and damage to the restart markers
themselves (renumbered, removed, duplicated, inserted).  For every class: how many mutants cv2
decodes, on how many of those the device returns the same pixels, and how the verdicts
(cv2 returns None  <->  status bit 1) line up.

    python scripts/jpeg_scan_fuzz.py [mutants_per_class] [seed]
"""

from __future__ import annotations

import os
import sys
from pathlib import Path

import numpy as np
import torch

REPO = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(REPO), str(REPO / "inference-arena_b200")]

import cv2  # noqa: E402

from shared.processing import jpeg as M  # noqa: E402
from workloads import jpeg_cases as W  # noqa: E402


def bases():
    out = []
    for i, (sampling, ri) in enumerate((("420", 0), ("420", 4), ("444", 1), ("422", 7), ("gray", 3), ("411", 0))):
        # IA_FUZZ_SIZE=HxW: larger frames put the damage into streams of several chained chunks
        h, w = (int(v) for v in os.environ.get("IA_FUZZ_SIZE", "96x152").split("x"))
        img = W.photo(h, w, seed=200 + i)
        data = W.encode(img[..., 0], 82, restart=ri) if sampling == "gray" else \
            W.encode(img, 82, sampling, restart=ri, optimize=bool(i & 1))
        out.append((data, M.parse_jpeg(data).scan_offset, ri))
    return out


def markers(data, start):
    return [i for i in range(start, len(data) - 2) if data[i] == 0xFF and 0xD0 <= data[i + 1] <= 0xD7]


def mutate(kind, rng, data, start, ri):
    b = bytearray(data)
    end = len(b) - 2  # keep the EOI
    pos = int(rng.integers(start, end))
    if kind == "byte":
        b[pos] = int(rng.integers(0, 256))
    elif kind == "bit":
        b[pos] ^= 1 << int(rng.integers(0, 8))
    elif kind == "ff":
        b[pos] = 0xFF
    elif kind == "delete":
        del b[pos:pos + int(rng.integers(1, 40))]
    elif kind == "insert":
        b[pos:pos] = bytes(rng.integers(0, 256, int(rng.integers(1, 40)), dtype=np.uint8))
    elif kind == "cut+eoi":
        b = b[:pos] + b"\xff\xd9"
    elif kind == "cut":
        b = b[:pos]
    else:
        rst = markers(b, start)
        if not rst:
            return None
        m = rst[int(rng.integers(0, len(rst)))]
        if kind == "rst renumber":
            b[m + 1] = 0xD0 + int(rng.integers(0, 8))
        elif kind == "rst remove":
            del b[m:m + 2]
        elif kind == "rst duplicate":
            b[m:m] = b[m:m + 2]
        elif kind == "rst insert":
            b[pos:pos] = bytes([0xFF, 0xD0 + int(rng.integers(0, 8))])
    return bytes(b)


KINDS = ("byte", "bit", "ff", "delete", "insert", "cut+eoi", "cut",
         "rst renumber", "rst remove", "rst duplicate", "rst insert")


def run(per_class=150, seed=1, verbose=True, dump=None):
    rng = np.random.default_rng(seed)
    files = bases()
    table = {}
    for kind in KINDS:
        blobs = []
        while len(blobs) < per_class:
            data, start, ri = files[int(rng.integers(0, len(files)))]
            m = mutate(kind, rng, data, start, ri)
            if m is not None:
                blobs.append(m)
        refs = [cv2.imdecode(np.frombuffer(b, np.uint8), cv2.IMREAD_COLOR) for b in blobs]
        batch = M.decode_jpeg_batch_cuda(blobs)
        torch.cuda.synchronize()
        status = batch.status.cpu().tolist()
        row = dict(n=len(blobs), cv2_decodes=0, same_pixels=0, flagged=0, differs_unflagged=0, cv2_fails=0, refused=0,
                   verdict_differs=0)
        for blob, ref, st, got in zip(blobs, refs, status, batch.frames):
            refused = bool(st & 2)
            if ref is None:
                row["cv2_fails"] += 1
                row["refused"] += refused
                row["verdict_differs"] += not refused
                continue
            row["cv2_decodes"] += 1
            row["verdict_differs"] += refused
            row["flagged"] += st != 0
            same = np.array_equal(got.cpu().numpy(), ref[..., ::-1])
            row["same_pixels"] += same
            if not same and dump is not None:
                dump.append((kind, blob))
            row["differs_unflagged"] += (not same) and st == 0  # must stay 0: status 0 promises cv2's pixels
        table[kind] = row
        if verbose:
            print(f"{kind:14s} {row}")
    return table


if __name__ == "__main__":
    differing = []
    run(int(sys.argv[1]) if len(sys.argv) > 1 else 150, int(sys.argv[2]) if len(sys.argv) > 2 else 1, dump=differing)
    out = REPO / "gpurun_out"
    out.mkdir(exist_ok=True)
    for i, (kind, blob) in enumerate(differing[:40]):  # for a look with the oracle on the CPU
        (out / f"scanfuzz_{i:02d}_{kind.replace(' ', '_').replace('+', '_')}.jpg").write_bytes(blob)
