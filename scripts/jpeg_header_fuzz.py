"""Differential fuzz of JPEG headers (CPU only, no CUDA): mutate bytes before the scan and
compare three verdicts on every mutant —

  cv2.imdecode        what the reference's load_image_from_bytes would do (None -> ValueError)
  oracle/jpeg.py      the restated decoder (pixels compared with cv2 when both decode)
  ia_jpeg_parse       the library's host parser (ok / invalid / unsupported)

A mutant cv2 refuses but the parser accepts would make the drop-in return pixels where the
reference raises; one cv2 decodes and the parser calls invalid (not merely unsupported) would
raise where the reference returns pixels.  Both are printed with a reproducer.

    python scripts/jpeg_header_fuzz.py [iterations] [seed]
"""

from __future__ import annotations

import random
import sys
from pathlib import Path

import numpy as np

REPO = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(REPO), str(REPO / "inference-arena_b200")]

import cv2  # noqa: E402

from oracle import jpeg as J  # noqa: E402
from shared.processing import jpeg as P  # noqa: E402
from workloads import jpeg_cases as W  # noqa: E402


def bases():
    return [
        W.encode(W.photo(16, 24, seed=1), 80, "420"),
        W.encode(W.photo(16, 24, seed=2), 60, "444", optimize=True, restart=2),
        W.encode(W.photo(9, 11, seed=3)[..., 0], 70),
        W.with_exif_orientation(W.encode(W.photo(16, 24, seed=4), 80, "422"), 6),
        W.with_16bit_quant_tables(W.encode(W.photo(12, 20, seed=5), 50, "440")),
    ]


def verdict(fn, unsupported):
    try:
        return "ok", fn()
    except unsupported:
        return "unsupported", None
    except ValueError:
        return "invalid", None
    except Exception as e:  # a crash of the checker itself is a finding
        return "crash:" + type(e).__name__, None


def mutate(rng, base, scan):
    b = bytearray(base)
    for _ in range(rng.choice([1, 1, 1, 2, 3])):
        pos = rng.randrange(2, scan)
        b[pos] = rng.choice([rng.randrange(256), b[pos] ^ (1 << rng.randrange(8)), 0, 0xFF])
    return bytes(b)


def run(iterations=20000, seed=1, verbose=True):
    rng = random.Random(seed)
    files = [(f, P.parse_jpeg(f).scan_offset) for f in bases()]
    stats, findings = {}, []
    for _ in range(iterations):
        base, scan = rng.choice(files)
        data = mutate(rng, base, scan)
        ref = cv2.imdecode(np.frombuffer(data, np.uint8), cv2.IMREAD_COLOR)
        ours, _ = verdict(lambda: P.parse_jpeg(data), P.UnsupportedJpeg)
        orc, pixels = verdict(lambda: J.load_image_from_bytes(data), J.UnsupportedJpeg)
        key = ("cv2 ok" if ref is not None else "cv2 fails", "parser " + ours, "oracle " + orc)
        stats[key] = stats.get(key, 0) + 1
        problem = None
        if orc.startswith("crash") or ours.startswith("crash"):
            problem = "crash"
        elif ours != orc and ours != "unsupported":  # the parser may refuse more layouts than the oracle
            problem = "parser and oracle disagree"
        elif ref is None and ours == "ok":
            problem = "accepted a file cv2 refuses"
        elif ref is not None and ours == "invalid":
            problem = "called invalid a file cv2 decodes"
        elif ref is not None and orc == "ok":
            want = cv2.cvtColor(ref, cv2.COLOR_BGR2RGB)
            if pixels.shape != want.shape or not np.array_equal(pixels, want):
                problem = "pixels differ from cv2"
        if problem:
            findings.append((problem, key, data))
    if verbose:
        for k, v in sorted(stats.items(), key=lambda kv: -kv[1]):
            print(f"{v:7d}  {' | '.join(k)}")
        kinds = {}
        for p, _, _ in findings:
            kinds[p] = kinds.get(p, 0) + 1
        print("findings:", kinds or "none")
    return stats, findings


if __name__ == "__main__":
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
    s = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    _, found = run(n, s)
    seen = set()
    for problem, key, data in found:
        if (problem, key) in seen:
            continue
        seen.add((problem, key))
        print(problem, key, data.hex()[:0], len(data))
        out = REPO / "gpurun_out" / f"fuzz_{len(seen):02d}.jpg"
        out.parent.mkdir(exist_ok=True)
        out.write_bytes(data)
        print("  ->", out)
