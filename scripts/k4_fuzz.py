"""Ad-hoc differential fuzz of the tile kernel: random frame shapes (odd pitches, 1-pixel sides, wide
frames that get fewer teams), random rectangles and output sizes through ia_resize_normalize_rois and
the detection-driven ia_crop_resize_normalize, plus K1's linear mode on random shapes / targets /
batch sizes, all compared bit for bit with the oracle.  argv[1] = number of rounds (default 40),
argv[2] = seed (default 2026)."""
import sys
from pathlib import Path
R = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(R)); sys.path.insert(0, str(R / "inference-arena_b200"))
import numpy as np
import torch
from oracle import pipeline as O
from shared.processing import device as D
from workloads import synth

rounds = int(sys.argv[1]) if len(sys.argv) > 1 else 40
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 2026)
dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
bad = 0
for it in range(rounds):
    h = int(rng.choice([1, 2, 3, 37, 224, 448, 480, 1080, 2160, int(rng.integers(4, 1500))]))
    w = int(rng.choice([1, 2, 5, 61, 224, 448, 640, 1920, 3840, 5000, int(rng.integers(4, 2500))]))
    n = int(rng.integers(1, 4))
    frames = synth.frames(n, h, w, seed=7000 + it)
    S = int(rng.choice([224, 224, 224, 96, 225, 300]))
    rects = []
    for _ in range(int(rng.integers(1, 20))):
        x1, y1 = int(rng.integers(0, w)), int(rng.integers(0, h))
        x2 = min(w, x1 + int(rng.integers(1, max(2, w))))
        y2 = min(h, y1 + int(rng.integers(1, max(2, h))))
        empty = int(rng.random() < 0.1)
        rects.append([int(rng.integers(0, n)), x1, y1, x2, y2, empty, 0, 0])
    rois = np.array(rects, np.int32)
    got = D.resize_normalize_rois_cuda(dev(frames), dev(rois), S).cpu().numpy()
    for r, (img, x1, y1, x2, y2, empty, *_r) in enumerate(rois):
        crop = np.zeros((1, 1, 3), np.uint8) if empty else np.ascontiguousarray(frames[img, y1:y2, x1:x2])
        want = O.mobilenet_preprocess(crop, size=S)
        if not np.array_equal(got[r:r + 1], want):
            bad += 1
            print("K4 MISMATCH", (h, w, S), rois[r].tolist(), flush=True)
    # K1 on the same frames (linear / copy / box / gather as the shape decides)
    target = int(rng.choice([640, 640, 320, 416, 641, 1280]))
    sc = min(target / h, target / w)
    if int(h * sc) >= 1 and int(w * sc) >= 1:
        t = D.preprocess_batch_cuda(dev(frames), target=target).cpu().numpy()
        for i in range(n):
            want, *_ = O.yolo_preprocess(frames[i], target=target)
            if not np.array_equal(t[i:i + 1], want):
                bad += 1
                print("K1 MISMATCH", (h, w, target, i), flush=True)
torch.cuda.synchronize()
print(f"k4_fuzz: {rounds} rounds, {bad} mismatches")
sys.exit(1 if bad else 0)
