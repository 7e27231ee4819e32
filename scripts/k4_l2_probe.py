"""Ad-hoc: the C5-like ROI set through ia_resize_normalize_rois with the frames spread over
256 images (HBM-resident) or 8 images (L2-resident): how much of K4's time is source-fetch latency."""
import sys
from pathlib import Path
R = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(R)); sys.path.insert(0, str(R / "inference-arena_b200"))
import numpy as np
import torch
from shared.processing import device as D
dev = torch.device("cuda")
rng = np.random.default_rng(5)
R_ = 1016
for n in (256, 8):
    frames = torch.randint(0, 256, (n, 1080, 1920, 3), dtype=torch.uint8, device=dev)
    wh = (180 + rng.random((R_, 2)) * 480).astype(np.int32)
    x1 = (rng.random(R_) * (1920 - wh[:, 0])).astype(np.int32)
    y1 = (rng.random(R_) * (1080 - wh[:, 1])).astype(np.int32)
    rois = np.zeros((R_, 8), np.int32)
    rois[:, 0] = rng.integers(0, n, R_)
    rois[:, 1], rois[:, 2], rois[:, 3], rois[:, 4] = x1, y1, x1 + wh[:, 0], y1 + wh[:, 1]
    rois_d = torch.from_numpy(rois).to(dev)
    for _ in range(3):
        out = D.resize_normalize_rois_cuda(frames, rois_d)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        out = D.resize_normalize_rois_cuda(frames, rois_d)
    e1.record(); torch.cuda.synchronize()
    print(f"{n} frames: {e0.elapsed_time(e1) / 20 * 1e3:.1f} us", flush=True)
    del frames
