"""Ad-hoc: K4 alone on the C4 ROI set (640x480 frames, ~2031 ROIs) and on a C5-like one
(1080p frames, ~1035 ROIs of 180-660 px), device-resident; argv[1] = c4 | c5 | both, argv[2] = reps."""
import json
import sys
from pathlib import Path
R = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(R)); sys.path.insert(0, str(R / "inference-arena_b200"))
import numpy as np
import torch
from workloads import synth
from shared.processing import device as D
dev = torch.device("cuda")
which = sys.argv[1] if len(sys.argv) > 1 else "both"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 30
peak = json.loads((R / "MEASURED_PEAKS.json").read_text())["hbm_gbs"] if (R / "MEASURED_PEAKS.json").exists() else 6650.0


def run(name, n, h, w, dets_np, count_np):
    frames = torch.randint(0, 256, (n, h, w, 3), dtype=torch.uint8, device=dev)
    dets, count = torch.from_numpy(dets_np).to(dev), torch.from_numpy(count_np).to(dev)
    geom = D.letterbox_geometry(h, w, 640)
    crops = torch.empty((n * 8, 3, 224, 224), dtype=torch.float32, device=dev)
    _, offsets, rois = D.crop_batch_cuda(frames, dets, count, geom, out=crops)
    r = rois[: int(offsets[-1])].cpu().numpy()
    area = ((r[:, 3] - r[:, 1]) * (r[:, 4] - r[:, 2]) * (1 - r[:, 5]) + r[:, 5]).sum()
    by = int(area) * 3 + len(r) * 3 * 224 * 224 * 4
    for _ in range(3):
        D.crop_batch_cuda(frames, dets, count, geom, out=crops)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        D.crop_batch_cuda(frames, dets, count, geom, out=crops)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    print(f"{name}: {len(r)} ROIs, mean {np.sqrt(area / len(r)):.0f} px, {ms*1e3:.1f} us, "
          f"{by/ms/1e6:.0f} GB/s algorithmic = {by/ms/1e6/peak:.3f} of peak", flush=True)


if which in ("c4", "both"):
    d, c = synth.detections(512, seed=33, kmax=8)
    run("C4", 512, 480, 640, d, c)
if which in ("c5", "both"):
    rng = np.random.default_rng(5)
    n = 256
    count = rng.choice([3, 4, 5], size=n, p=[0.25, 0.5, 0.25]).astype(np.int32)
    dets = np.zeros((n, 8, 6), np.float32)
    ctr = 80 + rng.random((n, 8, 2)) * 480
    wh = 60 + rng.random((n, 8, 2)) * 160
    dets[..., 0:2] = ctr - wh / 2
    dets[..., 2:4] = ctr + wh / 2
    dets[..., 4] = 0.9
    run("C5-like", n, 1080, 1920, dets, count)
