"""Ad-hoc: K1 throughput across frame shapes / resize modes (batch 64, device-resident)."""
import sys, json
from pathlib import Path
R = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(R)); sys.path.insert(0, str(R / "inference-arena_b200"))
import torch
from shared.processing import device as D
dev = torch.device("cuda")
peak = json.loads((R / "MEASURED_PEAKS.json").read_text())["hbm_gbs"] if (R / "MEASURED_PEAKS.json").exists() else 6650.0
for name, h, w in [("copy 640x480", 480, 640), ("gather 1080p", 1080, 1920), ("box2 720p", 720, 1280),
                   ("linear up 500x375", 375, 500), ("linear down 1024x768", 768, 1024),
                   ("linear 6x 4K", 2160, 3840), ("copy 640x427", 427, 640), ("portrait 640x480", 640, 480)]:
    n = 64
    sets = [torch.randint(0, 256, (n, h, w, 3), dtype=torch.uint8, device=dev) for _ in range(3)]
    out = torch.empty((n, 3, 640, 640), dtype=torch.float32, device=dev)
    for i in range(5): D.preprocess_batch_cuda(sets[i % 3], out=out)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(30): D.preprocess_batch_cuda(sets[i % 3], out=out)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 30
    by = n * (h * w * 3 + 3 * 640 * 640 * 4)
    print(f"{name:24s} {ms*1e3:8.1f} us  {n/ms*1e3/1e3:8.1f} k img/s  algorithmic {by/ms/1e6:7.0f} GB/s = {by/ms/1e6/peak:.2f} of peak")
