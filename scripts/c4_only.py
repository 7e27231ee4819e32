"""Ad-hoc: BASELINE config C4 alone (512 frames 640x480, ~2031 ROIs) for profiling."""
import sys
from pathlib import Path
R = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(R)); sys.path.insert(0, str(R / "inference-arena_b200"))
import torch
from workloads import synth
from shared.processing import device as D
dev = torch.device("cuda")
frames = torch.randint(0, 256, (512, 480, 640, 3), dtype=torch.uint8, device=dev)
dets_np, count_np = synth.detections(512, seed=33, kmax=8)
dets, count = torch.from_numpy(dets_np).to(dev), torch.from_numpy(count_np).to(dev)
geom = D.letterbox_geometry(480, 640, 640)
crops = torch.empty((512 * 8, 3, 224, 224), dtype=torch.float32, device=dev)
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 5):
    D.crop_batch_cuda(frames, dets, count, geom, out=crops)
torch.cuda.synchronize()
print("c4 done")
