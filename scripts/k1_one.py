"""Ad-hoc: one K1 shape, a few launches (for ncu): argv = h w [reps]."""
import sys
from pathlib import Path
R = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(R)); sys.path.insert(0, str(R / "inference-arena_b200"))
import torch
from shared.processing import device as D
h, w = int(sys.argv[1]), int(sys.argv[2])
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 4
dev = torch.device("cuda")
frames = torch.randint(0, 256, (64, h, w, 3), dtype=torch.uint8, device=dev)
out = torch.empty((64, 3, 640, 640), dtype=torch.float32, device=dev)
for _ in range(reps):
    D.preprocess_batch_cuda(frames, out=out)
torch.cuda.synchronize()
print("k1 done")
