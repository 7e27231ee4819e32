"""Ad-hoc: K1 on a ragged batch with a COCO-like mix of frame sizes (one launch per distinct shape)."""
import sys, torch, numpy as np
from pathlib import Path
R = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(R)); sys.path.insert(0, str(R / "inference-arena_b200"))
from shared.processing import device as D
dev = torch.device("cuda")
rng = np.random.default_rng(0)
shapes = [(480, 640), (427, 640), (640, 480), (426, 640), (375, 500), (640, 427), (500, 375), (333, 500),
          (480, 640), (428, 640), (612, 612), (360, 640), (640, 640), (424, 640), (478, 640), (500, 333)]
n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
distinct = int(sys.argv[2]) if len(sys.argv) > 2 else 16
picks = [shapes[int(rng.integers(0, distinct))] for _ in range(n)]
frames = [torch.randint(0, 256, (h, w, 3), dtype=torch.uint8, device=dev) for h, w in picks]
for _ in range(5): D.preprocess_ragged_cuda(frames)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
reps = 50
e0.record()
for _ in range(reps): D.preprocess_ragged_cuda(frames)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
by = sum(h * w * 3 for h, w in picks) + n * 3 * 640 * 640 * 4
print(f"ragged K1: {n} frames, {len(set(picks))} distinct shapes: {ms*1e3:.1f} us/call -> {n/ms*1e3:.0f} img/s, "
      f"{by/ms/1e6:.0f} GB/s algorithmic ({by/ms/1e6/6537:.2f} of HBM peak)")
