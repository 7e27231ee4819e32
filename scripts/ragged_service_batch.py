"""Ad-hoc: a service-like batch — 64 COCO-sized JPEG files of mixed shapes — through
RaggedPipeline (device decode -> K1 ragged -> K2+K3 -> K4 ragged), kernels only and from host bytes."""
import sys, time, torch, numpy as np
from pathlib import Path
R = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(R)); sys.path.insert(0, str(R / "inference-arena_b200"))
from shared.processing.pipeline import RaggedPipeline
from workloads import jpeg_cases as W, synth
dev = torch.device("cuda")
shapes = [(480, 640), (427, 640), (640, 480), (426, 640), (375, 500), (640, 427), (500, 375), (333, 500),
          (428, 640), (612, 612), (360, 640), (640, 640), (424, 640)]
n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
rng = np.random.default_rng(0)
base = [W.encode(W.photo(h, w, seed=i), 85, "420") for i, (h, w) in enumerate(shapes)]
files = [base[int(rng.integers(0, len(base)))] for _ in range(n)]
heads = torch.from_numpy(synth.head_v5(n, seed=1)).to(dev)
pipe = RaggedPipeline(dev)
for _ in range(3): out = pipe.run(files, heads)
torch.cuda.synchronize()
t0 = time.perf_counter(); reps = 10
for _ in range(reps): out = pipe.run(files, heads)
torch.cuda.synchronize()
ms = (time.perf_counter() - t0) / reps * 1e3
import cv2
t0 = time.perf_counter()
for f in files[:16]: cv2.cvtColor(cv2.imdecode(np.frombuffer(f, np.uint8), cv2.IMREAD_COLOR), cv2.COLOR_BGR2RGB)
cpu = (time.perf_counter() - t0) / 16 * 1e3
print(f"{n} mixed JPEG files ({sum(map(len, files))/n/1024:.0f} KB avg, {len(set(files))} shapes) -> chain: "
      f"{ms:.2f} ms/batch = {n/ms*1e3:.0f} requests/s from host bytes; cv2.imdecode alone {cpu:.2f} ms/file on one core; "
      f"kept {int(out.count.sum())} detections")
