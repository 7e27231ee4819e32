"""Ad-hoc: host-side cost of one C-ABI call (launch overhead) and small-batch latency."""
import sys, time
from pathlib import Path
R = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(R)); sys.path.insert(0, str(R / "inference-arena_b200"))
import torch
from shared.processing import device as D
from shared.processing.pipeline import DevicePipeline
dev = torch.device("cuda")
f = torch.randint(0, 256, (1, 480, 640, 3), dtype=torch.uint8, device=dev)
out = torch.empty((1, 3, 640, 640), dtype=torch.float32, device=dev)
lib = D.lib()
s = torch.cuda.current_stream().cuda_stream
for _ in range(100): lib.ia_yolo_preprocess(f.data_ptr(), 1, 480, 640, 640, 255.0, out.data_ptr(), s)
torch.cuda.synchronize()
t0 = time.perf_counter()
N = 5000
for _ in range(N): lib.ia_yolo_preprocess(f.data_ptr(), 1, 480, 640, 640, 255.0, out.data_ptr(), s)
t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
print(f"ia_yolo_preprocess n=1: host {1e6*(t1-t0)/N:.2f} us/call, incl. drain {1e6*(t2-t0)/N:.2f} us/call")
for B in (1, 8, 32):
    frames = torch.randint(0, 256, (B, 1080, 1920, 3), dtype=torch.uint8, device=dev)
    head = torch.rand((B, 25200, 85), device=dev) * 0.2
    pipe = DevicePipeline(B, 1080, 1920, dev)
    for _ in range(20): pipe.run(frames, head)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(200): pipe.run(frames, head)
    torch.cuda.synchronize(); t1 = time.perf_counter()
    print(f"DevicePipeline B={B}: {1e6*(t1-t0)/200:.1f} us/step -> {B*200/(t1-t0):.0f} img/s")
from shared.processing.pipeline import GraphedPipeline
for B in (1, 8):
    gp = GraphedPipeline(DevicePipeline(B, 1080, 1920, dev), B)
    for _ in range(20): gp.replay()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(500): gp.replay()
    torch.cuda.synchronize(); t1 = time.perf_counter()
    print(f"GraphedPipeline B={B}: {1e6*(t1-t0)/500:.1f} us/step -> {B*500/(t1-t0):.0f} img/s")
