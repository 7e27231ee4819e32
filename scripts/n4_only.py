"""Ad-hoc: the N4 softmax/top-5 stage alone (rows x 1000 logits), for tuning/profiling."""
import sys, torch
from pathlib import Path
R = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(R)); sys.path.insert(0, str(R / "inference-arena_b200"))
from shared.processing import device as D
dev = torch.device("cuda")
torch.manual_seed(0)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 50
rows = int(sys.argv[2]) if len(sys.argv) > 2 else 65536
sets = [torch.randn((rows, 1000), device=dev) * 4 for _ in range(2)]
for i in range(5): D.classify_topk_cuda(sets[i & 1], 5)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(n): D.classify_topk_cuda(sets[i & 1], 5)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / n
print(f"softmax top-5 {rows}x1000: {ms*1e3:.1f} us/call, {rows*4000/ms/1e6:.0f} GB/s")
