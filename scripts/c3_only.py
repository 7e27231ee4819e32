"""Ad-hoc: the C3 stage alone (256 v5 heads -> K2 + K3) and the v8 one, for tuning/profiling.
argv[1] = repetitions (default 50)."""
import sys, torch
from pathlib import Path
R = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(R)); sys.path.insert(0, str(R / "inference-arena_b200"))
import bench
from shared.processing import device as D
dev = torch.device("cuda")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 50
_, head = bench.synth_inputs_device(256, dev, seed=2)
ws = torch.empty(D.lib().ia_decode_nms_workspace_bytes(256, bench.ANCHORS, 4096), dtype=torch.uint8, device=dev)
run = lambda: D.decode_nms_cuda(head, bench.CONF, bench.IOU, max_candidates=4096, max_det=32, workspace=ws)
for _ in range(5): run()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(n): run()
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / n
by = head.numel() * 4
print(f"C3 v5 decode+nms 256x25200x85: {ms*1e3:.1f} us/call, {by/ms/1e6:.0f} GB/s; dets/image {run()[2].float().mean().item():.2f}")
