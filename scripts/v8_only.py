"""Ad-hoc: the v8-layout decode+NMS stage alone (256 x 84 x 8400), for tuning/profiling."""
import sys, torch
from pathlib import Path
R = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(R)); sys.path.insert(0, str(R / "inference-arena_b200"))
from shared.processing import device as D
dev = torch.device("cuda")
torch.manual_seed(0)
h8 = torch.rand((256, 84, 8400), device=dev) * 0.3
h8[:, 0:2] = torch.rand((256, 2, 8400), device=dev) * 640
h8[:, 2:4] = 8 + torch.rand((256, 2, 8400), device=dev) * 192
sel = torch.rand((256, 8400), device=dev) < 0.012
cls = torch.randint(4, 84, (256, 8400), device=dev)
bi, ai = sel.nonzero(as_tuple=True)
h8[bi, cls[bi, ai], ai] = 0.55 + 0.4 * torch.rand(bi.numel(), device=dev)
ws = torch.empty(D.lib().ia_decode_nms_workspace_bytes(256, 8400, 4096), dtype=torch.uint8, device=dev)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 50
for _ in range(5): D.decode_nms_cuda(h8, 0.5, 0.45, layout="v8", max_candidates=4096, max_det=128, workspace=ws)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(n): D.decode_nms_cuda(h8, 0.5, 0.45, layout="v8", max_candidates=4096, max_det=128, workspace=ws)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / n
print(f"v8 decode+nms: {ms*1e3:.1f} us/call, {256*84*8400*4/ms/1e6:.0f} GB/s")
