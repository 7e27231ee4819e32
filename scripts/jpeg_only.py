"""Ad-hoc: JPEG decode stage alone — kernels on resident bytes, the whole call from host
bytes, and cv2.imdecode on one host core, for a batch of equal frames."""
import sys, time, torch, numpy as np
from pathlib import Path
R = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(R)); sys.path.insert(0, str(R / "inference-arena_b200"))
from shared.processing import jpeg as M
from workloads import jpeg_cases as W
import cv2
h, w = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (1080, 1920)
batch_n = int(sys.argv[3]) if len(sys.argv) > 3 else 64
q = int(sys.argv[4]) if len(sys.argv) > 4 else 90
reps = int(sys.argv[5]) if len(sys.argv) > 5 else 20
samp = sys.argv[6] if len(sys.argv) > 6 else "420"
blobs = [(W.encode(W.photo(h, w, seed=i)[..., 0], q) if samp == "gray" else W.encode(W.photo(h, w, seed=i), q, samp))
         for i in range(min(batch_n, 8))]
blobs = [blobs[i % len(blobs)] for i in range(batch_n)]
kb = sum(len(b) for b in blobs) / batch_n / 1024
batch = M.decode_jpeg_batch_cuda(blobs); torch.cuda.synchronize()
assert batch.status.cpu().sum().item() == 0
assert np.array_equal(batch.frames[0].cpu().numpy(), W.reference_decode(blobs[0]))
for _ in range(3): batch.rerun()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps): batch.rerun()
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
for _ in range(3):
    b2 = M.decode_jpeg_batch_cuda(blobs); torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(5):
    b2 = M.decode_jpeg_batch_cuda(blobs)
torch.cuda.synchronize()
call_ms = (time.perf_counter() - t0) / 5 * 1e3
t0 = time.perf_counter()
for b in blobs[:8]: cv2.cvtColor(cv2.imdecode(np.frombuffer(b, np.uint8), cv2.IMREAD_COLOR), cv2.COLOR_BGR2RGB)
cpu_ms = (time.perf_counter() - t0) / 8 * 1e3
print(f"{batch_n} x {w}x{h} q{q} ({kb:.0f} KB/file): kernels {ms:.3f} ms -> {batch_n/ms*1e3:.0f} img/s; "
      f"whole call {call_ms:.1f} ms -> {batch_n/call_ms*1e3:.0f} img/s; cv2 one core {cpu_ms:.2f} ms/img")
rec = batch.chain_records(0).cpu().numpy()
live = rec[rec[:, 3] == 1]
print("chunks", len(live), "rounds", (live[:, 10] & 0xffff).tolist(), "tasks", ((live[:, 10] >> 16) & 0xffff).tolist())
print("phase us [guess, repeats, repair, chain wait, write] per chunk:", live[:, 11:16].tolist())
