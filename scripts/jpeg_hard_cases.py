"""Ad-hoc: streams that are hard for the self-synchronising decode — flat and periodic images whose
codes repeat with a short period — checked against cv2 and timed; prints the repeat rounds per chunk."""
import sys, time, torch, numpy as np
from pathlib import Path
R = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(R)); sys.path.insert(0, str(R / "inference-arena_b200"))
from shared.processing import jpeg as M
from workloads import jpeg_cases as W
h, w = 4096, 4096
yy, xx = np.mgrid[0:h, 0:w]
cases = {
    "flat": np.full((h, w, 3), 200, np.uint8),
    "checker8": (((xx // 8 + yy // 8) & 1) * 255).astype(np.uint8)[..., None].repeat(3, -1),
    "stripes16": (((xx // 16) & 1) * 180 + 40).astype(np.uint8)[..., None].repeat(3, -1),
    "gradient": ((xx + yy) // 32).astype(np.uint8)[..., None].repeat(3, -1),
    "period-noise": np.tile(np.random.default_rng(0).integers(0, 256, (16, 16, 3), dtype=np.uint8), (h // 16, w // 16, 1)),
}
for name, img in cases.items():
    for samp in ("420", "444"):
        data = W.encode(img, 90, samp)
        batch = M.decode_jpeg_batch_cuda([data]); torch.cuda.synchronize()
        ok = np.array_equal(batch.frames[0].cpu().numpy(), W.reference_decode(data))
        rec = batch.chain_records(0).cpu().numpy(); live = rec[rec[:, 3] == 1]
        rounds = (live[:, 10] & 0xffff).tolist()
        t0 = time.perf_counter(); batch.rerun(); torch.cuda.synchronize(); ms = (time.perf_counter() - t0) * 1e3
        print(f"{name:13s} {samp} {len(data)/1024:7.0f} KB  match={ok} status={batch.status.cpu().tolist()} "
              f"chunks={len(live)} rounds max={max(rounds)} {ms:.2f} ms")
