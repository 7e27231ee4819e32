"""Ad-hoc: coefficient-level comparison of cut-short scans, device vs oracle."""
import sys
from pathlib import Path
import numpy as np, torch
R = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(R), str(R / "inference-arena_b200"), str(R / "tests")]
from shared.processing import jpeg as M
from workloads import jpeg_cases as W
from oracle import jpeg as J
from test_gpu_jpeg import _scan_order_coefs
import cv2

samp, kw = sys.argv[1], ({"restart": int(sys.argv[2])} if len(sys.argv) > 2 and int(sys.argv[2]) else {})
img = W.photo(56, 88, seed=31)
data = W.encode(img[..., 0], 85) if samp == "gray" else W.encode(img, 85, samp, **kw)
scan = J.parse(data)["scan_start"]
body = len(data) - 2 - scan
bad = 0
for k in range(40):
    cut = scan + 1 + (k * 37) % (body - 1) if k < 30 else scan + body - (k - 29)
    tail = (b"\xff\xd9", b"\xff\xc4\x00\x02", b"\xff\xfe")[k % 3]
    blob = data[:cut] + tail
    ref = cv2.imdecode(np.frombuffer(blob, np.uint8), cv2.IMREAD_COLOR)
    if ref is None:
        continue
    batch = M.decode_jpeg_batch_cuda([blob]); torch.cuda.synchronize()
    got = batch.coefficients(0).cpu().numpy().astype(np.int64)
    want = _scan_order_coefs(blob)
    px = np.array_equal(batch.frames[0].cpu().numpy(), ref[..., ::-1])
    if not np.array_equal(got, want) or not px:
        bad += 1
        blocks = np.argwhere((got != want).any(-1))[:, 0]
        print(f"k {k} cut {cut - scan} status {batch.status.cpu().tolist()} pixels_equal {px} differing blocks {blocks[:8].tolist()} of {len(got)}")
        if len(blocks):
            b0 = blocks[0]
            print("  got ", got[b0][:16].tolist(), "\n  want", want[b0][:16].tolist())
print("bad", bad)
