"""Ad-hoc: index arithmetic at large batch (2048 images) — spot-check against the oracle."""
import sys
from pathlib import Path
R = Path(__file__).resolve().parent.parent.parent
sys.path.insert(0, str(R)); sys.path.insert(0, str(R / "inference-arena_b200"))
import numpy as np, torch
from oracle import pipeline as O, synth
from shared.processing.pipeline import DevicePipeline
n, h, w = 2048, 480, 640
dev = torch.device("cuda")
base_f, base_h = synth.frames(8, h, w, seed=5), synth.head_v5(8, seed=6)
frames = torch.from_numpy(base_f).to(dev).repeat(n // 8, 1, 1, 1)
heads = torch.from_numpy(base_h).to(dev).repeat(n // 8, 1, 1)
pipe = DevicePipeline(n, h, w, dev)
out = pipe.run(frames, heads); torch.cuda.synchronize()
offs = out.roi_offsets.cpu().numpy()
want = [O.run_image(base_f[i], base_h[i], 0.25, 0.45) for i in range(8)]
for i in (0, 7, 1023, 1024, 2040, 2047):
    t, d, k, _, c = want[i % 8]
    kk = int(out.count[i])
    assert kk == len(k) and out.keep[i, :kk].cpu().tolist() == k
    assert np.array_equal(out.tensor[i:i+1].cpu().numpy(), t)
    assert np.array_equal(out.dets[i, :kk].cpu().numpy(), d)
    assert np.array_equal(out.crops[offs[i]:offs[i+1]].cpu().numpy(), c)
print("large batch ok:", n, "images,", int(offs[-1]), "rois, status max", int(out.status.max()))
