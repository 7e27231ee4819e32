"""Ad-hoc: one request at a time through HostPipeline(max_batch=1) — host buffers in, host buffers
out, everything in one submit/wait — against the per-function numpy drop-in calls
(scripts/dropin_latency.py) on the same request.  Inputs: pageable numpy arrays (what a service
holds) and, for comparison, arrays that already live in pinned memory."""
import sys, time
from pathlib import Path
R = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(R)); sys.path.insert(0, str(R / "inference-arena_b200"))
import numpy as np, torch
from workloads import synth
from shared.processing.pipeline import HostPipeline
def t(fn, n=100):
    for _ in range(10): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): r = fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e3
for (h, w) in [(480, 640), (1080, 1920)]:
    frames = synth.frames(1, h, w, seed=0); head = synth.head_v5(1, seed=1, objects=4)
    pipe = HostPipeline(1, h, w, chunk=1, slots=2)
    fp = torch.from_numpy(frames).pin_memory(); hp = torch.from_numpy(head).pin_memory()
    def run(f, hd):
        res = pipe.submit(f, hd).wait()
        return res
    ft, ht = torch.from_numpy(frames), torch.from_numpy(head)  # pageable
    res = run(ft, ht)
    k = int(res["count"][0])
    a = t(lambda: run(ft, ht)); b = t(lambda: run(fp, hp))
    print(f"{h}x{w}: HostPipeline(max_batch=1) submit+wait {a:.3f} ms from pageable numpy, {b:.3f} ms from pinned "
          f"buffers ({k} detections; model input + {k} crops back on the host)")
