#!/usr/bin/env python
"""The e2e leg through ShardedHostPipeline: ONE process, one host batch of G x 256 1080p frames +
heads spread over G GPUs (threads, no torch.distributed), beside the copy ceiling.  Prints images/s
for G = 1, 2, 4, 8 (as many as the box has)."""
import json
import sys
import time
from pathlib import Path
R = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(R)); sys.path.insert(0, str(R / "inference-arena_b200"))
import torch
from shared.processing.pipeline import ShardedHostPipeline

have = torch.cuda.device_count()
per_gpu, steps = 256, 4
res = {}
from workloads import synth
frames = torch.randint(0, 256, (per_gpu * have, 1080, 1920, 3), dtype=torch.uint8).pin_memory()
# the bench's heads: ~4 detections per image, so the crops come back too (7.35 MB down per image)
one = torch.from_numpy(synth.head_v5(per_gpu, seed=5))
heads = torch.empty((per_gpu * have, 25200, 85), dtype=torch.float32).pin_memory()
for g in range(have):
    heads[g * per_gpu:(g + 1) * per_gpu].copy_(one)
for g in (1, 2, 4, 8):
    if g > have:
        break
    n = per_gpu * g
    sp = ShardedHostPipeline(n, 1080, 1920, devices=list(range(g)), chunk=32, slots=3, max_det=32,
                             max_candidates=4096, rois_per_image=8)
    for tth in (True, False):
        for _ in range(2):
            sp.run(frames[:n], heads[:n], tensors_to_host=tth)
        t0 = time.perf_counter()
        tk = None
        for _ in range(steps):
            nx = sp.submit(frames[:n], heads[:n], tensors_to_host=tth)
            if tk is not None:
                tk.wait()
            tk = nx
        tk.wait()
        wall = time.perf_counter() - t0
        res[f"{g}_{'host_out' if tth else 'device_out'}"] = n * steps / wall
        print(g, "GPUs", "tensors_to_host" if tth else "models_on_gpu", f"{n * steps / wall:.0f} images/s", flush=True)
    sp.close()
    del sp
(R / "gpurun_out").mkdir(exist_ok=True)
(R / "gpurun_out" / "sharded_e2e.json").write_text(json.dumps(res, indent=1))
