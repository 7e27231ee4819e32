#!/usr/bin/env python
"""What the box's pinned-memory copies alone sustain with the end-to-end byte mix, on 1/2/4/8 GPUs
at once: per GPU one thread issuing, on two streams, H2D copies of a frame+head chunk and D2H
copies of a tensor+crops chunk (the sizes bench.py's e2e leg moves per 32-image chunk), no
kernels.  The aggregate is the ceiling `e2e` can reach on this host; bench.py reports
`e2e.frac_of_copy_ceiling` against profiles/copy_ceiling.json.

    python scripts/copy_ceiling.py [--seconds 2] [--out profiles/copy_ceiling.json]
"""
import argparse
import json
import threading
import time
from pathlib import Path

import torch

R = Path(__file__).resolve().parent.parent
CHUNK = 32
H2D_PER_IMAGE = 1080 * 1920 * 3 + 25200 * 85 * 4      # frame + head
D2H_PER_IMAGE = 3 * 640 * 640 * 4 + 4 * 3 * 224 * 224 * 4  # model input + ~4 crops


def worker(dev: int, seconds: float, start: threading.Barrier, out: dict) -> None:
    torch.cuda.set_device(dev)
    up_h = torch.empty(CHUNK * H2D_PER_IMAGE, dtype=torch.uint8, pin_memory=True)
    dn_h = torch.empty(CHUNK * D2H_PER_IMAGE, dtype=torch.uint8, pin_memory=True)
    up_d = torch.empty_like(up_h, device=f"cuda:{dev}")
    dn_d = torch.empty_like(dn_h, device=f"cuda:{dev}")
    s_up, s_dn = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    for _ in range(2):
        with torch.cuda.stream(s_up):
            up_d.copy_(up_h, non_blocking=True)
        with torch.cuda.stream(s_dn):
            dn_h.copy_(dn_d, non_blocking=True)
    torch.cuda.synchronize(dev)
    start.wait()
    t0 = time.perf_counter()
    n = 0
    while time.perf_counter() - t0 < seconds:
        for _ in range(4):
            with torch.cuda.stream(s_up):
                up_d.copy_(up_h, non_blocking=True)
            with torch.cuda.stream(s_dn):
                dn_h.copy_(dn_d, non_blocking=True)
            n += 1
        s_up.synchronize()
        s_dn.synchronize()
    wall = time.perf_counter() - t0
    out[dev] = (n * CHUNK * H2D_PER_IMAGE / wall / 1e9, n * CHUNK * D2H_PER_IMAGE / wall / 1e9)


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--seconds", type=float, default=2.0)
    ap.add_argument("--out", default=str(R / "gpurun_out" / "copy_ceiling.json"))
    a = ap.parse_args()
    have = torch.cuda.device_count()
    result = {"chunk_images": CHUNK, "h2d_bytes_per_image": H2D_PER_IMAGE, "d2h_bytes_per_image": D2H_PER_IMAGE,
              "host_cpus": len(__import__("os").sched_getaffinity(0)), "by_gpus": {}}
    for g in (1, 2, 4, 8):
        if g > have:
            break
        out: dict = {}
        bar = threading.Barrier(g)
        ts = [threading.Thread(target=worker, args=(d, a.seconds, bar, out)) for d in range(g)]
        [t.start() for t in ts]
        [t.join() for t in ts]
        h2d, d2h = sum(v[0] for v in out.values()), sum(v[1] for v in out.values())
        result["by_gpus"][str(g)] = {"h2d_gbs": h2d, "d2h_gbs": d2h, "aggregate_gbs": h2d + d2h,
                                      "images_per_s_at_e2e_mix": (h2d + d2h) * 1e9 / (H2D_PER_IMAGE + D2H_PER_IMAGE)}
        print(g, "GPUs:", result["by_gpus"][str(g)], flush=True)
    Path(a.out).parent.mkdir(parents=True, exist_ok=True)
    Path(a.out).write_text(json.dumps(result, indent=1))


if __name__ == "__main__":
    main()
