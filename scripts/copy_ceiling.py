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


def worker(dev: int, direction: str, seconds: float, start: threading.Barrier, out: dict) -> None:
    """One direction of one GPU: its own thread, its own stream, eight copies queued between
    synchronisations (the two directions of a GPU and all GPUs run at the same time)."""
    torch.cuda.set_device(dev)
    nbytes = CHUNK * (H2D_PER_IMAGE if direction == "h2d" else D2H_PER_IMAGE)
    host = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
    devb = torch.empty(nbytes, dtype=torch.uint8, device=f"cuda:{dev}")
    stream = torch.cuda.Stream(dev)
    src, dst = (host, devb) if direction == "h2d" else (devb, host)
    with torch.cuda.stream(stream):
        for _ in range(2):
            dst.copy_(src, non_blocking=True)
    stream.synchronize()
    start.wait()
    t0 = time.perf_counter()
    n = 0
    while time.perf_counter() - t0 < seconds:
        with torch.cuda.stream(stream):
            for _ in range(8):
                dst.copy_(src, non_blocking=True)
        n += 8
        stream.synchronize()
    wall = time.perf_counter() - t0
    out[(dev, direction)] = n * nbytes / wall / 1e9


def mix_worker(dev: int, seconds: float, start: threading.Barrier, out: dict) -> None:
    """The e2e leg minus its kernels: three streams per GPU, on each a chunk's inputs go up and then
    its outputs come down (the order HostPipeline issues them), chunks rotating over the streams."""
    torch.cuda.set_device(dev)
    up_h = [torch.empty(CHUNK * H2D_PER_IMAGE, dtype=torch.uint8, pin_memory=True) for _ in range(2)]
    dn_h = [torch.empty(CHUNK * D2H_PER_IMAGE, dtype=torch.uint8, pin_memory=True) for _ in range(2)]
    up_d = [torch.empty(CHUNK * H2D_PER_IMAGE, dtype=torch.uint8, device=f"cuda:{dev}") for _ in range(3)]
    dn_d = [torch.empty(CHUNK * D2H_PER_IMAGE, dtype=torch.uint8, device=f"cuda:{dev}") for _ in range(3)]
    streams = [torch.cuda.Stream(dev) for _ in range(3)]

    def chunk(i):
        with torch.cuda.stream(streams[i % 3]):
            up_d[i % 3].copy_(up_h[i % 2], non_blocking=True)
            dn_h[i % 2].copy_(dn_d[i % 3], non_blocking=True)

    for i in range(3):
        chunk(i)
    torch.cuda.synchronize(dev)
    start.wait()
    t0 = time.perf_counter()
    n = 0
    while time.perf_counter() - t0 < seconds:
        for _ in range(6):
            chunk(n)
            n += 1
        streams[n % 3].synchronize()  # keeps about two rounds of chunks queued
    torch.cuda.synchronize(dev)
    wall = time.perf_counter() - t0
    out[dev] = n * CHUNK / wall  # images/s of this GPU


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
        row = {}
        for mode, dirs in (("both", ("h2d", "d2h")), ("h2d_only", ("h2d",)), ("d2h_only", ("d2h",))):
            out: dict = {}
            bar = threading.Barrier(g * len(dirs))
            ts = [threading.Thread(target=worker, args=(d, k, a.seconds, bar, out)) for d in range(g) for k in dirs]
            [t.start() for t in ts]
            [t.join() for t in ts]
            row[mode] = {k: sum(v for (_, kk), v in out.items() if kk == k) for k in dirs}
        h2d, d2h = row["both"]["h2d"], row["both"]["d2h"]
        # the e2e leg needs both directions at once, in the ratio of its bytes: the slower one bounds it
        images = min(h2d * 1e9 / H2D_PER_IMAGE, d2h * 1e9 / D2H_PER_IMAGE)
        # each direction alone can go faster; with the e2e byte ratio the pair sustains at most:
        alone = min(row["h2d_only"]["h2d"] * 1e9 / H2D_PER_IMAGE, row["d2h_only"]["d2h"] * 1e9 / D2H_PER_IMAGE)
        out = {}
        bar = threading.Barrier(g)
        ts = [threading.Thread(target=mix_worker, args=(d, a.seconds, bar, out)) for d in range(g)]
        [t.start() for t in ts]
        [t.join() for t in ts]
        mix = sum(out.values())
        result["by_gpus"][str(g)] = {"e2e_mix_images_per_s": mix, "e2e_mix_h2d_gbs": mix * H2D_PER_IMAGE / 1e9,
                                      "e2e_mix_d2h_gbs": mix * D2H_PER_IMAGE / 1e9,
                                      "h2d_gbs": h2d, "d2h_gbs": d2h, "aggregate_gbs": h2d + d2h,
                                      "h2d_alone_gbs": row["h2d_only"]["h2d"], "d2h_alone_gbs": row["d2h_only"]["d2h"],
                                      "images_per_s_both_saturated": images,
                                      "images_per_s_upper_bound": alone}
        print(g, "GPUs:", result["by_gpus"][str(g)], flush=True)
    Path(a.out).parent.mkdir(parents=True, exist_ok=True)
    Path(a.out).write_text(json.dumps(result, indent=1))


if __name__ == "__main__":
    main()
