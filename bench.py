#!/usr/bin/env python
"""bench.py — images/sec of the shared.processing hot path on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--batch B] [--impl reference]

A step is one pass of the whole path (K1 letterbox+normalise+CHW -> K2 decode+filter ->
K3 NMS -> K4 ROI crop-resize-normalise) over one batch of B synthetic 1920x1080 frames and
(B,25200,85) YOLOv5n heads per GPU (config C5 of BASELINE.json — the configuration the
headline metric "preprocess+NMS+crop" is quoted on; it fits one GPU).  Prints ONE JSON
line (rank 0).  See DESIGN.md §6 for the definition of every key.

  value      images/s with inputs resident in HBM (CUDA events, max over ranks)
  e2e        images/s through HostPipeline: pinned HOST buffers in/out, H2D + D2H timed
  roofline   dominant kernel: algorithmic bytes / CUDA-event duration vs MEASURED_PEAKS.json
  stages     the single-stage configs C2 / C3 / C4 of BASELINE.json, each vs the HBM roofline
  cpu_baseline  the CPU oracle port of the reference path (numpy + cv2) on the host cores
  --impl reference   times that CPU path alone, same metric/config, "impl": "reference"
"""

from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time
from pathlib import Path

REPO = Path(__file__).resolve().parent
PKG = REPO / "inference-arena_b200"
for _p in (str(REPO), str(PKG)):
    if _p not in sys.path:
        sys.path.insert(0, _p)

METRIC = "images/sec preprocess+NMS+crop"
UNIT = "images/s"
SRC_H, SRC_W = 1080, 1920
ANCHORS, CLASSES = 25200, 80
CONF, IOU = 0.25, 0.45
FALLBACK_HBM_GBS = 6650.0  # B200_PROFILING.md fallback when MEASURED_PEAKS.json is absent


def workload_config(batch: int, n_gpus: int) -> dict:
    return {
        "workload": f"C5 full chain: {batch} x 1920x1080 uint8 frames + ({batch},25200,85) f32 heads per GPU "
                    f"-> letterbox640+/255+CHW -> decode+NMS(conf {CONF}, IoU {IOU}) -> ~4 crops/image at 224^2",
        "batch_per_gpu": batch, "global_batch": batch * n_gpus, "frame": [SRC_H, SRC_W, 3],
        "head": [ANCHORS, 5 + CLASSES], "sharding": f"dp{n_gpus}, independent images, no collective",
        "l2_policy": "inputs+outputs per step (>6 GB at B=256) exceed the 126 MB L2; stage benches rotate buffers",
    }


def hbm_peak() -> tuple[float, str]:
    p = REPO / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return float(json.loads(p.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md)"


# =============================================================================== CPU arm
_W = {}


def _cpu_worker_init(images: int, seed_base: int) -> None:
    """Each worker owns `images` synthetic requests; cv2 single-threaded per worker
    (BASELINE.md §3 mode ii: one process per host core).  The worker runs the reference's own
    modules when they are staged (baseline/_ref), else the oracle port of them."""
    import cv2

    from workloads import synth

    cv2.setNumThreads(1)
    try:
        from oracle import refrun
        _W["reference"] = refrun.Reference()
    except Exception:
        _W["reference"] = None
    seed = seed_base + os.getpid() % 100000
    _W["frames"] = synth.frames(images, SRC_H, SRC_W, seed=seed)
    _W["heads"] = synth.head_v5(images, seed=seed + 1)


def _cpu_worker_step(reps: int):
    from oracle import pipeline as O

    ref = _W.get("reference")
    t0 = time.perf_counter()
    done = 0
    for _ in range(reps):
        for f, h in zip(_W["frames"], _W["heads"]):
            if ref is not None:
                ref.run_image(f, h, CONF, IOU)
            else:
                O.run_image(f, h, CONF, IOU, resize=O.cv2_resize)
            done += 1
    return done, time.perf_counter() - t0


class CpuArm:
    """The reference's CPU implementation of the path — the oracle port (numpy + cv2, the
    same library calls the reference makes) — on all host cores, one process per core."""

    def __init__(self, images_per_worker: int = 4) -> None:
        import multiprocessing as mp

        self.cores = len(os.sched_getaffinity(0))
        self.images_per_worker = images_per_worker
        from oracle import refrun
        self.reference_root = refrun.find_reference()
        ctx = mp.get_context("fork")
        self.pool = ctx.Pool(self.cores, initializer=_cpu_worker_init, initargs=(images_per_worker, 1000))

    def step(self, reps: int = 1) -> tuple[int, float]:
        t0 = time.perf_counter()
        res = self.pool.map(_cpu_worker_step, [reps] * self.cores, chunksize=1)
        wall = time.perf_counter() - t0
        return sum(r[0] for r in res), wall

    def close(self) -> None:
        self.pool.close()
        self.pool.join()

    def describe(self, images: int, wall: float) -> dict:
        import cv2
        import numpy as np

        if self.reference_root is not None:
            kind, what = "reference", (f"the reference's own modules from {self.reference_root.name} "
                                       "(YOLOPreprocessor, curator.py:295-304 decode restated + its apply_nms, "
                                       "scale_boxes, extract_crop, MobileNetPreprocessor) via oracle/refrun.py")
        else:
            kind, what = "port", "oracle/pipeline.py run_image with cv2.resize (reference files not staged)"
        return {"value": images / wall, "unit": UNIT, "cores": self.cores, "kind": kind,
                "sample": f"{images} requests of the same C5 workload (1080p frame + 25200x85 head -> full chain) "
                          f"in {wall:.2f} s wall over {self.cores} processes, cv2 {cv2.__version__} (1 thread each), "
                          f"numpy {np.__version__}; {what}"}


def run_reference_arm(args) -> None:
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    arm = CpuArm(images_per_worker=4)
    reps = 2  # each worker runs its 4 requests twice per step: pool dispatch stays negligible
    for _ in range(max(args.warmup, 1)):
        arm.step(reps)
    total, wall_total, per_step = 0, 0.0, []
    for _ in range(args.steps):
        n, wall = arm.step(reps)
        total += n
        wall_total += wall
        per_step.append(wall)
    arm_desc = arm.describe(total, wall_total)
    arm.close()
    value = total / wall_total
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * wall_total / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8/f32",
        "data": "synthetic", "config": workload_config(args.batch, args.gpus),
        "cpu_baseline": arm_desc | {"value": value, "cpu_step": f"{arm.cores * 4 * reps} requests per timed step"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


# ============================================================================== GPU arm
def synth_inputs_device(batch: int, dev, seed: int):
    """Device-side twin of workloads/synth.py (frames: uniform noise; heads: low-confidence
    background + k in {3,4,5} planted objects x 30 jittered anchors per image)."""
    import torch

    g = torch.Generator(device=dev)
    g.manual_seed(seed)
    frames = torch.randint(0, 256, (batch, SRC_H, SRC_W, 3), dtype=torch.uint8, device=dev, generator=g)
    head = torch.empty((batch, ANCHORS, 5 + CLASSES), dtype=torch.float32, device=dev)
    head[..., 0:2] = torch.rand((batch, ANCHORS, 2), device=dev, generator=g) * 640
    head[..., 2:4] = 8 + torch.rand((batch, ANCHORS, 2), device=dev, generator=g) * 192
    head[..., 4] = torch.rand((batch, ANCHORS), device=dev, generator=g) * 0.05
    head[..., 5:] = torch.rand((batch, ANCHORS, CLASSES), device=dev, generator=g) * 0.3
    dup = 30
    k = torch.multinomial(torch.tensor([0.25, 0.5, 0.25], device=dev), batch, replacement=True, generator=g) + 3
    slots = torch.rand((batch, ANCHORS), device=dev, generator=g).topk(5 * dup, dim=1).indices  # distinct anchors
    for o in range(5):
        rows = slots[:, o * dup:(o + 1) * dup]
        active = (k > o)
        centre = 80 + torch.rand((batch, 1, 2), device=dev, generator=g) * 480
        wh = 60 + torch.rand((batch, 1, 2), device=dev, generator=g) * 160
        cls = torch.randint(0, CLASSES, (batch,), device=dev, generator=g)
        box = torch.cat([centre + 3 * torch.randn((batch, dup, 2), device=dev, generator=g),
                         wh + 4 * torch.randn((batch, dup, 2), device=dev, generator=g)], dim=2)
        obj = 0.5 + 0.49 * torch.rand((batch, dup), device=dev, generator=g)
        score = 0.6 + 0.39 * torch.rand((batch, dup), device=dev, generator=g)
        bi = torch.arange(batch, device=dev)[:, None].expand(-1, dup)
        sel = active[:, None].expand(-1, dup)
        head[bi[sel], rows[sel], 0:4] = box[sel]
        head[bi[sel], rows[sel], 4] = obj[sel]
        head[bi[sel], rows[sel], (5 + cls)[:, None].expand(-1, dup)[sel]] = score[sel]
    return frames, head


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md)."""

    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int) -> None:
        self.proc = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={gpu_index}", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                 "-lms", "10"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except Exception:
            self.proc.kill()
            out = ""
        sm, smax, power, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in out.strip().splitlines():
            parts = [p.strip() for p in ln.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0]))
                smax.append(float(parts[1]))
                power.append(float(parts[2]))
            except ValueError:
                continue
            for name, val in zip(names, parts[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


def time_loop(fn, steps: int, warmup: int, dev):
    import torch

    for _ in range(warmup):
        fn()
    torch.cuda.synchronize(dev)
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(steps):
        fn()
    t1.record()
    torch.cuda.synchronize(dev)
    return t0.elapsed_time(t1) / steps  # ms per call


def stage_benchmarks(dev, steps: int, warmup: int, peak_gbs: float) -> dict:
    """BASELINE.json configs C2, C3, C4 as single-stage lines, each against the HBM roofline
    (algorithmic bytes per SURVEY §8d).  Buffers rotate so no input is re-read from L2."""
    import torch

    from workloads import synth  # seeded input generators (host side); not on the timed path
    from shared.processing import device as D

    out = {}
    # ---- C2: 64 x 640x480 -> K1
    sets = [torch.randint(0, 256, (64, 480, 640, 3), dtype=torch.uint8, device=dev) for _ in range(4)]
    outs = [torch.empty((64, 3, 640, 640), dtype=torch.float32, device=dev) for _ in range(2)]
    it = {"i": 0}

    def c2():
        i = it["i"] = it["i"] + 1
        D.preprocess_batch_cuda(sets[i % 4], 640, out=outs[i % 2])

    ms = time_loop(c2, steps, warmup, dev)
    by = 64 * (480 * 640 * 3 + 3 * 640 * 640 * 4)
    out["C2_preprocess_64x640x480"] = {"images_per_s": 64 / ms * 1e3, "ms": ms, "algorithmic_bytes": by,
                                       "achieved_gbs": by / ms / 1e6, "frac_of_hbm_peak": by / ms / 1e6 / peak_gbs}
    del sets, outs
    # ---- K1's general linear mode on the shapes COCO actually has: 500x375 (upscale to 640x480)
    # and 1024x768 (downscale), 64 frames each; algorithmic bytes per SURVEY 8d
    for name, h, w in (("K1_linear_up_64x500x375", 375, 500), ("K1_linear_down_64x1024x768", 768, 1024)):
        sets = [torch.randint(0, 256, (64, h, w, 3), dtype=torch.uint8, device=dev) for _ in range(4)]
        outs = [torch.empty((64, 3, 640, 640), dtype=torch.float32, device=dev) for _ in range(2)]
        it["i"] = 0

        def k1_linear():
            i = it["i"] = it["i"] + 1
            D.preprocess_batch_cuda(sets[i % 4], 640, out=outs[i % 2])

        ms = time_loop(k1_linear, steps, warmup, dev)
        by = 64 * (h * w * 3 + 3 * 640 * 640 * 4)
        out[name] = {"images_per_s": 64 / ms * 1e3, "ms": ms, "algorithmic_bytes": by,
                     "achieved_gbs": by / ms / 1e6, "frac_of_hbm_peak": by / ms / 1e6 / peak_gbs}
        del sets, outs
    # ---- C3: 256 heads -> K2+K3
    _, head = synth_inputs_device(256, dev, seed=2)
    ws = torch.empty(D.lib().ia_decode_nms_workspace_bytes(256, ANCHORS, 4096), dtype=torch.uint8, device=dev)
    ms = time_loop(lambda: D.decode_nms_cuda(head, CONF, IOU, max_candidates=4096, max_det=32, workspace=ws),
                   steps, warmup, dev)
    by = 256 * ANCHORS * (5 + CLASSES) * 4
    out["C3_decode_nms_256x25200x85"] = {"images_per_s": 256 / ms * 1e3, "ms": ms, "algorithmic_bytes": by,
                                         "achieved_gbs": by / ms / 1e6, "frac_of_hbm_peak": by / ms / 1e6 / peak_gbs}
    del head
    # ---- N1: the v8 channel-major head the shipped services feed, 256 x 84 x 8400
    h8 = torch.rand((256, 84, 8400), device=dev) * 0.3
    h8[:, 0:2] = torch.rand((256, 2, 8400), device=dev) * 640
    h8[:, 2:4] = 8 + torch.rand((256, 2, 8400), device=dev) * 192
    sel = torch.rand((256, 8400), device=dev) < 0.012  # ~100 candidates per image at conf 0.5
    cls = torch.randint(4, 84, (256, 8400), device=dev)
    bi, ai = sel.nonzero(as_tuple=True)
    h8[bi, cls[bi, ai], ai] = 0.55 + 0.4 * torch.rand(bi.numel(), device=dev)
    ws8 = torch.empty(D.lib().ia_decode_nms_workspace_bytes(256, 8400, 4096), dtype=torch.uint8, device=dev)
    ms = time_loop(lambda: D.decode_nms_cuda(h8, 0.5, IOU, layout="v8", max_candidates=4096, max_det=128,
                                             workspace=ws8), steps, warmup, dev)
    by = 256 * 84 * 8400 * 4
    out["N1_v8_decode_nms_256x84x8400"] = {"images_per_s": 256 / ms * 1e3, "ms": ms, "algorithmic_bytes": by,
                                           "achieved_gbs": by / ms / 1e6, "frac_of_hbm_peak": by / ms / 1e6 / peak_gbs}
    del h8
    # ---- C1-sized request (one 640x480 frame + one head) through the whole chain: launch-bound,
    # so the six launches are also replayed from a CUDA graph
    from shared.processing.pipeline import DevicePipeline, GraphedPipeline
    f1 = torch.randint(0, 256, (1, 480, 640, 3), dtype=torch.uint8, device=dev)
    _, h1 = synth_inputs_device(1, dev, seed=1)
    p1 = DevicePipeline(1, 480, 640, dev)
    eager_ms = time_loop(lambda: p1.run(f1, h1), max(steps, 50), warmup, dev)
    g1 = GraphedPipeline(p1, 1)
    g1.frames.copy_(f1)
    g1.head.copy_(h1)
    graph_ms = time_loop(g1.replay, max(steps, 50), warmup, dev)
    out["C1_single_request_640x480"] = {"eager_us": eager_ms * 1e3, "cuda_graph_us": graph_ms * 1e3,
                                        "requests_per_s_graph": 1e3 / graph_ms}
    del g1, p1
    # ---- C4: 512 frames 640x480, ~4 ROIs each -> K4
    frames = torch.randint(0, 256, (512, 480, 640, 3), dtype=torch.uint8, device=dev)
    dets_np, count_np = synth.detections(512, seed=33, kmax=8)
    dets, count = torch.from_numpy(dets_np).to(dev), torch.from_numpy(count_np).to(dev)
    geom = D.letterbox_geometry(480, 640, 640)
    crops = torch.empty((512 * 8, 3, 224, 224), dtype=torch.float32, device=dev)
    _, offsets, rois = D.crop_batch_cuda(frames, dets, count, geom, out=crops)
    r = rois[: int(offsets[-1])].cpu().numpy()
    area = ((r[:, 3] - r[:, 1]) * (r[:, 4] - r[:, 2]) * (1 - r[:, 5]) + r[:, 5]).sum()
    by = int(area) * 3 + len(r) * 3 * 224 * 224 * 4
    ms = time_loop(lambda: D.crop_batch_cuda(frames, dets, count, geom, out=crops), steps, warmup, dev)
    out["C4_crop_512img"] = {"images_per_s": 512 / ms * 1e3, "rois": len(r), "ms": ms, "algorithmic_bytes": by,
                             "achieved_gbs": by / ms / 1e6, "frac_of_hbm_peak": by / ms / 1e6 / peak_gbs}
    del frames, crops
    # ---- N4: softmax + top-5 of MobileNetV2 logits — one C5 step's worth of crops
    # (launch-bound) and a batch larger than L2 (HBM-bound); algorithmic bytes = the logits read
    for rows in (1035, 65536):
        lg = torch.randn((rows, 1000), device=dev) * 4
        ms = time_loop(lambda: D.classify_topk_cuda(lg, 5), steps, warmup, dev)
        by = rows * 1000 * 4
        out[f"N4_softmax_top5_{rows}x1000"] = {"rows_per_s": rows / ms * 1e3, "ms": ms, "algorithmic_bytes": by,
                                               "achieved_gbs": by / ms / 1e6,
                                               "frac_of_hbm_peak": by / ms / 1e6 / peak_gbs}
        del lg
    # ---- N3: baseline JPEG decode on the device, 64 x 1080p 4:2:0 quality 90.  Not an
    # HBM-bound stage (Huffman decoding is latency-bound); algorithmic bytes = file in + RGB out
    try:
        import cv2  # encoder for the synthetic files, and the host decoder timed beside the GPU
        import numpy as np

        from shared.processing import jpeg as J
        from workloads import jpeg_cases as W
        files = [W.encode(W.photo(1080, 1920, seed=i), 90, "420") for i in range(8)]
        blobs = [files[i % 8] for i in range(64)]
        batch = J.decode_jpeg_batch_cuda(blobs, dev)
        torch.cuda.synchronize(dev)
        ok = int(batch.status.sum().item()) == 0
        ms = time_loop(batch.rerun, max(steps // 4, 10), 3, dev)
        calls = []  # host wall clock per call (parse, upload, six launches): the median of seven
        for _ in range(7):
            t0 = time.perf_counter()
            last = J.decode_jpeg_batch_cuda(blobs, dev)
            torch.cuda.synchronize(dev)
            calls.append((time.perf_counter() - t0) * 1e3)
        call_ms = sorted(calls)[len(calls) // 2]
        t0 = time.perf_counter()
        for b in files:
            cv2.cvtColor(cv2.imdecode(np.frombuffer(b, np.uint8), cv2.IMREAD_COLOR), cv2.COLOR_BGR2RGB)
        cpu_ms = (time.perf_counter() - t0) / len(files) * 1e3
        # the same call on every host core (cv2.imdecode releases the GIL: threads, no fork after CUDA init)
        from concurrent.futures import ThreadPoolExecutor
        cores = len(os.sched_getaffinity(0))

        def host_decode(b):
            return cv2.cvtColor(cv2.imdecode(np.frombuffer(b, np.uint8), cv2.IMREAD_COLOR), cv2.COLOR_BGR2RGB).shape

        with ThreadPoolExecutor(cores) as pool:
            list(pool.map(host_decode, blobs[:cores]))
            t0 = time.perf_counter()
            list(pool.map(host_decode, blobs * 2))
            cpu_all = len(blobs) * 2 / (time.perf_counter() - t0)
        by = sum(len(b) for b in blobs) + 64 * 1080 * 1920 * 3
        out["N3_jpeg_decode_64x1080p"] = {
            "images_per_s": 64 / ms * 1e3, "ms": ms, "status_ok": ok, "file_kb": sum(map(len, files)) / 8 / 1024,
            "algorithmic_bytes": by, "achieved_gbs": by / ms / 1e6, "frac_of_hbm_peak": by / ms / 1e6 / peak_gbs,
            "from_host_bytes_images_per_s": 64 / call_ms * 1e3, "from_host_bytes_ms": call_ms,
            "cv2_imdecode_one_core_ms_per_image": cpu_ms,
            "cv2_imdecode_all_cores_images_per_s": cpu_all, "host_cores": cores}
        del batch, last
        # the same frames at quality 75 (files of the size photographs usually have)
        files75 = [W.encode(W.photo(1080, 1920, seed=i), 75, "420") for i in range(8)]
        batch = J.decode_jpeg_batch_cuda([files75[i % 8] for i in range(64)], dev)
        torch.cuda.synchronize(dev)
        ms75 = time_loop(batch.rerun, max(steps // 4, 10), 3, dev)
        out["N3_jpeg_decode_64x1080p"]["quality75"] = {
            "images_per_s": 64 / ms75 * 1e3, "ms": ms75, "file_kb": sum(map(len, files75)) / 8 / 1024,
            "status_ok": int(batch.status.sum().item()) == 0}
        del batch
    except ImportError as e:  # no encoder for the synthetic files on this box
        out["N3_jpeg_decode_64x1080p"] = {"skipped": str(e)}
    return out


def run_gpu_arm(args) -> None:
    import numpy as np
    import torch
    import torch.distributed as dist

    from shared.processing import device as D
    from shared.processing.pipeline import KERNELS_PER_STEP, DevicePipeline, HostPipeline

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    cpu = cpu_baseline_leg() if (world == 1 and not args.no_cpu) else None  # before CUDA init (fork)
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device (the product path has no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa = None
    if not args.no_numa_bind:  # pinned staging memory should live on the GPU's own NUMA node
        from shared.processing.numa import bind_to_gpu_numa
        numa = bind_to_gpu_numa(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    D.lib()
    B = args.batch
    peak_gbs, peak_src = hbm_peak()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    frames, head = synth_inputs_device(B, dev, seed=4 + rank)
    pipe = DevicePipeline(B, SRC_H, SRC_W, dev, max_det=32, max_candidates=4096, rois_per_image=8)

    # ---------------- device-resident throughput (`value`) + per-kernel events -----------
    # the clock sampler runs from before the warm-up to the end of the timed region; the warm-up
    # (the same steps, untimed) lasts at least 0.4 s so that a 10 ms sampler sees the load
    sampler = ClockSampler(local) if rank == 0 else None
    t_warm = time.perf_counter()
    done = 0
    while done < args.warmup or time.perf_counter() - t_warm < 0.4:
        out = pipe.run(frames, head)
        done += 1
        if done % 16 == 0:
            torch.cuda.synchronize(dev)
    torch.cuda.synchronize(dev)
    assert int(out.status.max()) == 0, "synthetic workload overflowed the candidate/detection budget"
    total_rois = int(out.roi_offsets[-1])
    rois_np = out.rois[:total_rois].cpu().numpy()
    crop_px = int(((rois_np[:, 3] - rois_np[:, 1]) * (rois_np[:, 4] - rois_np[:, 2]) * (1 - rois_np[:, 5])
                   + rois_np[:, 5]).sum())
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(args.steps)]
    barrier()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for s in range(args.steps):
        pipe.run(frames, head, events=ev[s])
    t1.record()
    barrier()
    clocks = sampler.stop() if sampler else None
    elapsed_ms = max_over_ranks(t0.elapsed_time(t1))
    ms_per_step = elapsed_ms / args.steps
    value = world * B * args.steps / (elapsed_ms / 1e3)

    k_ms = {"K1_letterbox_normalize_chw": statistics.mean(e[0].elapsed_time(e[1]) for e in ev),
            "K2K3_decode_filter_nms": statistics.mean(e[1].elapsed_time(e[2]) for e in ev),
            "K4_roi_crop_resize_normalize": statistics.mean(e[2].elapsed_time(e[3]) for e in ev)}
    k_bytes = {"K1_letterbox_normalize_chw": B * (SRC_H * SRC_W * 3 + 3 * 640 * 640 * 4),
               "K2K3_decode_filter_nms": B * ANCHORS * (5 + CLASSES) * 4,
               "K4_roi_crop_resize_normalize": crop_px * 3 + total_rois * 3 * 224 * 224 * 4}
    kernels = {k: {"ms": k_ms[k], "algorithmic_bytes": k_bytes[k], "achieved_gbs": k_bytes[k] / k_ms[k] / 1e6,
                   "frac": k_bytes[k] / k_ms[k] / 1e6 / peak_gbs, "share_of_step": k_ms[k] / sum(k_ms.values())}
               for k in k_ms}
    # K1 at 1080p is a 3x gather: it reads only the source rows it needs (one in three), so the
    # algorithmic figure above exceeds the HBM peak; the bytes it physically moves are the honest
    # roofline fraction (the ncu DRAM counters in profiles/ agree with this count)
    g = D.letterbox_geometry(SRC_H, SRC_W, 640)
    if SRC_H == 3 * g.new_h and SRC_W == 3 * g.new_w:
        k1 = kernels["K1_letterbox_normalize_chw"]
        k1["physical_bytes"] = B * (g.new_h * SRC_W * 3 + 3 * 640 * 640 * 4)
        k1["frac_physical"] = k1["physical_bytes"] / k1["ms"] / 1e6 / peak_gbs
        k1["note"] = "frac = algorithmic bytes (SURVEY 8d) / time / peak; frac_physical counts only the source rows the 3x gather reads"
    dom = max(k_ms, key=k_ms.get)
    roofline = {"bound": "hbm", "kernel": dom, "achieved": kernels[dom]["achieved_gbs"], "peak": peak_gbs,
                "unit": "GB/s", "frac": kernels[dom]["frac"], "traffic": None, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": k_bytes[dom], "avg_launch_ms": k_ms[dom]}
    traffic_file = REPO / "profiles" / "traffic.json"
    if traffic_file.exists():
        try:
            roofline["traffic"] = json.loads(traffic_file.read_text()).get(dom)
        except Exception:
            pass

    # extra, not the headline: the same steps with K1 on a side stream (see DevicePipeline.run)
    for _ in range(3):
        pipe.run(frames, head, overlap=True)
    barrier()
    t0.record()
    for _ in range(args.steps):
        pipe.run(frames, head, overlap=True)
    t1.record()
    barrier()
    overlapped_value = world * B * args.steps / (max_over_ranks(t0.elapsed_time(t1)) / 1e3)

    # ---------------- end-to-end: host buffers in and out (`e2e`) -----------------------
    if args.no_e2e:  # profiling runs (ncu) only; the e2e key is then null
        if rank == 0:
            print(json.dumps({"metric": METRIC, "value": value, "unit": UNIT, "ms_per_step": ms_per_step,
                              "kernels": kernels, "roofline": roofline, "e2e": None, "note": "--no-e2e profiling run"}))
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return
    hp = HostPipeline(B, SRC_H, SRC_W, dev, chunk=min(32, B), slots=3, max_det=32, max_candidates=4096,
                      rois_per_image=8)
    h_frames = torch.empty(frames.shape, dtype=torch.uint8, pin_memory=True).copy_(frames)
    h_heads = torch.empty(head.shape, dtype=torch.float32, pin_memory=True).copy_(head)
    for _ in range(max(1, min(args.warmup, 3))):
        res = hp.run(h_frames, h_heads)
    e2e_steps = max(3, min(args.steps, 10))
    barrier()
    t0.record()
    ticket = None
    for _ in range(e2e_steps):  # asynchronous submits, two batches in flight; every copy is timed
        nxt = hp.submit(h_frames, h_heads)
        if ticket is not None:
            res = ticket.wait()
        ticket = nxt
    res = ticket.wait()
    t1.record()
    barrier()
    e2e_ms = max_over_ranks(t0.elapsed_time(t1))
    e2e_value = world * B * e2e_steps / (e2e_ms / 1e3)
    assert int(res["status"].max()) == 0
    h2d = B * hp.h2d_bytes_per_image
    d2h = hp.d2h_bytes(B)

    # ---------------- the same leg with the frames arriving as JPEG files ----
    # What the services actually receive.  Extra information, not the `e2e` key: the boundary
    # moves one step earlier (the reference decodes with cv2 on the host before the path).
    # Every rank runs these legs (they move 0.8 MB per image instead of 14.8 MB: the legs that
    # can scale on a host whose copy bandwidth is shared); times are the max over ranks.
    def timed_leg(pipe_, src_, heads_, **kw):
        for _ in range(2):
            r = pipe_.run(src_, heads_, **kw)
        barrier()
        t0.record()
        tk = None
        for _ in range(e2e_steps):
            nx = pipe_.submit(src_, heads_, **kw)
            if tk is not None:
                r = tk.wait()
            tk = nx
        r = tk.wait()
        t1.record()
        barrier()
        assert int(r["status"].max()) == 0
        return max_over_ranks(t0.elapsed_time(t1))

    e2e_jpeg = None
    if not args.no_stages:
        try:
            from workloads import jpeg_cases as W
            files = [W.encode(W.photo(SRC_H, SRC_W, seed=i), 90, "420") for i in range(8)]
            blobs = [files[i % 8] for i in range(B)]
            jm = timed_leg(hp, blobs, h_heads)
            e2e_jpeg = {"value": world * B * e2e_steps / (jm / 1e3), "unit": UNIT, "h2d_bytes_per_step": hp.last_h2d_bytes,
                        "d2h_bytes_per_step": hp.d2h_bytes(B), "ms_per_step": jm / e2e_steps,
                        "jpeg_kb_per_frame": sum(map(len, files)) / 8 / 1024,
                        "api": "HostPipeline.submit(list_of_jpeg_bytes, heads): files decoded on the device"}
            # and with both models on the GPU: the head never leaves the device, the
            # model inputs are not copied back; only frames go up and detections come down
            on_gpu = {}
            # third leg: 64-file chunks on two streams — decode-bound work likes larger launches
            # (the Huffman kernel switches to its 128 KB chunks), PCIe-bound work smaller ones
            legs = [("raw_frames", h_frames, hp), ("jpeg_files", blobs, hp)]
            if B >= 128:
                legs.append(("jpeg_files_chunk64", blobs, None))
            for name, src, pipe_ in legs:
                if pipe_ is None:
                    pipe_ = HostPipeline(B, SRC_H, SRC_W, dev, chunk=64, slots=2, max_det=32, max_candidates=4096,
                                         rois_per_image=8)
                gm = timed_leg(pipe_, src, head, tensors_to_host=False)
                on_gpu[name] = {"value": world * B * e2e_steps / (gm / 1e3), "unit": UNIT, "ms_per_step": gm / e2e_steps,
                                "h2d_bytes_per_step": pipe_.last_h2d_bytes, "d2h_bytes_per_step": pipe_.d2h_bytes(B)}
            e2e_jpeg["models_on_gpu"] = on_gpu
        except ImportError as e:
            e2e_jpeg = {"skipped": str(e)}

    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    # ---------------- rank 0 only: stage lines, CPU baseline, parity spot check ----------
    del hp, h_frames, h_heads, frames, head, pipe
    torch.cuda.empty_cache()
    stages = stage_benchmarks(dev, max(args.steps, 10), max(args.warmup, 3), peak_gbs) if not args.no_stages else {}
    parity = parity_spot_check(dev) if cpu is not None else None

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8/f32", "data": "synthetic", "config": workload_config(B, world),
        "roofline": roofline, "kernels": kernels, "stages": stages,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "steps": e2e_steps, "ms_per_step": e2e_ms / e2e_steps, **copy_ceiling(world, e2e_value, h2d, d2h, B),
                "api": "shared.processing.pipeline.HostPipeline.submit/wait (pinned host buffers in and out, 3 streams x 32-image chunks, 2 batches in flight)"},
        "e2e_jpeg": e2e_jpeg,
        "gpu_launches": KERNELS_PER_STEP * args.steps,
        "clocks": clocks,
        "rois_per_step": total_rois,
        "value_two_streams": overlapped_value,
        "numa_bind": numa,
    }
    if cpu is not None:
        line["cpu_baseline"] = cpu
        line["parity_spot_check"] = parity
    print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def copy_ceiling(world: int, e2e_value: float, h2d_bytes: int, d2h_bytes: int, batch: int) -> dict:
    """How close the host-buffer leg is to what the box's pinned-memory copies alone sustain on
    `world` GPUs at once (profiles/copy_ceiling.json, measured by scripts/copy_ceiling.py with no
    kernels): each direction's best rate divided by the bytes this leg moves per image in that
    direction; the slower direction is the ceiling."""
    path = REPO / "profiles" / "copy_ceiling.json"
    try:
        row = json.loads(path.read_text())["by_gpus"][str(world)]
        up = row["h2d_alone_gbs"] * 1e9 / (h2d_bytes / batch)
        down = row["d2h_alone_gbs"] * 1e9 / (d2h_bytes / batch)
        mix = row["e2e_mix_images_per_s"]
    except Exception:
        return {}
    # two yardsticks: the same copies with no kernels in between (what this software structure can
    # reach on this host), and each direction's best rate measured alone (an upper bound that the
    # host does not sustain when both directions run)
    return {"copy_ceiling_images_per_s": mix, "frac_of_copy_ceiling": e2e_value / mix,
            "copy_upper_bound_images_per_s": min(up, down), "frac_of_copy_upper_bound": e2e_value / min(up, down)}


def cpu_baseline_leg() -> dict:
    """cpu_baseline leg: the oracle port timed on the host cores over a bounded sample
    (~20 core-seconds).  Runs BEFORE CUDA is initialised (the pool forks)."""
    arm = CpuArm(images_per_worker=4)
    arm.step(1)  # warm-up
    _, w1 = arm.step(1)
    reps = max(1, min(50, int(20.0 / max(w1 * arm.cores, 1e-3))))
    n, wall = arm.step(reps)
    desc = arm.describe(n, wall)
    arm.close()
    return desc


def parity_spot_check(dev) -> str:
    """The oracle in its role as checker (part of the cpu_baseline leg): four requests of
    the bench workload through the GPU path, compared bit for bit."""
    import numpy as np
    import torch

    from oracle import pipeline as O
    from workloads import synth
    from shared.processing.pipeline import DevicePipeline

    frames = synth.frames(4, SRC_H, SRC_W, seed=77)
    heads = synth.head_v5(4, seed=78)
    pipe = DevicePipeline(4, SRC_H, SRC_W, dev, max_det=32, max_candidates=4096, rois_per_image=8)
    out = pipe.run(torch.from_numpy(frames).to(dev), torch.from_numpy(heads).to(dev))
    torch.cuda.synchronize(dev)
    offs = out.roi_offsets.cpu().numpy()
    ok = True
    for i in range(4):
        t, dets, keep, _, crops = O.run_image(frames[i], heads[i], CONF, IOU)
        k = int(out.count[i])
        ok &= k == len(keep) and out.keep[i, :k].cpu().tolist() == keep
        ok &= np.array_equal(out.tensor[i:i + 1].cpu().numpy(), t)
        ok &= np.array_equal(out.dets[i, :k].cpu().numpy(), dets)
        ok &= np.array_equal(out.crops[offs[i]:offs[i + 1]].cpu().numpy(), crops)
    return "bit-exact vs oracle on 4 requests" if ok else "MISMATCH vs oracle"


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--batch", type=int, default=256, help="images per GPU per step")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-stages", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-numa-bind", action="store_true", help="do not pin the process to the GPU's NUMA node")
    ap.add_argument("--no-e2e", action="store_true", help="profiling runs only: skip the host-buffer leg")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup

    if args.impl == "reference":
        run_reference_arm(args)
        return
    if args.gpus > 1 and "RANK" not in os.environ:
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", "29531", __file__] + sys.argv[1:]
        raise SystemExit(subprocess.call(cmd))
    run_gpu_arm(args)


if __name__ == "__main__":
    main()
