"""Ad-hoc (torchrun): the host-buffer leg of bench.py alone, for several (chunk, slots) settings of
HostPipeline, max over ranks.  torchrun --nproc-per-node N scripts/e2e_sweep.py [CHUNKxSLOTS ...]"""
import os
import sys
from pathlib import Path
R = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(R)); sys.path.insert(0, str(R / "inference-arena_b200"))
import torch
import torch.distributed as dist
from shared.processing.pipeline import HostPipeline

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
B, steps = 256, 4
frames = torch.randint(0, 256, (B, 1080, 1920, 3), dtype=torch.uint8).pin_memory()
from workloads import synth
heads = torch.from_numpy(synth.head_v5(16, seed=5 + rank)).repeat(B // 16, 1, 1).pin_memory()  # ~4 detections per image, as in bench.py
settings = [tuple(int(v) for v in a.split("x")) for a in sys.argv[1:]] or [(32, 3), (64, 2), (16, 4), (64, 3), (128, 2)]
for chunk, slots in settings:
    hp = HostPipeline(B, 1080, 1920, dev, chunk=chunk, slots=slots, max_det=32, max_candidates=4096, rois_per_image=8)
    hp.run(frames, heads)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    tk = None
    for _ in range(steps):
        nx = hp.submit(frames, heads)
        if tk is not None:
            tk.wait()
        tk = nx
    tk.wait()
    t1.record()
    torch.cuda.synchronize(dev)
    ms = torch.tensor([t0.elapsed_time(t1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    if rank == 0:
        print(f"chunk {chunk} slots {slots}: {world * B * steps / (ms.item() / 1e3):.0f} images/s", flush=True)
    hp.drain()
    del hp
    torch.cuda.empty_cache()
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
