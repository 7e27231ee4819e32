"""Ad-hoc: models-on-GPU leg from JPEG files for different HostPipeline chunk sizes."""
import sys, torch
from pathlib import Path
R = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(R), str(R / "inference-arena_b200")]
from shared.processing.pipeline import HostPipeline
from workloads import jpeg_cases as W, synth
B, H, Wd = 256, 1080, 1920
files = [W.encode(W.photo(H, Wd, seed=i), 90, "420") for i in range(8)]
blobs = [files[i % 8] for i in range(B)]
head = torch.from_numpy(synth.head_v5(8, seed=3)).cuda().repeat(B // 8, 1, 1)
for chunk, slots in ((32, 3), (64, 3), (64, 2), (128, 2)):
    hp = HostPipeline(B, H, Wd, "cuda", chunk=chunk, slots=slots, max_det=32, max_candidates=4096, rois_per_image=8)
    for _ in range(2):
        hp.run(blobs, head, tensors_to_host=False)
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record(); ticket = None
    for _ in range(8):
        nxt = hp.submit(blobs, head, tensors_to_host=False)
        if ticket is not None: ticket.wait()
        ticket = nxt
    ticket.wait(); t1.record(); torch.cuda.synchronize()
    ms = t0.elapsed_time(t1) / 8
    print(f"chunk {chunk} slots {slots}: {ms:.2f} ms/step -> {B / ms * 1e3:.0f} img/s")
    del hp
