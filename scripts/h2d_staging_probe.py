"""Ad-hoc: how to get a pageable numpy array onto the device fastest — torch's plain .to(device)
(the driver's own bounce buffer) against a copy into a pinned block by 1 / 2 / 4 / 8 host threads
followed by one asynchronous DMA."""
import sys, time
from concurrent.futures import ThreadPoolExecutor
import numpy as np, torch
dev = torch.device("cuda")
pool = ThreadPoolExecutor(8)
def plain(a): return torch.from_numpy(a).to(dev)
def staged(a, threads):
    flat = a.reshape(-1)
    pin = torch.empty(flat.shape, dtype=torch.from_numpy(flat[:1]).dtype, pin_memory=True)
    dst = pin.numpy()
    if threads == 1:
        np.copyto(dst, flat)
    else:
        step = -(-flat.size // threads)
        list(pool.map(lambda k: np.copyto(dst[k * step:(k + 1) * step], flat[k * step:(k + 1) * step]), range(threads)))
    out = pin.to(dev, non_blocking=True)
    torch.cuda.current_stream().synchronize()
    return out.view(a.shape)
def t(fn, n=30):
    for _ in range(5): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e3
for name, a in [("head 25200x85 f32", np.random.rand(1, 25200, 85).astype(np.float32)),
                ("frame 1080p u8", np.random.randint(0, 256, (1080, 1920, 3), dtype=np.uint8)),
                ("frame 640x480 u8", np.random.randint(0, 256, (480, 640, 3), dtype=np.uint8))]:
    line = f"{name} ({a.nbytes/1e6:.2f} MB): .to(device) {t(lambda: plain(a)):.3f} ms"
    for th in (1, 2, 4, 8):
        line += f", staged x{th} {t(lambda: staged(a, th)):.3f}"
    assert torch.equal(staged(a, 4).cpu(), torch.from_numpy(a))
    print(line)
