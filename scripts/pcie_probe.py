"""Ad-hoc: host<->device copy bandwidth for pinned vs write-combined pinned memory."""
import ctypes, time, torch, numpy as np
cudart = ctypes.CDLL("libcudart.so.12") if False else None
try:
    cudart = ctypes.CDLL([l for l in open("/proc/self/maps").read().split() if "libcudart" in l][0])
except Exception:
    import glob
    cands = glob.glob("/opt/prime-rl/.venv/lib/python3.12/site-packages/nvidia/cuda_runtime/lib/libcudart.so*") + glob.glob("/usr/local/cuda/lib64/libcudart.so*")
    cudart = ctypes.CDLL(cands[0])
dev = torch.device("cuda")
n = 1 << 30
d = torch.empty(n, dtype=torch.uint8, device=dev)
def bw(h, label, both=False):
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    d2 = torch.empty(n, dtype=torch.uint8, device=dev) if both else None
    h2 = torch.empty(n, dtype=torch.uint8, pin_memory=True) if both else None
    for _ in range(2): d.copy_(h, non_blocking=True)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(5):
        with torch.cuda.stream(s1): d.copy_(h, non_blocking=True)
        if both:
            with torch.cuda.stream(s2): h2.copy_(d2, non_blocking=True)
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    print(f"{label}: H2D {5*n/dt/1e9:.1f} GB/s" + (" (with concurrent D2H of the same size)" if both else ""))
h = torch.empty(n, dtype=torch.uint8, pin_memory=True); h.fill_(1)
bw(h, "pinned"); bw(h, "pinned", both=True)
ptr = ctypes.c_void_p()
rc = cudart.cudaHostAlloc(ctypes.byref(ptr), ctypes.c_size_t(n), ctypes.c_uint(0x04))  # cudaHostAllocWriteCombined
print("cudaHostAlloc WC rc", rc)
if rc == 0:
    buf = (ctypes.c_uint8 * n).from_address(ptr.value)
    hw = torch.frombuffer(buf, dtype=torch.uint8)
    hw.fill_(1)
    print("is_pinned:", hw.is_pinned())
    bw(hw, "write-combined"); bw(hw, "write-combined", both=True)
