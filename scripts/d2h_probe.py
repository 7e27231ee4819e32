"""Ad-hoc: device->host copy bandwidth alone and with a concurrent host->device copy."""
import time, torch
dev = torch.device("cuda")
n = 1 << 30
d_src = torch.empty(n, dtype=torch.uint8, device=dev)
d_dst = torch.empty(n, dtype=torch.uint8, device=dev)
h_out = torch.empty(n, dtype=torch.uint8, pin_memory=True)
h_in = torch.empty(n, dtype=torch.uint8, pin_memory=True); h_in.fill_(1)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def run(both, reps=5, parts=1):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    step = n // parts
    for _ in range(reps):
        for k in range(parts):
            with torch.cuda.stream(s1): h_out[k*step:(k+1)*step].copy_(d_src[k*step:(k+1)*step], non_blocking=True)
            if both:
                with torch.cuda.stream(s2): d_dst[k*step:(k+1)*step].copy_(h_in[k*step:(k+1)*step], non_blocking=True)
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    return reps * n / dt / 1e9
run(False, 2)
print(f"D2H alone: {run(False):.1f} GB/s")
print(f"D2H with concurrent H2D: {run(True):.1f} GB/s each direction")
print(f"D2H alone in 8 MB pieces: {run(False, parts=128):.1f} GB/s")
print(f"both in 8 MB pieces: {run(True, parts=128):.1f} GB/s each direction")
