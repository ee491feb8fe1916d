"""PCIe ceiling of the e2e arm: bare pinned H2D copies of one bench step (384 C1 frames = 138.6 MB), alone and
with the step's 24.5 MB D2H beside it.  Measured on the B200 box: 2.85 ms (48.6 GB/s) / 2.65 ms (52.2 GB/s);
bench.py's e2e step takes 2.65 ms."""
import torch, time
n = 138611712
h = torch.empty(n, dtype=torch.uint8).pin_memory()
d = torch.empty(n, dtype=torch.uint8, device="cuda")
h2 = torch.empty(24520704, dtype=torch.uint8).pin_memory()
d2 = torch.empty(24520704, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
for both in (False, True):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(20):
        with torch.cuda.stream(s1):
            d.copy_(h, non_blocking=True)
        if both:
            with torch.cuda.stream(s2):
                h2.copy_(d2, non_blocking=True)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / 20
    print("with D2H" if both else "H2D only", round(dt * 1e3, 3), "ms per 138.6 MB ->", round(n / dt / 1e9, 1), "GB/s")
