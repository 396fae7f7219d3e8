"""Host<->device copy rates of the box (pinned memory, CUDA events): what bounds the e2e line of bench.py."""
import torch
dev = torch.device("cuda", 0)
def rate(nbytes, d2h=False, both=False, reps=10):
    h = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True); d = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    h2 = torch.empty(nbytes // 2, dtype=torch.uint8, pin_memory=True); d2 = torch.empty(nbytes // 2, dtype=torch.uint8, device=dev)
    s2 = torch.cuda.Stream(dev)
    for _ in range(2):
        d.copy_(h, non_blocking=True)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        if d2h: h.copy_(d, non_blocking=True)
        else: d.copy_(h, non_blocking=True)
        if both:
            with torch.cuda.stream(s2):
                h2.copy_(d2, non_blocking=True)
    b.record(); torch.cuda.synchronize(); s2.synchronize()
    return nbytes * reps / (a.elapsed_time(b) * 1e-3) / 1e9
for mb in (25, 100, 400, 1600):
    n = mb << 20
    print(f"{mb:5d} MB  H2D {rate(n):6.1f} GB/s   D2H {rate(n, d2h=True):6.1f} GB/s   H2D with concurrent D2H of half the size {rate(n, both=True):6.1f} GB/s")
