"""Host<->device copy bandwidth of the box (pinned memory), alone and both directions at once: the ceiling of bench.py's
e2e number (3.02 GB in + 1.51 GB out per 120-frame step)."""
import time
import torch

n = 1 << 30
h_in = torch.empty(n, dtype=torch.uint8, pin_memory=True)
h_out = torch.empty(n, dtype=torch.uint8, pin_memory=True)
d_in = torch.empty(n, dtype=torch.uint8, device="cuda")
d_out = torch.empty(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def timeit(fn, reps=5):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps


def h2d():
    with torch.cuda.stream(s1):
        d_in.copy_(h_in, non_blocking=True)


def d2h():
    with torch.cuda.stream(s2):
        h_out.copy_(d_out, non_blocking=True)


def both():
    h2d(); d2h()


print("H2D GB/s", n / timeit(h2d) / 1e9)
print("D2H GB/s", n / timeit(d2h) / 1e9)
t = timeit(both)
print("both: H2D+D2H GB/s each", n / t / 1e9)
import subprocess
print(subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True).stdout[:1500])
import os
print("cpus", os.cpu_count(), "affinity", len(os.sched_getaffinity(0)))
