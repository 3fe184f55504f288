#!/usr/bin/env python
"""Concurrent H2D + D2H bandwidth of every visible GPU at once (pinned host buffers, one stream per direction and device):
what the box's PCIe / host memory allows the partitioned e2e leg.  Run under gpurun --gpus N."""
import time, torch
n = torch.cuda.device_count()
GB = 1 << 30
size = 2 * GB
hs = [(torch.empty(size, dtype=torch.uint8, pin_memory=True), torch.empty(size, dtype=torch.uint8, pin_memory=True)) for _ in range(n)]
ds = [(torch.empty(size, dtype=torch.uint8, device=f"cuda:{i}"), torch.empty(size, dtype=torch.uint8, device=f"cuda:{i}")) for i in range(n)]
ss = [(torch.cuda.Stream(device=i), torch.cuda.Stream(device=i)) for i in range(n)]
def run(devs, reps=3, both=True):
    for i in devs: torch.cuda.synchronize(i)
    t0 = time.perf_counter()
    for _ in range(reps):
        for i in devs:
            with torch.cuda.stream(ss[i][0]): ds[i][0].copy_(hs[i][0], non_blocking=True)
            if both:
                with torch.cuda.stream(ss[i][1]): hs[i][1].copy_(ds[i][1], non_blocking=True)
    for i in devs: torch.cuda.synchronize(i)
    dt = time.perf_counter() - t0
    return size * reps / dt / 1e9
run(range(n), 1)
print("gpus", n)
print("H2D alone, GPU 0           : %.1f GB/s" % run([0], both=False))
print("H2D + D2H, GPU 0           : %.1f GB/s per direction" % run([0]))
if n > 1:
    print("H2D + D2H, all %d GPUs      : %.1f GB/s per direction per GPU" % (n, run(range(n))))
    print("H2D alone, all %d GPUs      : %.1f GB/s per GPU" % (n, run(range(n), both=False)))
