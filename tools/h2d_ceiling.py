#!/usr/bin/env python
"""What the box can move from host memory to its GPUs at once: N concurrent pinned cudaMemcpyAsync streams of the
size one config-2 step uploads per GPU (631 MB), one stream per GPU, for N = 1 .. all visible GPUs.  The end-to-end
numbers of bench.py at N GPUs are read against this ceiling.  One JSON line."""
import json
import sys
import time

import torch

nbytes = int(sys.argv[1]) if len(sys.argv) > 1 else 631_000_000
reps = 5
ndev = torch.cuda.device_count()
host = [torch.empty(nbytes, dtype=torch.uint8).pin_memory() for _ in range(ndev)]
dev = [torch.empty(nbytes, dtype=torch.uint8, device=f"cuda:{d}") for d in range(ndev)]
streams = [torch.cuda.Stream(device=d) for d in range(ndev)]
rows = []
for n in [k for k in (1, 2, 4, 8, 16) if k <= ndev]:
    def once():
        for d in range(n):
            with torch.cuda.stream(streams[d]):
                dev[d].copy_(host[d], non_blocking=True)
        for d in range(n):
            streams[d].synchronize()
    once()
    t0 = time.perf_counter()
    for _ in range(reps):
        once()
    dt = (time.perf_counter() - t0) / reps
    rows.append({"gpus": n, "ms_per_round": dt * 1e3, "aggregate_GBps": n * nbytes / dt / 1e9, "per_gpu_GBps": nbytes / dt / 1e9})
# and the other direction at full width (results come home over the same links)
def back():
    for d in range(ndev):
        with torch.cuda.stream(streams[d]):
            host[d].copy_(dev[d], non_blocking=True)
    for d in range(ndev):
        streams[d].synchronize()
back()
t0 = time.perf_counter()
for _ in range(reps):
    back()
dt = (time.perf_counter() - t0) / reps
print(json.dumps({"what": "concurrent pinned H2D copies, one stream per GPU", "bytes_per_gpu": nbytes, "h2d": rows,
                  "d2h_all_gpus": {"gpus": ndev, "aggregate_GBps": ndev * nbytes / dt / 1e9}}))
