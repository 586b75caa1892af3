"""Host<->device copy bandwidth probe (pinned torch tensor vs the Arrow-buffer view the
staging path uses)."""
import time
import warnings

import numpy as np
import pyarrow as pa
import torch

n = 2 << 30
dev = torch.device("cuda")
d = torch.empty(n, dtype=torch.uint8, device=dev)
h = torch.empty(n, dtype=torch.uint8, pin_memory=True)
h.fill_(1)


def t(fn, reps=3):
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        torch.cuda.synchronize()
        best = min(best, time.perf_counter() - t0)
    return n / best / 1e9


print("H2D pinned torch tensor        GB/s", round(t(lambda: d.copy_(h, non_blocking=True)), 1))
print("D2H pinned torch tensor        GB/s", round(t(lambda: h.copy_(d, non_blocking=True)), 1))
buf = pa.py_buffer(h.numpy())
with warnings.catch_warnings():
    warnings.simplefilter("ignore")
    v = torch.from_numpy(np.frombuffer(buf, dtype=np.uint8))
print("arrow view is_pinned:", v.is_pinned())
print("H2D arrow-buffer view          GB/s", round(t(lambda: d.copy_(v, non_blocking=True)), 1))
# chunked (12 blocks) as staging does
k = 12
per = n // k
print("H2D 12 chunks                  GB/s",
      round(t(lambda: [d[i * per:(i + 1) * per].copy_(v[i * per:(i + 1) * per], non_blocking=True) for i in range(k)]), 1))
h2 = torch.empty(n, dtype=torch.uint8, pin_memory=True)
s2 = torch.cuda.Stream()


def both():
    d2 = d  # same buffer: bandwidth only
    with torch.cuda.stream(s2):
        h2.copy_(d2, non_blocking=True)
    d.copy_(h, non_blocking=True)


torch.cuda.synchronize()
t0 = time.perf_counter()
both()
torch.cuda.synchronize()
print("H2D + D2H concurrent, each     GB/s", round(n / (time.perf_counter() - t0) / 1e9, 1))
t0 = time.perf_counter()
x = torch.empty(4 << 30, dtype=torch.uint8, pin_memory=True)
print("pinned alloc 4 GiB s", round(time.perf_counter() - t0, 2))
