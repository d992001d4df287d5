#!/usr/bin/env python
"""HBM calibration on the GPU box: what do plain torch kernels reach for read-only, write-only and copy traffic at
the bench's footprint (310 MB) and at 2 GB?  Prints GB/s (algorithmic bytes / CUDA-event time, best of 10)."""
import json
import torch

dev = torch.device("cuda", 0)


def best(fn, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e-3)
    return min(ts)


out = {}
for name, n in (("310MB", 77_414_400), ("2GB", 512 * 1024 * 1024)):
    x = torch.randn(n, device=dev)
    y = torch.empty_like(x)
    flush = torch.empty(64 * 1024 * 1024, device=dev)
    nb = n * 4
    out[name] = {
        "copy_rw": 2 * nb / best(lambda: y.copy_(x)) / 1e9,
        "read_sum": nb / best(lambda: x.sum()) / 1e9,
        "write_fill": nb / best(lambda: y.fill_(1.0)) / 1e9,
        "sigmoid_rw": 2 * nb / best(lambda: torch.sigmoid(x, out=y)) / 1e9,
        "add_2r1w": 3 * nb / best(lambda: torch.add(x, y, out=y)) / 1e9,
    }
    del x, y, flush
print(json.dumps(out))
