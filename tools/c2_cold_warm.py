"""C2 device time with and without the L2 flush between timed steps (how much of the device p50 is cold code / data)."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from quadruped_pympc_tamols_b200 import _cabi  # noqa: E402

dev = torch.device("cuda:0")
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
small = torch.empty(1 << 10, dtype=torch.uint8, device=dev)
# the floor of the measurement itself: the same event pair around a one-block torch kernel on a handle's stream
sb = bench.SolveBench(torch, _cabi, dev, 0, "c2", 0)
one = torch.zeros(32, device=dev)
with torch.cuda.stream(sb.stream):
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(300)]
    for k in range(330):
        flush.fill_(k & 0xFF)  # a long kernel ahead, as in the bench: the host is never the pacer
        if k >= 30:
            ev[k - 30][0].record(sb.stream)
        one.add_(1.0)
        if k >= 30:
            ev[k - 30][1].record(sb.stream)
    torch.cuda.synchronize()
fl = np.array([a.elapsed_time(b) for a, b in ev]) * 1e3
print(f"event pair around a one-block kernel: p50 {np.percentile(fl, 50):.2f} us  p10 {np.percentile(fl, 10):.2f}  p90 {np.percentile(fl, 90):.2f}")
sb.close()
for key in sys.argv[1:] or ["c2", "c1"]:
    sb = bench.SolveBench(torch, _cabi, dev, 0, key, 0)
    for name, buf in (("flushed", flush), ("warm", small), ("flushed", flush), ("warm", small)):
        ms, _ = sb.time_device(300, 30, buf)
        print(f"{key} {name}: p50 {np.percentile(ms, 50) * 1e3:.2f} us  p10 {np.percentile(ms, 10) * 1e3:.2f}  p90 {np.percentile(ms, 90) * 1e3:.2f}")
    sb.close()
