"""C5 through host buffers: BatchedHostSolver.solve wall time against the number of chunks."""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from quadruped_pympc_tamols_b200.batched_host import BatchedHostSolver  # noqa: E402

R, n = 8192, 1024
rng = np.random.default_rng(77)
probs = [bench.synth_problem(rng, 0.33, (0.24, 0.15), gait=([0.5, 1.0, 1.0, 0.5], 1.4, 0.65)) for _ in range(64)]
idx = rng.integers(0, 64, R)
st = np.stack([probs[i][0] for i in idx]).astype(np.float32)
ref, seq = np.stack([probs[i][1] for i in idx]).astype(np.float32), np.stack([probs[i][2] for i in idx]).astype(np.float32)
for chunks in (1, 2, 4, 8, 16):
    hs = BatchedHostSolver(R, chunks=chunks, sampling_method="mppi", parametrization="zero_order", num_parallel_computations=n, **bench.ALIENGO)
    hs.np_in["state"][...], hs.np_in["reference"][...], hs.np_in["contact"][...] = st, ref, seq
    w = []
    for it in range(13):
        t0 = time.perf_counter()
        hs.solve(None, None, None, None, 7, it)
        w.append(time.perf_counter() - t0)
    print(f"chunks {chunks:2d}: {np.median(w[3:]) * 1e3:.3f} ms per step (launches per step {hs.launch_count // 13})")
    hs.close()
