#!/usr/bin/env python
"""How many samples of a C5 robot carry a non-zero softmin weight (the rows k_rollout_pair draws again), per step."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from quadruped_pympc_tamols_b200 import _cabi
dev = torch.device("cuda", 0)
R, n = 8192, 1024
rng = np.random.default_rng(77)
h = _cabi.Handle(_cabi.make_config(sampling_method="mppi", parametrization="zero_order", num_parallel_computations=n, max_robots=R, **bench.ALIENGO))
P = h.P
probs = [bench.synth_problem(rng, 0.33, (0.24, 0.15), gait=([0.5, 1.0, 1.0, 0.5], 1.4, 0.65)) for _ in range(64)]
idx = rng.integers(0, 64, R)
t_ = lambda a: torch.as_tensor(np.ascontiguousarray(a), dtype=torch.float32, device=dev)
d_state = t_(np.stack([probs[i][0] for i in idx]) + rng.normal(0, 1e-3, (R, 24)).astype(np.float32))
d_ref, d_seq = t_(np.stack([probs[i][1] for i in idx])), t_(np.stack([probs[i][2] for i in idx]))
d_best = [torch.zeros((R, P), device=dev), torch.zeros((R, P), device=dev)]
d_res = torch.zeros((R, _cabi.RESULT_FLOATS), device=dev)
d_costs = torch.zeros((R, n), device=dev)
stream = torch.cuda.ExternalStream(h.stream, device=dev)
h.set_profiling(True)
with torch.cuda.stream(stream):
    for it in range(13):
        b = _cabi.QpmpcBatch(num_robots=R, d_state=d_state.data_ptr(), d_reference=d_ref.data_ptr(), d_contact_sequence=d_seq.data_ptr(),
                             d_best_in=d_best[it & 1].data_ptr(), d_result=d_res.data_ptr(), d_best_out=d_best[(it + 1) & 1].data_ptr(), d_costs=d_costs.data_ptr())
        h.solve_batched(b, 7, it)
        ms = h.profile_last()
        torch.cuda.synchronize()
        c = d_costs.double()
        w = torch.exp(-(c - c.min(1, keepdim=True).values).float())
        live = (w > 0).sum(1).float()
        d = (c - c.min(1, keepdim=True).values)
        l24, l16 = (d < 24).sum(1).float(), (d < 16.6).sum(1).float()
        print(f"      d<24: mean {l24.mean():.1f} p50 {l24.median():.0f} p99 {l24.quantile(0.99):.0f} max {l24.max():.0f} | d<16.6: mean {l16.mean():.1f} p99 {l16.quantile(0.99):.0f} max {l16.max():.0f} | median d {d.median():.0f}")
        print(f"it {it}: {ms[0]:.3f}+{ms[1]:.3f} ms  kernel {h.last_kernel}  live per robot: mean {live.mean():.1f} p50 {live.median():.0f} p99 {live.quantile(0.99):.0f} max {live.max():.0f}  mean cost {c.mean():.0f} min {c.min(1).values.mean():.0f}")
h.close()
