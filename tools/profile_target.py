#!/usr/bin/env python
"""Small ncu target: a few launches of the hot kernels at bench shapes (C2 single solve, C5 batched, TAMOLS batched).
Usage: python tools/profile_target.py [c2|c3|c5|tamols] [reps]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from quadruped_pympc_tamols_b200 import _cabi  # noqa: E402

what = sys.argv[1] if len(sys.argv) > 1 else "c5"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
dev = torch.device("cuda", 0)
t_ = lambda a, dt=torch.float32: torch.as_tensor(np.ascontiguousarray(a), dtype=dt, device=dev)  # noqa: E731
rng = np.random.default_rng(1)
if what == "c2":
    st, ref, seq, timing, nominal = bench.synth_problem(rng, 0.27, (0.19, 0.13))
    h = _cabi.Handle(_cabi.make_config(sampling_method="mppi", parametrization="cubic_spline", gait_adaptive=True,
                                       num_parallel_computations=10000, **bench.GO2))
    best = np.zeros(h.P, np.float32)
    for it in range(reps):
        r = h.solve(st, ref, seq, best, 42, it, timing=timing, nominal_freq=nominal, optimize_swing=1, want_costs=False)
        best = r["best"]
    print("c2 best cost", r["best_cost"])
elif what == "c3":
    st, ref, seq, timing, nominal = bench.config_problem("c3", 0)
    h = _cabi.Handle(_cabi.make_config(sampling_method="cem_mppi", parametrization="cubic_spline", gait_adaptive=False,
                                       num_parallel_computations=10000, **bench.HYQREAL1))
    best, sig = np.zeros(h.P, np.float32), np.full(h.P, 3.0, np.float32)
    for it in range(reps):
        r = h.solve(st, ref, seq, best, 42, it, timing=timing, nominal_freq=nominal, optimize_swing=1, sigma_in=sig, want_costs=False)
        best, sig = r["best"], r["sigma"]
    print("c3 best cost", r["best_cost"])
elif what == "c5":
    R, n = 8192, 1024
    h = _cabi.Handle(_cabi.make_config(sampling_method="mppi", parametrization="zero_order", num_parallel_computations=n,
                                       max_robots=R, **bench.ALIENGO))
    P = h.P
    probs = [bench.synth_problem(rng, 0.33, (0.24, 0.15), gait=([0.5, 1.0, 1.0, 0.5], 1.4, 0.65)) for _ in range(64)]
    idx = rng.integers(0, 64, R)
    d_state = t_(np.stack([probs[i][0] for i in idx]))
    d_ref, d_seq = t_(np.stack([probs[i][1] for i in idx])), t_(np.stack([probs[i][2] for i in idx]))
    d_best = [torch.zeros((R, P), device=dev), torch.zeros((R, P), device=dev)]
    d_res = torch.zeros((R, _cabi.RESULT_FLOATS), device=dev)
    for it in range(reps):
        b = _cabi.QpmpcBatch(num_robots=R, d_state=d_state.data_ptr(), d_reference=d_ref.data_ptr(),
                             d_contact_sequence=d_seq.data_ptr(), d_best_in=d_best[it & 1].data_ptr(),
                             d_result=d_res.data_ptr(), d_best_out=d_best[(it + 1) & 1].data_ptr())
        h.solve_batched(b, 7, it)
    torch.cuda.synchronize()
    print("c5 mean best cost", float(d_res[:, 48].mean()))
else:
    import synthetic as sy
    from quadruped_pympc_tamols_b200 import _tamols_cabi as tc
    R, w = 2048, 64
    p = tc.make_params(sy.tamols_params_for("go2"), "go2")
    sc = sy.sweep_scene("go2", np.random.default_rng(90), "stepping_stones_tall", window=w, seed=4, robots=R)
    d_h = torch.from_numpy(np.ascontiguousarray(sc["heights"])).to(dev)
    d_jobs = torch.from_numpy(np.frombuffer(sc["jobs"].tobytes(), np.uint8).copy()).to(dev)
    d_res = torch.zeros(4 * R * tc.RESULT_DTYPE.itemsize, dtype=torch.uint8, device=dev)
    h = tc.Handle(p, max_jobs=4 * R)
    for it in range(reps):
        h.score_device(d_h.data_ptr(), tc.F64, 1, 2048, 2048, 2048, d_jobs.data_ptr(), 4 * R, w, w, d_res.data_ptr())
    torch.cuda.synchronize()
    print("tamols done", h.launch_count)
h.close()
