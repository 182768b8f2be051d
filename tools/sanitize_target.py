#!/usr/bin/env python
"""compute-sanitizer target: every kernel family once at a reduced size (the tool slows kernels 10-100x).
Usage: compute-sanitizer --tool {memcheck,racecheck,synccheck,initcheck} python tools/sanitize_target.py"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import synthetic as sy  # noqa: E402
from quadruped_pympc_tamols_b200 import _cabi  # noqa: E402
from quadruped_pympc_tamols_b200 import _tamols_cabi as tc  # noqa: E402

dev = torch.device("cuda", 0)
t_ = lambda a, dt=torch.float32: torch.as_tensor(np.ascontiguousarray(a), dtype=dt, device=dev)  # noqa: E731
rng = np.random.default_rng(1)

# single-robot solves: quad kernel + fused finalize, all three methods (CEM: radix select + elite rows), chained iterations
for key in ("c1", "c2", "c3"):
    c = bench.CONFIGS[key]
    st, ref, seq, timing, nominal = bench.config_problem(key, 5)
    h = _cabi.Handle(_cabi.make_config(sampling_method=c["method"], parametrization=c["param"], gait_adaptive=c["ga"],
                                       num_parallel_computations=10000, **c["robot"]))
    best = np.zeros(h.P, np.float32)
    sig = np.full(h.P, 3.0, np.float32) if c["method"] == "cem_mppi" else None
    for it in range(2):
        r = h.solve(st, ref, seq, best, 42, it, timing=timing, nominal_freq=nominal, optimize_swing=1, sigma_in=sig,
                    iterations=c["iterations"])
        best = r["best"]
    print(key, h.last_kernel, "best cost", r["best_cost"])
    h.close()

# batched robots: pair kernel fused (one block per robot), pair kernel + k_pair_update, multi-block pair with ticket,
# thread kernel (cubic) with ticket and with k_finalize
for name, env, param, R, n in (("pair fused", {}, "zero_order", 96, 1024), ("pair + update", {"QPMPC_FUSED": "0"}, "zero_order", 96, 1024),
                               ("pair multi-block", {"QPMPC_KERNEL": "thread"}, "zero_order", 2, 4097),
                               ("thread cubic", {"QPMPC_KERNEL": "thread"}, "cubic_spline", 8, 1000),
                               ("thread cubic 2 launches", {"QPMPC_KERNEL": "thread", "QPMPC_FUSED": "0"}, "cubic_spline", 8, 1000)):
    for k in ("QPMPC_FUSED", "QPMPC_KERNEL"):
        os.environ.pop(k, None)
    os.environ.update(env)
    for method in ("mppi", "cem_mppi", "random_sampling"):
        h = _cabi.Handle(_cabi.make_config(sampling_method=method, parametrization=param, num_parallel_computations=n,
                                           max_robots=R, **bench.ALIENGO))
        P = h.P
        probs = [bench.synth_problem(rng, 0.33, (0.24, 0.15), gait=([0.5, 1.0, 1.0, 0.5], 1.4, 0.65)) for _ in range(R)]
        d_state, d_ref, d_seq = t_(np.stack([p[0] for p in probs])), t_(np.stack([p[1] for p in probs])), t_(np.stack([p[2] for p in probs]))
        d_best = [torch.zeros((R, P), device=dev), torch.zeros((R, P), device=dev)]
        d_sig = [torch.full((R, P), 3.0, device=dev), torch.full((R, P), 3.0, device=dev)]
        d_res = torch.zeros((R, _cabi.RESULT_FLOATS), device=dev)
        for it in range(2):
            b = _cabi.QpmpcBatch(num_robots=R, d_state=d_state.data_ptr(), d_reference=d_ref.data_ptr(),
                                 d_contact_sequence=d_seq.data_ptr(), d_best_in=d_best[it & 1].data_ptr(),
                                 d_result=d_res.data_ptr(), d_best_out=d_best[(it + 1) & 1].data_ptr(),
                                 d_sigma_in=d_sig[it & 1].data_ptr() if method == "cem_mppi" else None,
                                 d_sigma_out=d_sig[(it + 1) & 1].data_ptr() if method == "cem_mppi" else None)
            h.solve_batched(b, 7, it)
        torch.cuda.synchronize()
        print(name, method, h.last_kernel, "mean best cost", float(d_res[:, 48].mean()))
        h.close()
for k in ("QPMPC_FUSED", "QPMPC_KERNEL"):
    os.environ.pop(k, None)

# TAMOLS: multi-tile jobs (ticket + merge), single-tile host call (mapped result pairs), both tile heights
p = tc.make_params(sy.tamols_params_for("go2"), "go2")
for rows_env in ("8", "32"):
    os.environ["TAMOLS_TILE_ROWS"] = rows_env
    sc = sy.sweep_scene("go2", np.random.default_rng(90), "stepping_stones_tall", window=64, seed=4, robots=8)
    d_h = torch.from_numpy(np.ascontiguousarray(sc["heights"])).to(dev)
    d_jobs = torch.from_numpy(np.frombuffer(sc["jobs"].tobytes(), np.uint8).copy()).to(dev)
    d_res = torch.zeros(32 * tc.RESULT_DTYPE.itemsize, dtype=torch.uint8, device=dev)
    h = tc.Handle(p, max_jobs=32)
    for it in range(2):
        h.score_device(d_h.data_ptr(), tc.F64, 1, 2048, 2048, 2048, d_jobs.data_ptr(), 32, 64, 64, d_res.data_ptr())
    torch.cuda.synchronize()
    s = sy.robot_scene("hyqreal1", np.random.default_rng(3), "stepping_stones")
    hh = tc.Handle(tc.make_params(sy.tamols_params_for("hyqreal1"), "hyqreal1"), max_jobs=4)
    res, _ = hh.score_host(s["heights"], s["jobs"], xy=s["xy"])
    print("tamols tile rows", rows_env, "best", res["best_index"].tolist())
    h.close()
    hh.close()
print("sanitize target done")
