import sys, os, zlib
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import srbd_oracle as so
from quadruped_pympc_tamols_b200 import _cabi
from tests.test_sampling_parity_gpu import ROBOTS, synth_problem, CASES
for case in CASES[:4]:
    robot, gait, method, param, ga, n = case
    rng = np.random.default_rng(zlib.crc32(repr(case).encode()))
    mass, inertia, _, _ = ROBOTS[robot]
    kw = dict(mass=mass, inertia=np.array(inertia), parametrization=param, gait_adaptive=ga)
    h = _cabi.Handle(_cabi.make_config(sampling_method=method, num_parallel_computations=n, **kw))
    ocfg = so.make_cfg(**kw)
    P = h.P
    st, ref, seq, timing, nominal = synth_problem(robot, gait, rng)
    best = (2.0 * rng.normal(size=P)).astype(np.float32)
    sigma = np.full(P, 3.0, np.float32)
    d_eps = torch.empty((n, P), dtype=torch.float32, device="cuda"); d_freqs = torch.empty(n, dtype=torch.float32, device="cuda")
    h.generate_noise(42, 0, d_eps.data_ptr(), d_freqs.data_ptr(), sigma_in=sigma if method == "cem_mppi" else None, nominal_freq=nominal, optimize_swing=1)
    eps = d_eps.cpu().numpy(); freqs = d_freqs.cpu().numpy() if ga else None
    r = h.solve(st, ref, seq, best, seed=42, offset=0, timing=timing, nominal_freq=nominal, optimize_swing=1, sigma_in=sigma if method == "cem_mppi" else None)
    o = so.solve(ocfg, method, st, ref, seq, best, eps, timing=timing, freqs=freqs, fast=False)
    c, oc = r["costs"].astype(np.float64), o["costs"].astype(np.float64)
    w = np.exp(-(oc - oc.min())); w /= w.sum()
    live = w > 1e-6
    # update computed in float64 from the GPU costs and from the oracle costs
    wg = np.exp(-(c - c.min())); wg /= wg.sum()
    upd_o64 = best + w @ eps.astype(np.float64); upd_g64 = best + wg @ eps.astype(np.float64)
    print(case, "\n  rel cost err max %.2e, on live samples abs %.2e (cost min %.1f, live %d)" % (np.max(np.abs(c-oc)/np.abs(oc)), np.max(np.abs(c-oc)[live]), oc.min(), live.sum()),
          "\n  best err gpu-vs-oracle %.2e | oracle-vs-f64(oracle costs) %.2e | gpu-vs-f64(gpu costs) %.2e | f64(gpu costs)-vs-f64(oracle costs) %.2e" % (
          np.max(np.abs(r["best"]-o["best"])), np.max(np.abs(o["best"]-upd_o64)), np.max(np.abs(r["best"]-upd_g64)), np.max(np.abs(upd_g64-upd_o64))))
    h.close()
