import sys, os
sys.path.insert(0, '/root/repo')
import numpy as np, bench
from quadruped_pympc_tamols_b200 import _cabi
_cabi.LIB_PATH = os.path.join(os.path.dirname(_cabi.LIB_PATH), 'timing', 'libqpmpc_b200.so')  # -DQPMPC_TIMING build
st, ref, seq, timing, nominal = bench.config_problem("c3", 0)
c = bench.CONFIGS["c3"]
h = _cabi.Handle(_cabi.make_config(sampling_method="cem_mppi", parametrization="cubic_spline", gait_adaptive=False, num_parallel_computations=10000, **c["robot"]))
best = np.zeros(h.P, np.float32)
sig = np.full(h.P, 3.0, np.float32)
for it in range(130):
    r = h.solve(st, ref, seq, best, 42, it, timing=timing, nominal_freq=nominal, optimize_swing=1, sigma_in=sig, want_costs=False, iterations=1)
    best = r["best"]
    if it % 32 == 31: print("phases", r["footholds"][1:5], "block: rollout, ticket, finalize", r["footholds"][6:9], "stage, noise, loop", r["footholds"][9:12])
