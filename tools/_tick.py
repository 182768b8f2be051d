import sys, os
sys.path.insert(0, '/root/repo')
import numpy as np, bench
from quadruped_pympc_tamols_b200 import _cabi
_cabi.LIB_PATH = os.path.join(os.path.dirname(_cabi.LIB_PATH), 'timing', 'libqpmpc_b200.so')  # -DQPMPC_TIMING build
rng = np.random.default_rng(1)
st, ref, seq, timing, nominal = bench.synth_problem(rng, 0.27, (0.19, 0.13))
h = _cabi.Handle(_cabi.make_config(sampling_method="mppi", parametrization="cubic_spline", gait_adaptive=True, num_parallel_computations=10000, **bench.GO2))
best = np.zeros(h.P, np.float32)
for it in range(120):
    r = h.solve(st, ref, seq, best, 42, it, timing=timing, nominal_freq=nominal, optimize_swing=1, want_costs=False)
    best = r["best"]
    if it % 20 == 19: print("phases", r["footholds"][1:5], "block: rollout, ticket, finalize", r["footholds"][6:9], "stage, noise, loop", r["footholds"][9:12])
