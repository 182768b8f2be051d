"""GPU: BatchedHostSolver (quadruped_pympc_tamols_b200/batched_host.py) -- host buffers in and out, robots cut into
chunks that alternate between two handles / streams so that copies overlap the rollouts -- against ONE
qpmpc_solve_batched launch over all robots on device tensors.

Robot r draws noise subsequence robot_offset + r whatever the chunking.  The solver must equal, bit for bit, the
same chunks launched one after the other with device tensors; against ONE launch over all robots the rollouts are the
same but the launch plan (rollout kernel, blocks per robot) may differ with the number of robots in a launch, so only
the best cost is compared there, to float32 round-off."""
import numpy as np
import pytest
import torch

from quadruped_pympc_tamols_b200 import _cabi
from quadruped_pympc_tamols_b200.batched_host import BatchedHostSolver
from tests.test_sampling_parity_gpu import ROBOTS, synth_problem

pytestmark = pytest.mark.gpu


def dev(a, dt=np.float32):
    return torch.as_tensor(np.ascontiguousarray(a, dt), device="cuda")


@pytest.mark.parametrize("method,param,ga,R,n,chunks", [
    ("mppi", "zero_order", False, 37, 1024, 4),       # the C5 shape (pair kernel), ragged chunks
    ("cem_mppi", "cubic_spline", False, 9, 600, 3),   # sigma in and out
    ("mppi", "cubic_spline", True, 10, 512, 4),       # per-sample gait: timing / nominal frequency / swing flag
    ("random_sampling", "linear_spline", False, 5, 300, 1),  # one chunk = the plain sequence on one handle
])
def test_chunked_host_solver_equals_one_device_launch(method, param, ga, R, n, chunks):
    rng = np.random.default_rng(R * 1000 + n)
    mass, inertia, _, _ = ROBOTS["aliengo"]
    cfg = dict(mass=mass, inertia=np.array(inertia), sampling_method=method, parametrization=param, gait_adaptive=ga,
               num_parallel_computations=n)
    probs = [synth_problem("aliengo", "trot", rng) for _ in range(R)]
    state, ref = np.stack([p[0] for p in probs]).astype(np.float32), np.stack([p[1] for p in probs]).astype(np.float32)
    seq, timing = np.stack([p[2] for p in probs]).astype(np.float32), np.stack([p[3] for p in probs]).astype(np.float32)
    nominal = np.array([p[4] for p in probs], np.float32)
    oswing = (np.arange(R) % 2).astype(np.int32)
    offset0 = 11  # robot_offset: a rank that owns robots 11 .. 11 + R - 1

    s = BatchedHostSolver(R, chunks=chunks, robot_offset=offset0, **cfg)
    P = s.P
    best = rng.normal(size=(R, P)).astype(np.float32)
    sigma = np.full((R, P), 3.0, np.float32)
    assert len(s.bounds) == chunks and s.bounds[0][0] == 0 and s.bounds[-1][1] == R
    outs = []
    for it in range(2):  # twice: the pinned arrays and the alternation of the handles are reused
        out = s.solve(state, ref, seq, best, 5, 20 + it, sigma=sigma if method == "cem_mppi" else None,
                      timing=timing if ga else None, nominal_freq=nominal if ga else None, optimize_swing=oswing if ga else None)
        outs.append([o.copy() for o in out])
    assert s.launch_count >= 2 * chunks
    bounds = list(s.bounds)
    h2d, d2h = s.io_bytes()
    assert h2d >= R * 4 * (48 + 4 * 12 + P) and d2h >= R * 4 * (_cabi.RESULT_FLOATS + P)
    s.close()

    h = _cabi.Handle(_cabi.make_config(max_robots=R, **cfg))
    d_res = torch.zeros((R, _cabi.RESULT_FLOATS), device="cuda")
    d_bo, d_so = torch.zeros((R, P), device="cuda"), torch.zeros((R, P), device="cuda")
    keep = [dev(state), dev(ref), dev(seq), dev(best), dev(sigma), dev(timing), dev(nominal), dev(oswing, np.int32)]

    def launch(lo, hi, it):
        sl = lambda t: t[lo:hi].data_ptr()  # noqa: E731
        b = _cabi.QpmpcBatch(num_robots=hi - lo, robot_offset=offset0 + lo, d_state=sl(keep[0]), d_reference=sl(keep[1]),
                             d_contact_sequence=sl(keep[2]), d_best_in=sl(keep[3]), d_result=sl(d_res), d_best_out=sl(d_bo),
                             d_sigma_in=sl(keep[4]) if method == "cem_mppi" else None,
                             d_sigma_out=sl(d_so) if method == "cem_mppi" else None,
                             d_timing=sl(keep[5]) if ga else None, d_nominal_freq=sl(keep[6]) if ga else None,
                             d_optimize_swing=sl(keep[7]) if ga else None)
        h.solve_batched(b, 5, 20 + it)
        torch.cuda.synchronize()

    for it in range(2):
        res, bo = outs[it][0], outs[it][1]
        # (1) the same chunks launched one after the other on one handle with device tensors: bit for bit
        for lo, hi in bounds:
            launch(lo, hi, it)
        assert np.array_equal(res.view(np.int32), d_res.cpu().numpy().view(np.int32))
        assert np.array_equal(bo, d_bo.cpu().numpy())
        if method == "cem_mppi":
            assert np.array_equal(outs[it][2], d_so.cpu().numpy())
        # (2) ONE launch over all robots: the same rollouts (a launch of another size may run another rollout kernel
        # and merges in another order: best cost to float32 round-off, record column 48)
        launch(0, R, it)
        wres = d_res.cpu().numpy()
        assert np.max(np.abs(res[:, 48] - wres[:, 48]) / np.abs(wres[:, 48])) < 5e-6
        assert np.array_equal(res[:, 49], wres[:, 49])  # best_freq
    h.close()


def test_carry_best_swaps_the_page_locked_arrays():
    """carry_best(): the returned mean is the next input without a host copy -- same results as passing a copy back in;
    a view of the outputs that outlives close() stays readable (the arrays are registered numpy memory)."""
    rng = np.random.default_rng(4)
    mass, inertia, _, _ = ROBOTS["go2"]
    cfg = dict(mass=mass, inertia=np.array(inertia), sampling_method="cem_mppi", parametrization="cubic_spline",
               num_parallel_computations=512)
    R = 6
    probs = [synth_problem("go2", "trot", rng) for _ in range(R)]
    state, ref, seq = (np.stack([p[k] for p in probs]).astype(np.float32) for k in range(3))
    runs = []
    for carry in (False, True):
        s = BatchedHostSolver(R, chunks=3, **cfg)
        best, sigma = np.zeros((R, s.P), np.float32), np.full((R, s.P), 3.0, np.float32)
        for it in range(3):
            if carry and it > 0:
                s.carry_best()
                res, best, sigma = s.solve(state, ref, seq, None, 3, it)
            else:
                res, best, sigma = (o.copy() for o in s.solve(state, ref, seq, best, 3, it, sigma=sigma))
        runs.append((res.copy(), best.copy(), sigma.copy()))
        s.close()
        assert np.isfinite(best).all() and float(np.abs(best).max()) > 0  # still readable after close()
    for a, b in zip(*runs):
        assert np.array_equal(a, b)
