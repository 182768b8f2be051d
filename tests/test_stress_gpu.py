"""GPU stress tests of the hand-rolled synchronisation (compute-sanitizer is closed on this pool, so
races are hunted by determinism instead): the arrival tickets + fences of the fused finalize, the shared-memory scratch
the finalize overlays on live columns, and the {word, sequence} result pairs the host polls in mapped memory.
100,000 back-to-back host-buffer solves alternate between two different problems; every call must return bit-for-bit
the outputs its problem produced the first time (a stale pair, a lost ticket or a scratch race would show up as the
other problem's words or as run-to-run differences)."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
import synthetic as sy  # noqa: E402
from quadruped_pympc_tamols_b200 import _cabi  # noqa: E402
from tests.test_sampling_parity_gpu import ROBOTS, synth_problem  # noqa: E402

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("method,param,ga,calls", [("mppi", "cubic_spline", True, 100000), ("cem_mppi", "cubic_spline", False, 8000),
                                                   ("random_sampling", "zero_order", False, 20000)])
def test_alternating_solves_never_read_stale_results(method, param, ga, calls):
    mass, inertia, _, _ = ROBOTS["go2"]
    h = _cabi.Handle(_cabi.make_config(mass=mass, inertia=np.array(inertia), sampling_method=method, parametrization=param,
                                       gait_adaptive=ga, num_parallel_computations=10000))
    rng = np.random.default_rng(8)
    probs = [synth_problem("go2", g, rng) for g in ("trot", "bound")]
    bests = [rng.normal(size=h.P).astype(np.float32) for _ in probs]
    sig = np.full(h.P, 3.0, np.float32) if method == "cem_mppi" else None

    def call(k):
        st, ref, seq, timing, nominal = probs[k]
        r = h.solve(st, ref, seq, bests[k], 11, 3 + k, timing=timing, nominal_freq=nominal, optimize_swing=1, sigma_in=sig,
                    want_costs=False)
        return np.concatenate([r["grf"], r["predicted_state"], r["best"], [r["best_cost"], r["best_freq"], float(r["best_index"])],
                               r["sigma"] if sig is not None else []]).astype(np.float32)

    want = [call(0), call(1)]
    assert not np.array_equal(want[0], want[1])
    bad = 0
    for it in range(calls):
        k = (it * 7 + it // 3) & 1  # an irregular A / B pattern
        got = call(k)
        if got.tobytes() != want[k].tobytes():
            bad += 1
    assert bad == 0, f"{bad} of {calls} calls returned other words than the first call of their problem"
    h.close()


def test_alternating_tamols_host_calls_never_read_stale_results():
    from quadruped_pympc_tamols_b200 import _tamols_cabi as tc
    rng = np.random.default_rng(12)
    scenes = [sy.robot_scene("hyqreal1", rng, "stepping_stones") for _ in range(2)]
    h = tc.Handle(tc.make_params(sy.tamols_params_for("hyqreal1"), "hyqreal1"), max_jobs=4)
    want = [h.score_host(s["heights"], s["jobs"], xy=s["xy"])[0].tobytes() for s in scenes]
    assert want[0] != want[1]
    bad = 0
    for it in range(50000):
        k = (it * 5 + it // 7) & 1
        s = scenes[k]
        if h.score_host(s["heights"], s["jobs"], xy=s["xy"])[0].tobytes() != want[k]:
            bad += 1
    assert bad == 0
    h.close()


def test_batched_pair_kernel_is_deterministic():
    """k_rollout_pair (fused, one block per robot) and k_rollout_pair + k_pair_update: 8 repetitions of one 448-robot
    launch (enough robots for one block per robot) give identical bytes -- block scan, live-row compaction, slice sums
    and the scratch overlay have no race -- and the two variants, which run the same arithmetic in the same order, agree
    bit for bit with each other."""
    import os
    mass, inertia, _, _ = ROBOTS["aliengo"]
    rng = np.random.default_rng(4)
    R, n = 448, 1024
    best_h = rng.normal(size=(R, 144))
    outs = []
    for fused in ("1", "0"):
        os.environ["QPMPC_FUSED"] = fused
        try:
            h = _cabi.Handle(_cabi.make_config(mass=mass, inertia=np.array(inertia), sampling_method="mppi",
                                               parametrization="zero_order", num_parallel_computations=n, max_robots=R))
        finally:
            os.environ.pop("QPMPC_FUSED", None)
        P = h.P
        prob = [synth_problem("aliengo", "trot", np.random.default_rng(100 + r)) for r in range(R)]
        dev = lambda a: torch.as_tensor(np.ascontiguousarray(a, np.float32), device="cuda")  # noqa: E731
        d_state, d_ref, d_seq = dev(np.stack([p[0] for p in prob])), dev(np.stack([p[1] for p in prob])), dev(np.stack([p[2] for p in prob]))
        d_best = dev(best_h)
        first = None
        for rep in range(8):
            d_res = torch.zeros((R, _cabi.RESULT_FLOATS), device="cuda")
            d_bo = torch.zeros((R, P), device="cuda")
            b = _cabi.QpmpcBatch(num_robots=R, d_state=d_state.data_ptr(), d_reference=d_ref.data_ptr(),
                                 d_contact_sequence=d_seq.data_ptr(), d_best_in=d_best.data_ptr(), d_result=d_res.data_ptr(),
                                 d_best_out=d_bo.data_ptr())
            h.solve_batched(b, 3, 9)
            torch.cuda.synchronize()
            got = d_res.cpu().numpy().tobytes() + d_bo.cpu().numpy().tobytes()
            assert h.last_kernel == "pair"
            if first is None:
                first = got
            assert got == first, f"repetition {rep} differs (QPMPC_FUSED={fused})"
        outs.append(first)
        h.close()
    # fused and two-launch variants run the same arithmetic in the same order
    assert outs[0] == outs[1]
