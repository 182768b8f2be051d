"""GPU: the drop-in classes driven exactly the way the reference's callers drive them
(interfaces/srbd_controller_interface.py:113-180 for Sampling_MPC, interfaces/wb_interface.py:230-246 for
VisualFootholdAdaptation), checked against the CPU oracle fed with the very same noise and against the fixtures
produced by the reference's own TAMOLS code."""
import os

import numpy as np
import pytest

torch = pytest.importorskip("torch")

from oracle import srbd_oracle as so  # noqa: E402  (checker only)
from oracle.tamols_reference import SensorMap  # noqa: E402  (numpy stand-in for gym_quadruped's HeightMap)
from quadruped_pympc_tamols_b200._config import make_config  # noqa: E402
from quadruped_pympc_tamols_b200.helpers.legs_attr import LegsAttr  # noqa: E402
from quadruped_pympc_tamols_b200.helpers.visual_foothold_adaptation import VisualFootholdAdaptation  # noqa: E402
from tests.test_sampling_parity_gpu import compare, synth_problem  # noqa: E402
from tests.util import tamols_fixtures  # noqa: E402

pytestmark = pytest.mark.gpu
LEGS = ("FL", "FR", "RL", "RR")


def _dicts(st, ref):
    state = {"position": st[0:3], "linear_velocity": st[3:6], "orientation": st[6:9], "angular_velocity": st[9:12]}
    refd = {"ref_position": ref[0:3], "ref_linear_velocity": ref[3:6], "ref_orientation": ref[6:9],
            "ref_angular_velocity": ref[9:12]}
    for k, f in enumerate(LEGS):
        state["foot_" + f] = st[12 + 3 * k:15 + 3 * k]
        refd["ref_foot_" + f] = ref[12 + 3 * k:15 + 3 * k].reshape(1, 3)
    return state, refd


@pytest.mark.parametrize("robot,gait,method,param,ga,iters", [
    ("aliengo", "trot", "random_sampling", "zero_order", False, 1),
    ("go2", "bound", "mppi", "cubic_spline", True, 1),
    ("hyqreal1", "pace", "cem_mppi", "cubic_spline", False, 2),
    ("aliengo", "trot", "mppi", "linear_spline", False, 3),
])
def test_sampling_mpc_driven_like_the_interface(robot, gait, method, param, ga, iters):
    if ga:
        from quadruped_pympc_tamols_b200.controllers.sampling.centroidal_nmpc_jax_gait_adaptive import Sampling_MPC
    else:
        from quadruped_pympc_tamols_b200.controllers.sampling.centroidal_nmpc_jax import Sampling_MPC
    n = 4096
    cfg = make_config(robot, sampling_method=method, control_parametrization=param, optimize_step_freq=ga,
                      num_parallel_computations=n, num_sampling_iterations=iters)
    controller = Sampling_MPC(config=cfg)
    ocfg = so.make_cfg(mass=cfg.mass, inertia=cfg.inertia, parametrization=param, gait_adaptive=ga)
    rng = np.random.default_rng(5)
    st, ref, seq, timing, freq = synth_problem(robot, gait, rng)
    state_current, ref_state = _dicts(st, ref)
    previous_contact = np.array([1, 1, 1, 1])
    P = controller.num_control_parameters
    controller.best_control_parameters = rng.normal(0, 1, P).astype(np.float32)
    d_eps = torch.empty((n, P), dtype=torch.float32, device="cuda")
    d_freqs = torch.empty(n, dtype=torch.float32, device="cuda")

    # ---- srbd_controller_interface.py:113-168
    current_contact = np.array([seq[0][0], seq[1][0], seq[2][0], seq[3][0]])
    state_jax, reference_jax = controller.prepare_state_and_reference(state_current, ref_state, current_contact,
                                                                      previous_contact)
    for it in range(controller.num_sampling_iterations):
        # each iteration is checked on its own: the oracle starts from the same best the CUDA path starts from
        o_best = np.array(controller.best_control_parameters, np.float32)
        controller = controller.with_newkey()
        key = controller.master_key
        if controller.sampling_method == "cem_mppi":
            if it == 0:
                controller = controller.with_newsigma(cfg.mpc_params["sigma_cem_mppi"])
            sigma_in = np.broadcast_to(np.asarray(controller.sigma_cem_mppi, np.float32), (P,))
            out = controller.jitted_compute_control(state_jax, reference_jax, seq, controller.best_control_parameters,
                                                    key, controller.sigma_cem_mppi)
            (grf, footholds, pred, controller.best_control_parameters, best_cost, best_freq, costs, sigma) = out
            controller = controller.with_newsigma(sigma)
        else:
            sigma_in = None
            out = controller.jitted_compute_control(state_jax, reference_jax, seq, controller.best_control_parameters,
                                                    key, timing, freq, 1)
            (grf, footholds, pred, controller.best_control_parameters, best_cost, best_freq, costs) = out
            sigma = None
        # ---- the oracle on the same noise (dumped from the same Philox key)
        controller._handle.generate_noise(int(key[0]), int(key[1]), d_eps.data_ptr(), d_freqs.data_ptr(),
                                          sigma_in=sigma_in, nominal_freq=freq, optimize_swing=1)
        o = so.solve(ocfg, method, state_jax, reference_jax, seq, o_best, d_eps.cpu().numpy(),
                     timing=timing if ga else None, freqs=d_freqs.cpu().numpy() if ga else None)
        r = dict(grf=np.asarray(grf), footholds=np.asarray(footholds), predicted_state=np.asarray(pred),
                 best=np.asarray(controller.best_control_parameters), best_cost=best_cost, best_freq=best_freq,
                 costs=np.asarray(costs), sigma=sigma, best_index=int(np.argmin(costs)))
        compare(r, o, method, tag=f"iteration {it}", eps=d_eps.cpu().numpy(), best_in=o_best)
    assert np.asarray(grf).shape == (12,) and np.asarray(pred).shape == (24,) and np.asarray(costs).shape == (n,)
    # swing legs of the passed-in sequence carry no force in the returned GRFs
    for leg in range(4):
        if current_contact[leg] == 0:
            assert np.all(np.asarray(grf)[3 * leg:3 * leg + 3] == 0)
    controller.close()


def test_vectorized_rollout_matches_oracle():
    from quadruped_pympc_tamols_b200.controllers.sampling.centroidal_nmpc_jax import Sampling_MPC
    cfg = make_config("go2", sampling_method="mppi", control_parametrization="cubic_spline",
                      num_parallel_computations=512)
    c = Sampling_MPC(config=cfg)
    rng = np.random.default_rng(11)
    st, ref, seq, _, _ = synth_problem("go2", "trot", rng)
    params = rng.normal(0, 3, (300, c.num_control_parameters)).astype(np.float32)
    costs = c.jit_vectorized_rollout(st, ref, params, seq)
    ocfg = so.make_cfg(mass=cfg.mass, inertia=cfg.inertia, parametrization="cubic_spline")
    oc = so.rollout_costs(ocfg, st, ref, seq, params)
    assert np.max(np.abs(costs - oc) / np.maximum(np.abs(oc), 1e-6)) < 1e-4
    assert abs(c.compute_rollout(st, ref, params[7], seq) - oc[7]) <= 1e-4 * abs(oc[7])
    c.close()


@pytest.mark.parametrize("path", tamols_fixtures(), ids=lambda p: os.path.basename(p)[7:-4])
def test_visual_foothold_adaptation_matches_reference_fixture(path):
    g = np.load(path)
    jobs, heights, xy = g["jobs"], g["heights"], g["xy"]
    if len(jobs) != 4:
        pytest.skip("fixture is not a 4-leg compute_adaptation call")
    robot = str(g["robot"])
    from quadruped_pympc_tamols_b200 import _tamols_cabi as tc
    ps = tc.TamolsParams.from_buffer_copy(g["params"].tobytes())
    tp = {k: getattr(ps, k) for k in ("gradient_delta", "slope_threshold", "weight_edge_avoidance", "weight_roughness",
                                      "weight_deviation", "weight_nominal_kinematic", "weight_reference_tracking",
                                      "weight_stability", "stability_margin", "estimated_swing_time", "h_des",
                                      "constraint_box_dx", "constraint_box_dy")}
    tp["l_min"], tp["l_max"] = {robot: ps.l_min}, {robot: ps.l_max}
    cfg = make_config(robot)
    cfg.simulation_params["tamols_params"] = tp
    vfa = VisualFootholdAdaptation(legs_order=LEGS, adaptation_strategy="tamols", config=cfg)
    maps, seeds, hips = {}, {}, {}
    for l, name in enumerate(LEGS):
        j = jobs[l]
        sl = (int(j["map"]), slice(int(j["r0"]), int(j["r0"] + j["nr"])), slice(int(j["c0"]), int(j["c0"] + j["nc"])))
        maps[name] = SensorMap(xy[sl], heights[sl])
        seeds[name], hips[name] = np.array(j["seed"], np.float64), np.array(j["hip"], np.float64)
    has_vel = bool(jobs[0]["flags"] & 1)
    vel = np.array(jobs[0]["vel"], np.float64) if has_vel else None
    stab = [bool(j["flags"] & 2) for j in jobs]
    base, feet, contact = None, None, None
    if any(stab):
        k = stab.index(True)
        base = np.array([jobs[k]["base_xy"][0], jobs[k]["base_xy"][1], 0.0])
        partner = {"FL": "RR", "FR": "RL", "RL": "FR", "RR": "FL"}
        fp = {n: np.zeros(3) for n in LEGS}
        for l, name in enumerate(LEGS):
            if stab[l]:
                fp[partner[name]] = np.array([jobs[l]["diag_foot"][0], jobs[l]["diag_foot"][1], 0.0])
        feet = LegsAttr(**fp)
        contact = np.array([0 if s else 1 for s in stab])
    elif "contact" in g:
        contact = g["contact"]
    ref = LegsAttr(**{n: seeds[n].copy() for n in LEGS})
    assert vfa.get_footholds_adapted(ref)[0] is ref
    ok = vfa.compute_adaptation(LEGS, ref, LegsAttr(**hips), LegsAttr(**maps), vel, np.zeros(3), np.zeros(3),
                                base_position=base, current_contact=contact, current_feet_pos=feet,
                                phase_signal=np.zeros(4))
    assert ok is True and vfa.initialized
    ref_scores, ref_idx = g["ref_scores"], g["ref_best_index"]
    for l, name in enumerate(LEGS):
        got = np.asarray(ref[name], np.float64)
        if np.array_equal(got, g["ref_best_xyz"][l]):
            if ref_idx[l] >= 0:
                assert np.array_equal(vfa.footholds_constraints[name][0], g["ref_vertex1"][l])
                assert np.array_equal(vfa.footholds_constraints[name][1], g["ref_vertex2"][l])
            continue
        # otherwise only a documented sub-1e-6 score tie may have moved the argmin
        assert ref_idx[l] >= 0
        nc = int(jobs[l]["nc"])
        cand = xy[int(jobs[l]["map"]), int(jobs[l]["r0"]):int(jobs[l]["r0"] + jobs[l]["nr"]),
                  int(jobs[l]["c0"]):int(jobs[l]["c0"]) + nc].reshape(-1, 2)
        k = int(np.argmin(np.sum((cand - got[:2]) ** 2, axis=1)))
        assert abs(ref_scores[l, k] - ref_scores[l, ref_idx[l]]) < 1e-6
    assert vfa.get_footholds_adapted(None)[0] is ref
    vfa.close()


@pytest.mark.parametrize("robot,gait,method,param,ga,iters", [
    ("go2", "bound", "mppi", "cubic_spline", True, 3),
    ("hyqreal1", "pace", "cem_mppi", "cubic_spline", False, 3),
    ("aliengo", "trot", "random_sampling", "zero_order", False, 2),
    ("aliengo", "trot", "mppi", "linear_spline", False, 1),
])
def test_interface_chained_iterations_equal_the_python_loop(robot, gait, method, param, ga, iters):
    """SRBDControllerInterface.compute_control with the num_sampling_iterations loop chained on the
    device (qpmpc_solve_iterated) against the upstream loop replayed call by call on a second controller
    (srbd_controller_interface.py:117-183): bit-identical GRFs, prediction, mean, sigma, key; two consecutive control
    steps, the second one with a lift-off edge (parameter block zeroing)."""
    from quadruped_pympc_tamols_b200.interfaces.srbd_controller_interface import SRBDControllerInterface
    if ga:
        from quadruped_pympc_tamols_b200.controllers.sampling.centroidal_nmpc_jax_gait_adaptive import Sampling_MPC
    else:
        from quadruped_pympc_tamols_b200.controllers.sampling.centroidal_nmpc_jax import Sampling_MPC
    cfg = make_config(robot, sampling_method=method, control_parametrization=param, optimize_step_freq=ga,
                      num_parallel_computations=4096, num_sampling_iterations=iters)
    itf = SRBDControllerInterface(config=cfg)
    loop = Sampling_MPC(config=cfg)
    rng = np.random.default_rng(11)
    previous_contact = np.array([1, 1, 1, 1])
    for step in range(2):
        st, ref, seq, timing, freq = synth_problem(robot, gait, rng)
        if step == 1:
            seq = np.array(seq)
            seq[1, 0] = 0.0  # FR lifts off between the two control steps
        state_current, ref_state = _dicts(st, ref)
        got = itf.compute_control(state_current, ref_state, seq, cfg.inertia, timing, freq, 1)

        current_contact = np.array([seq[0][0], seq[1][0], seq[2][0], seq[3][0]])
        s_, r_ = loop.prepare_state_and_reference(state_current, ref_state, current_contact, previous_contact)
        previous_contact = current_contact
        for it in range(loop.num_sampling_iterations):
            loop = loop.with_newkey()
            if method == "cem_mppi":
                if it == 0:
                    loop = loop.with_newsigma(cfg.mpc_params["sigma_cem_mppi"])
                out = loop.jitted_compute_control(s_, r_, seq, loop.best_control_parameters, loop.master_key,
                                                  loop.sigma_cem_mppi)
                grf, _, pred, loop.best_control_parameters, best_cost, best_freq, costs, sigma = out
                loop = loop.with_newsigma(sigma)
            else:
                out = loop.jitted_compute_control(s_, r_, seq, loop.best_control_parameters, loop.master_key, timing,
                                                  freq, 1)
                grf, _, pred, loop.best_control_parameters, best_cost, best_freq, costs = out
        grf = np.array(grf)
        for k, leg in enumerate(LEGS):
            assert np.array_equal(got[0][leg], grf[3 * k:3 * k + 3] * current_contact[k])
            assert np.array_equal(got[1][leg], ref_state["ref_foot_" + leg][0])
        assert got[2] is None and got[3] is None and got[4] is None
        assert got[5] == best_freq
        assert np.array_equal(got[6], pred)
        assert np.array_equal(itf.controller.best_control_parameters, loop.best_control_parameters)
        assert np.array_equal(itf.controller.master_key, loop.master_key)
        assert itf.last_best_cost == float(best_cost) and np.array_equal(itf.last_costs, costs)
        if method == "cem_mppi":
            assert np.array_equal(itf.controller.sigma_cem_mppi, loop.sigma_cem_mppi)
    assert itf.controller._handle.launch_count == 2 * iters  # one kernel launch per iteration, nothing else
    itf.close()
    loop.close()


def test_interface_rejects_what_upstream_cannot_run():
    from quadruped_pympc_tamols_b200.interfaces.srbd_controller_interface import SRBDControllerInterface
    with pytest.raises(NotImplementedError):
        SRBDControllerInterface(config=make_config("go2", type="nominal"))
    itf = SRBDControllerInterface(config=make_config("go2", sampling_method="cem_mppi", optimize_step_freq=True,
                                                     num_parallel_computations=256))
    rng = np.random.default_rng(0)
    st, ref, seq, timing, freq = synth_problem("go2", "trot", rng)
    state_current, ref_state = _dicts(st, ref)
    with pytest.raises(TypeError):
        itf.compute_control(state_current, ref_state, seq, None, timing, freq, 1)
    with pytest.raises(NotImplementedError):
        itf.compute_RTI()
    itf.close()


def test_costs_stay_on_the_device_until_read():
    """Sampling_MPC returns its costs like upstream's device array: fetched on first look, and before the next solve
    overwrites them; values equal the eagerly copied costs of the same solve."""
    from quadruped_pympc_tamols_b200 import _cabi
    from quadruped_pympc_tamols_b200.controllers.sampling.centroidal_nmpc_jax import Sampling_MPC
    cfg = make_config("aliengo", sampling_method="mppi", control_parametrization="linear_spline",
                      num_parallel_computations=4096)
    c = Sampling_MPC(config=cfg)
    rng = np.random.default_rng(3)
    st, ref, seq, timing, freq = synth_problem("aliengo", "trot", rng)
    best0 = np.zeros(c.num_control_parameters, np.float32)
    key1, key2 = np.array([42, 1], np.uint64), np.array([42, 2], np.uint64)
    eager1 = c._handle.solve(st, ref, seq, best0, 42, 1, want_costs=True)["costs"]
    eager2 = c._handle.solve(st, ref, seq, best0, 42, 2, want_costs=True)["costs"]
    out1 = c.jitted_compute_control(st, ref, seq, best0, key1, None, 0.0, 0)
    costs1 = out1[6]
    assert isinstance(costs1, _cabi.DeviceCosts) and costs1._host is None and len(costs1) == 4096
    out2 = c.jitted_compute_control(st, ref, seq, best0, key2, None, 0.0, 0)  # must not clobber costs1
    assert costs1._host is not None
    assert np.array_equal(np.asarray(costs1), eager1)
    assert float(np.min(costs1)) == float(out1[4]) and costs1[int(np.argmin(costs1))] == out1[4]
    assert np.array_equal(out2[6] - eager2, np.zeros(4096, np.float32))  # arithmetic goes through __array__
    dropped = c.jitted_compute_control(st, ref, seq, best0, key1, None, 0.0, 0)[6]
    del dropped  # nobody holds it: the next solve fetches nothing
    n0 = c._handle.launch_count
    c.jitted_compute_control(st, ref, seq, best0, key2, None, 0.0, 0)
    assert c._handle.launch_count == n0 + 1
    held = c.jitted_compute_control(st, ref, seq, best0, key1, None, 0.0, 0)[6]
    c.close()  # closing fetches what is still held
    assert np.array_equal(np.asarray(held), eager1)
