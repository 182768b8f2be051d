"""CPU: host-side logic of the drop-in classes (no GPU compute): config plumbing, the loud failure without a GPU,
prepare_state_and_reference / shift_solution / key and sigma state of Sampling_MPC, and the state machine of
VisualFootholdAdaptation.  Mirrors how interfaces/srbd_controller_interface.py:113-180 and
interfaces/wb_interface.py:230-246 drive the reference classes."""
import numpy as np
import pytest

from oracle import srbd_oracle as so  # checker only
from quadruped_pympc_tamols_b200 import _cabi, _tamols_cabi
from quadruped_pympc_tamols_b200._config import make_config
from quadruped_pympc_tamols_b200.controllers.sampling import centroidal_nmpc_jax as fixed
from quadruped_pympc_tamols_b200.controllers.sampling import centroidal_nmpc_jax_gait_adaptive as adaptive
from quadruped_pympc_tamols_b200.helpers.legs_attr import LegsAttr
from quadruped_pympc_tamols_b200.helpers.visual_foothold_adaptation import VisualFootholdAdaptation


class _NoDeviceHandle:
    """Stands in for the native handle so that the HOST logic can be exercised on a box without a GPU."""

    def __init__(self, cfg):
        self.cfg, self.N, self.H = cfg, cfg.num_parallel_computations, cfg.horizon
        self.P = so.num_params(so.make_cfg(mass=1.0, inertia=np.eye(3), horizon=cfg.horizon,
                                           parametrization=["zero_order", "linear_spline", "cubic_spline"][cfg.parametrization],
                                           num_splines=cfg.num_splines))

    def close(self):
        pass


@pytest.fixture()
def host_only(monkeypatch):
    monkeypatch.setattr(_cabi, "Handle", _NoDeviceHandle)


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:  # noqa: BLE001
        return False


@pytest.mark.skipif(_has_gpu(), reason="checks the failure mode of a box without a GPU")
def test_product_path_fails_loudly_without_gpu():
    with pytest.raises(_cabi.QpmpcError, match="no CPU fallback|no CUDA device"):
        fixed.Sampling_MPC(config=make_config("go2", sampling_method="mppi"))
    with pytest.raises(_tamols_cabi.TamolsError, match="no CPU fallback|no CUDA device"):
        VisualFootholdAdaptation(("FL", "FR", "RL", "RR"), "tamols", config=make_config("go2"))


@pytest.mark.parametrize("param,P", [("zero_order", 144), ("linear_spline", 36), ("cubic_spline", 96)])
def test_parameter_counts_and_attributes(host_only, param, P):
    c = fixed.Sampling_MPC(config=make_config("aliengo", control_parametrization=param, sampling_method="cem_mppi"))
    assert c.num_control_parameters == P and c.best_control_parameters.shape == (P,)
    assert c.sigma_cem_mppi.shape == (P,) and np.all(c.sigma_cem_mppi == 3)
    assert c.num_sampling_iterations == 1 and c.sampling_method == "cem_mppi"
    assert c.robot.mass == pytest.approx(24.637)
    assert c.jitted_compute_control == c.compute_control_cem_mppi
    k0 = c.get_key().copy()
    assert c.with_newkey() is c and not np.array_equal(c.get_key(), k0)
    assert c.with_newsigma(3) is c and c.get_sigma() == 3


def test_unknown_sampling_method_exits(host_only):
    with pytest.raises(SystemExit):  # centroidal_nmpc_jax.py:107-110
        fixed.Sampling_MPC(config=make_config("go2", sampling_method="simulated_annealing"))


def _dicts(rng):
    state = {"position": rng.normal(size=3), "linear_velocity": rng.normal(size=3), "orientation": rng.normal(size=3),
             "angular_velocity": rng.normal(size=3)}
    ref = {"ref_position": rng.normal(size=3), "ref_linear_velocity": rng.normal(size=3),
           "ref_orientation": rng.normal(size=3), "ref_angular_velocity": rng.normal(size=3)}
    for f in ("FL", "FR", "RL", "RR"):
        state["foot_" + f] = rng.normal(size=3)
        ref["ref_foot_" + f] = rng.normal(size=(1, 3))
    return state, ref


def test_prepare_state_and_reference(host_only):
    rng = np.random.default_rng(0)
    c = fixed.Sampling_MPC(config=make_config("go2", sampling_method="mppi", control_parametrization="linear_spline"))
    c.best_control_parameters = np.arange(36, dtype=np.float32) + 1
    state, ref = _dicts(rng)
    cur, prev = np.array([1.0, 0.0, 0.0, 1.0]), np.array([1.0, 1.0, 0.0, 1.0])
    s, r = c.prepare_state_and_reference(state, ref, cur, prev)
    assert s.shape == (24,) and r.shape == (24,) and s.dtype == np.float64
    assert np.array_equal(s[0:3], state["position"]) and np.array_equal(s[9:12], state["angular_velocity"])
    assert np.array_equal(s[12:15], state["foot_FL"])                      # stance: measured foot
    assert np.array_equal(s[15:18], ref["ref_foot_FR"].reshape(3))         # swing: reference foothold (:588-595)
    assert np.array_equal(s[18:21], ref["ref_foot_RL"].reshape(3))
    assert np.array_equal(r[12:24], np.concatenate([ref["ref_foot_" + f].reshape(3) for f in ("FL", "FR", "RL", "RR")]))
    b = c.best_control_parameters
    assert np.all(b[9:18] == 0) and np.all(b[0:9] != 0) and np.all(b[18:] != 0)  # only the lift-off leg (FR) is zeroed


@pytest.mark.parametrize("param", ["zero_order", "linear_spline", "cubic_spline"])
def test_host_splines_match_oracle_rollout_convention(host_only, param):
    """The host spline functions (used by shift_solution) agree with the formulas the oracle rolls out: evaluate the
    first-step force the way compute_control does (step 0, horizon_leg 1) and compare with the oracle's GRF tail."""
    rng = np.random.default_rng(3)
    cfgo = make_config("aliengo", control_parametrization=param, sampling_method="mppi")
    c = fixed.Sampling_MPC(config=cfgo)
    P = c.num_control_parameters
    best = rng.normal(0, 2, P).astype(np.float32)
    n = c.num_control_parameters_single_leg
    ocfg = so.make_cfg(mass=cfgo.mass, inertia=cfgo.inertia, parametrization=param)
    st = np.zeros(24); st[2] = 0.33
    st[12:24] = np.array([[0.24, 0.15, 0], [0.24, -0.15, 0], [-0.24, 0.15, 0], [-0.24, -0.15, 0]]).ravel()
    seq = np.ones((4, 12))
    o = so.solve(ocfg, "mppi", st, st.copy(), seq, best, np.zeros((4, P), np.float32))  # zero noise: best unchanged
    mg4 = np.float32(np.float32(cfgo.mass * 9.81) / np.float32(4))
    for leg in range(4):
        fx, fy, fz = c.spline_fun_FL(best[leg * n:(leg + 1) * n], 0.0, 1)
        fz = min(max(np.float32(fz + mg4), 0), np.float32(cfgo.mass * 9.81))
        lim = np.float32(0.5) * fz
        exp = [min(max(np.float32(fx / 3), -lim), lim), min(max(np.float32(fy / 3), -lim), lim), fz]
        assert np.allclose(o["grf"][3 * leg:3 * leg + 3], exp, rtol=2e-6, atol=1e-5)


def test_shift_solution_touches_entries_0_2_4(host_only):
    c = fixed.Sampling_MPC(config=make_config("go2", control_parametrization="linear_spline"))
    best = np.arange(36, dtype=np.float32)
    out = c.shift_solution(best, 0.01)
    changed = np.nonzero(out != best)[0]
    assert set(changed) <= {leg * 9 + k for leg in range(4) for k in (0, 2, 4)}
    assert np.array_equal(best, np.arange(36, dtype=np.float32))  # input untouched


def test_gait_adaptive_variant_signature(host_only):
    c = adaptive.Sampling_MPC(config=make_config("go2", sampling_method="cem_mppi", optimize_step_freq=True))
    assert c._gait_adaptive and c._handle.cfg.gait_adaptive == 1 and c._handle.cfg.cem_elite == 200
    with pytest.raises(TypeError):  # the interface's 6-argument cem call is broken upstream as well
        c.jitted_compute_control(np.zeros(24), np.zeros(24), np.ones((4, 12)), c.best_control_parameters, c.master_key, 3)


class _Map:
    def __init__(self, data):
        self.data = data

    def get_height(self, p):
        return 0.125


def test_vfa_state_machine_and_height_strategy():
    legs = ("FL", "FR", "RL", "RR")
    v = VisualFootholdAdaptation(legs, adaptation_strategy="height")
    ref = LegsAttr(*[np.array([0.1 * k, 0.0, 0.0]) for k in range(4)])
    got, cons = v.get_footholds_adapted(ref)
    assert got is ref and not v.initialized
    none_maps = LegsAttr(*[_Map(None) for _ in legs])
    assert v.compute_adaptation(legs, ref, ref, none_maps, None, None, None) is False and not v.initialized
    maps = LegsAttr(*[_Map(np.zeros((3, 3, 1, 3))) for _ in legs])
    assert v.compute_adaptation(legs, ref, ref, maps, None, None, None, phase_signal=np.zeros(4)) is True
    assert v.initialized and all(ref[l][2] == 0.125 for l in legs)
    assert v.get_footholds_adapted(None)[0] is ref
    v.reset()
    assert not v.initialized
    with pytest.raises(ImportError):
        VisualFootholdAdaptation(legs, adaptation_strategy="vfa")


class _RecordingHandle(_NoDeviceHandle):
    """Records the arguments of the chained solve and returns a recognisable result."""

    def solve(self, state, reference, contact_sequence, best_in, seed, offset, timing=None, nominal_freq=0.0,
              optimize_swing=0, sigma_in=None, d_eps=None, d_freqs=None, want_costs=True, iterations=1):
        self.call = dict(state=np.array(state), best_in=np.array(best_in), seed=seed, offset=offset, timing=timing,
                         nominal_freq=nominal_freq, optimize_swing=optimize_swing, sigma_in=sigma_in,
                         iterations=iterations)
        return dict(grf=np.arange(12, dtype=np.float32), footholds=np.zeros(12, np.float32),
                    predicted_state=np.full(24, 7.0, np.float32), best=np.full(self.P, 2.0, np.float32),
                    best_cost=np.float32(5.0), best_freq=np.float32(2.0), costs=np.ones(self.N, np.float32),
                    sigma=None if sigma_in is None else np.full(self.P, 0.5, np.float32), best_index=3)


@pytest.mark.parametrize("method,ga", [("mppi", True), ("cem_mppi", False)])
def test_interface_sampling_branch_plumbing(monkeypatch, method, ga):
    """SRBDControllerInterface (sampling branch, srbd_controller_interface.py:117-183, 224-243): ONE chained call per
    control step, key advanced by num_sampling_iterations, lift-off zeroing, contact-masked GRFs, reference footholds."""
    from quadruped_pympc_tamols_b200.interfaces.srbd_controller_interface import SRBDControllerInterface
    monkeypatch.setattr(_cabi, "Handle", _RecordingHandle)
    cfg = make_config("go2", sampling_method=method, optimize_step_freq=ga, num_sampling_iterations=4,
                      num_parallel_computations=64)
    itf = SRBDControllerInterface(config=cfg)
    c = itf.controller
    assert type(c).__module__.endswith("gait_adaptive") == ga
    c.best_control_parameters = np.ones(c.num_control_parameters, np.float32)
    state = {"position": np.zeros(3), "linear_velocity": np.zeros(3), "orientation": np.zeros(3),
             "angular_velocity": np.zeros(3)}
    ref = {"ref_position": np.zeros(3), "ref_linear_velocity": np.zeros(3), "ref_orientation": np.zeros(3),
           "ref_angular_velocity": np.zeros(3)}
    for k, f in enumerate(("FL", "FR", "RL", "RR")):
        state["foot_" + f] = np.full(3, float(k))
        ref["ref_foot_" + f] = np.full((1, 3), 10.0 + k)
    seq = np.ones((4, 12))
    seq[2, 0] = 0  # RL in swing now, was in stance (previous_contact_mpc starts at [1,1,1,1]): its block is zeroed
    out = itf.compute_control(state, ref, seq, cfg.inertia, np.array([0.1, 0.6, 0.6, 0.1]), 1.4, 1)
    call = c._handle.call
    n = c.num_control_parameters_single_leg
    assert call["iterations"] == 4 and call["seed"] == 42 and call["offset"] == 1
    assert np.array_equal(c.master_key, np.array([42, 4], dtype=np.uint64))
    assert np.all(call["best_in"][2 * n:3 * n] == 0) and np.all(call["best_in"][:2 * n] == 1)
    assert np.array_equal(call["state"][18:21], ref["ref_foot_RL"][0])  # swing foot takes its reference foothold
    if method == "cem_mppi":
        assert np.all(call["sigma_in"] == cfg.mpc_params["sigma_cem_mppi"]) and call["timing"] is None
        assert np.all(c.sigma_cem_mppi == 0.5)
    else:
        assert np.allclose(call["timing"], [0.1, 0.6, 0.6, 0.1]) and call["nominal_freq"] == 1.4
        assert call["optimize_swing"] == 1 and call["sigma_in"] is None
    grfs, footholds, jp, jv, ja, freq, pred = out
    assert np.array_equal(grfs.FL, [0, 1, 2]) and np.array_equal(grfs.RL, [0, 0, 0]) and np.array_equal(grfs.RR, [9, 10, 11])
    assert np.array_equal(footholds.FR, ref["ref_foot_FR"][0]) and jp is None and jv is None and ja is None
    assert freq == 2.0 and np.all(pred == 7.0) and np.all(c.best_control_parameters == 2.0)
    assert np.array_equal(itf.previous_contact_mpc, [1, 1, 0, 1])
    # second control step: the key keeps advancing, nothing is zeroed (no new lift-off)
    itf.compute_control(state, ref, seq, cfg.inertia, np.array([0.1, 0.6, 0.6, 0.1]), 1.4, 1)
    assert c._handle.call["offset"] == 5 and np.all(c._handle.call["best_in"] == 2.0)


def test_interface_rejects_non_sampling_types():
    from quadruped_pympc_tamols_b200.interfaces.srbd_controller_interface import SRBDControllerInterface
    with pytest.raises(NotImplementedError, match="sampling"):
        SRBDControllerInterface(config=make_config("go2", type="nominal"))
