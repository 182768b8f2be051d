"""CPU: the C oracle (oracle/srbd_oracle.c) against the fixtures produced by executing the reference's own
sampling-controller source under the numpy jax stand-in (oracle/make_golden_sampling.py)."""
import os

import numpy as np
import pytest

from oracle import srbd_oracle as so
from tests.util import fixture_kwargs, rel_err, sampling_fixtures


@pytest.mark.parametrize("path", sampling_fixtures(), ids=lambda p: os.path.basename(p)[9:-4])
def test_oracle_matches_reference_fixture(path):
    g = np.load(path)
    cfg = so.make_cfg(**fixture_kwargs(g))
    best = g["best_in"]
    for it in range(int(g["iterations"])):
        t = f"it{it}_"
        r = so.solve(cfg, str(g["method"]), g["state"], g["reference"], g["contact_sequence"], best, g[t + "eps"],
                     timing=g["timing"], freqs=g[t + "freqs"] if t + "freqs" in g else None)
        assert rel_err(r["costs"], g[t + "costs"]) < 1e-5
        assert np.max(np.abs(r["best"] - g[t + "best_out"])) < 1e-3
        assert np.max(np.abs(r["grf"] - g[t + "grf"])) < 1e-3
        assert rel_err(r["predicted_state"], g[t + "predicted_state"], 1e-3) < 1e-5
        assert abs(float(r["best_cost"]) - float(g[t + "best_cost"])) <= 1e-5 * abs(float(g[t + "best_cost"]))
        if t + "best_freq" in g:
            assert abs(float(r["best_freq"]) - float(g[t + "best_freq"])) < 1e-6
        if str(g["method"]) == "cem_mppi":
            assert rel_err(r["sigma"], g[t + "sigma_out"]) < 1e-5
        best = g[t + "best_out"]


def test_known_answer_all_stance_rest():
    """level body at rest, symmetric feet, zero parameters -> fz = mg/4, state stays put."""
    mass = 24.637
    cfg = so.make_cfg(mass=mass, inertia=np.diag([0.2, 1.4, 1.5]), parametrization="zero_order")
    st = np.zeros(24)
    st[2] = 0.33
    st[12:24] = np.array([[0.24, 0.15, 0], [0.24, -0.15, 0], [-0.24, 0.15, 0], [-0.24, -0.15, 0]]).ravel()
    ref = st.copy()
    eps = np.zeros((4, 144), np.float32)
    r = so.solve(cfg, "mppi", st, ref, np.ones((4, 12)), np.zeros(144), eps)
    assert np.allclose(r["grf"].reshape(4, 3)[:, 2], mass * 9.81 / 4, rtol=1e-6)
    assert np.allclose(r["grf"].reshape(4, 3)[:, :2], 0)
    assert np.all(r["costs"] < 1e-6)
    assert np.allclose(r["predicted_state"], st, atol=1e-6)


def test_known_answer_flight_phase_free_fall():
    """zero legs in stance -> all forces 0 (NaN path), v_z -= 9.81 dt each step."""
    cfg = so.make_cfg(mass=15.0, inertia=np.diag([0.1, 0.4, 0.5]), parametrization="linear_spline")
    st = np.zeros(24)
    st[2] = 0.3
    rng = np.random.default_rng(0)
    eps = rng.normal(size=(8, 36)).astype(np.float32)
    r = so.solve(cfg, "random_sampling", st, st.copy(), np.zeros((4, 12)), np.zeros(36), eps)
    assert np.all(r["grf"] == 0)
    assert abs(r["predicted_state"][5] + 9.81 * 0.02) < 1e-6
    assert np.allclose(r["costs"], r["costs"][0])  # parameters cannot matter in flight
