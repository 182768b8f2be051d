"""Shared helpers for the parity tests: fixture loading and config construction for oracle and CUDA path."""
import glob
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")


def sampling_fixtures():
    return sorted(glob.glob(os.path.join(GOLDEN, "sampling_*.npz")))


def fixture_kwargs(g):
    """config kwargs shared by oracle.srbd_oracle.make_cfg and _cabi.make_config"""
    return dict(mass=float(g["mass"]), inertia=np.asarray(g["inertia"]), horizon=int(g["horizon"]), dt=float(g["dt"]),
                parametrization=str(g["parametrization"]), mu=float(g["mu"]), gait_adaptive=bool(int(g["gait_adaptive"])),
                use_nonuniform_discretization=bool(int(g["use_nonuniform_discretization"])),
                horizon_fine_grained=int(g["horizon_fine_grained"]), dt_fine_grained=float(g["dt_fine_grained"]))


def rel_err(a, b, floor=1e-6):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), floor))) if a.size else 0.0


def tamols_fixtures():
    return sorted(glob.glob(os.path.join(GOLDEN, "tamols_*.npz")))


def load_tamols_fixture(path):
    """-> (TamolsParams, heights, xy, jobs, reference outputs dict)"""
    from quadruped_pympc_tamols_b200 import _tamols_cabi as tc
    g = np.load(path)
    p = tc.TamolsParams.from_buffer_copy(g["params"].tobytes())
    return p, g["heights"], g["xy"], g["jobs"].astype(tc.JOB_DTYPE), g


def check_tamols_against_reference(res, scores, g, score_tol=1e-9, tie_tol=1e-6):
    """Parity bar of the TAMOLS path (BASELINE.json north_star): per-candidate scores equal to the reference's within
    float64 round-off, identical finite/inf pattern, and the selected cell index bit-exact except for documented cost
    ties under 1e-6 (the chosen cell's reference score is then within 1e-6 of the reference minimum)."""
    ref_scores, ref_idx = g["ref_scores"], g["ref_best_index"]
    n = ref_scores.shape[1]
    fin = np.isfinite(ref_scores)
    assert np.array_equal(fin, np.isfinite(scores[:, :n])), "finite / rejected pattern differs from the reference"
    if fin.any():
        assert np.max(np.abs(scores[:, :n][fin] - ref_scores[fin])) < score_tol
    for l in range(len(ref_idx)):
        bi = int(res["best_index"][l])
        if bi != int(ref_idx[l]):
            assert bi >= 0 and ref_idx[l] >= 0, (l, bi, ref_idx[l])
            assert abs(ref_scores[l, bi] - ref_scores[l, ref_idx[l]]) < tie_tol, (l, bi, ref_idx[l])
        if ref_idx[l] >= 0 and bi == int(ref_idx[l]):
            assert np.array_equal(res["best_xyz"][l], g["ref_best_xyz"][l])  # same cell -> same float64 coordinates
            assert np.max(np.abs(res["vertex1"][l] - g["ref_vertex1"][l])) == 0.0
            assert np.max(np.abs(res["vertex2"][l] - g["ref_vertex2"][l])) == 0.0
        assert int(res["num_finite"][l]) == int(fin[l].sum())


def softmin_tolerance(costs_gpu, costs_oracle, eps, base=1e-3):
    """Tolerance on the MPPI / CEM-MPPI update against the ORACLE's update, accounting for the conditioning of the
    softmin at temperature 1 (centroidal_nmpc_jax.py:828-836).  A cost perturbation of delta on the samples that carry
    weight changes each weight by at most a factor exp(+-2 delta), hence the update by at most
    (exp(2 delta) - 1) * sum_i w_i |eps_i|.  Costs are 1e3..1e5, so one float32 ulp is 5e-4..8e-3: when two rollouts
    are within ~1 cost unit of the minimum, round-off alone (both sides are float32; neither is exact) moves the update
    by more than the plain 1e-3 bar of BASELINE.json, which therefore applies as is only while delta <~ 1e-4.  The
    arithmetic of the update itself is pinned separately and tightly by `softmin_update_f64`."""
    c = np.asarray(costs_oracle, np.float64)
    w = np.exp(-(c - c.min()))
    w /= w.sum()
    live = w > 1e-7
    delta = float(np.max(np.abs(np.asarray(costs_gpu, np.float64)[live] - c[live])))
    spread = float(np.max(w @ np.abs(np.asarray(eps, np.float64))))
    return base + (np.exp(2 * delta) - 1) * spread


PARITY_LOG = []  # rows printed by conftest.pytest_terminal_summary at the end of a -m gpu run


def record_parity(tag, method, r, o, eps):
    """Observed deviation of one CUDA solve from the oracle on the same noise, for the terminal summary: which configs
    meet the plain north-star bars (costs 1e-4 relative, update 1e-3), and where the conditioning-aware tolerance of
    softmin_tolerance was actually needed (delta = largest cost difference on the samples that carry weight)."""
    row = dict(tag=tag, cost_rel=rel_err(r["costs"], o["costs"]), best_abs=float(np.max(np.abs(r["best"] - o["best"]))))
    if method != "random_sampling":
        c = np.asarray(o["costs"], np.float64)
        w = np.exp(-(c - c.min()))
        w /= w.sum()
        live = w > 1e-7
        row.update(tol=float(softmin_tolerance(r["costs"], o["costs"], eps)), live=int(live.sum()),
                   delta=float(np.max(np.abs(np.asarray(r["costs"], np.float64)[live] - c[live]))))
    else:
        row.update(tol=1e-3, live=None, delta=None)
    PARITY_LOG.append(row)


def oracle_tail(so, ocfg, state, ref, contact_seq, best, timing=None):
    """First-step GRFs and one-step prediction of the ORACLE evaluated at a given mean `best`
    (centroidal_nmpc_jax.py:707-784): one solve with a single zero-noise sample.  Pins the tail of the CUDA solve
    tightly, independent of the conditioning of the update that produced `best`."""
    best = np.asarray(best, np.float32)
    o = so.solve(ocfg, "random_sampling", state, ref, contact_seq, best, np.zeros((1, best.size), np.float32),
                 timing=timing, freqs=np.full(1, 1.4, np.float32))
    return o["grf"], o["predicted_state"]


def softmin_update_f64(best_in, costs, eps):
    """best + sum_i softmin(costs)_i eps_i evaluated in float64 from the given (already saturated) costs."""
    c = np.asarray(costs, np.float64)
    w = np.exp(-(c - c.min()))
    w /= w.sum()
    return np.asarray(best_in, np.float64) + w @ np.asarray(eps, np.float64)
