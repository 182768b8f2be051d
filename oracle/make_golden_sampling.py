"""Generate tests/golden/sampling_*.npz by running the reference's OWN sampling controller source
(/root/reference/quadruped_pympc/controllers/sampling/*.py, unmodified) under the numpy-backed jax
stand-in of oracle/jax_numpy_standin.py.

TEST INFRASTRUCTURE.  Run in the build container only (needs /root/reference):
    python oracle/make_golden_sampling.py
Each fixture stores the inputs, the explicit noise matrix `eps` (= additional_random_parameters) and
per-sample step frequencies the reference drew, and every output of compute_control_*.
"""
import importlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
REF = os.environ.get("QPMPC_REFERENCE", "/root/reference")
sys.path = [p for p in sys.path if os.path.abspath(p or ".") != REPO]  # the repo's drop-in shim must not shadow
sys.path.insert(0, REF)
sys.path.insert(0, HERE)

import jax_numpy_standin as standin  # noqa: E402

standin.install()
import quadruped_pympc.config as config  # noqa: E402  (the reference's config)

ROBOTS = {
    "aliengo": (24.637, [[0.2310941359705289, -0.0014987128245817424, -0.021400468992761768],
                         [-0.0014987128245817424, 1.4485084687476608, 0.0004641447134275615],
                         [-0.021400468992761768, 0.0004641447134275615, 1.503217877350808]], 0.33, (0.24, 0.15)),
    "go2": (15.019, [[1.58460467e-01, 1.21660000e-04, -1.55444692e-02],
                     [1.21660000e-04, 4.68645637e-01, -3.12000000e-05],
                     [-1.55444692e-02, -3.12000000e-05, 5.24474661e-01]], 0.27, (0.19, 0.13)),
    "hyqreal1": (108.40, [[4.55031444e+00, 2.75249434e-03, -5.11957307e-01],
                          [2.75249434e-03, 2.02411774e+01, -7.38560592e-04],
                          [-5.11957307e-01, -7.38560592e-04, 2.14269772e+01]], 0.55, (0.44, 0.30)),
}
GAITS = {"trot": ([0.5, 1.0, 1.0, 0.5], 1.4, 0.65), "pace": ([0.8, 0.3, 0.8, 0.3], 1.4, 0.7),
         "bound": ([0.5, 0.5, 0.0, 0.0], 1.8, 0.65)}


def contact_sequence(gait, phase0, horizon, dt):
    """Host gait clock -> contact_sequence[4,H] (format of helpers/periodic_gait_generator.py:104-118):
    leg in stance while its phase < duty factor."""
    offs, freq, duty = GAITS[gait]
    seq = np.zeros((4, horizon))
    t = (np.array(offs) + phase0) % 1.0
    for n in range(horizon):
        seq[:, n] = (t < duty).astype(float)
        t = (t + dt * freq) % 1.0
    return seq, (np.array(offs) + phase0) % 1.0, freq


def synth_state(robot, rng):
    """SURVEY.md 8(d) C1 state/reference distribution."""
    _, _, h, (hx, hy) = ROBOTS[robot]
    st = np.zeros(24)
    st[0:3] = [0.0, 0.0, h + rng.normal(0, 0.02)]
    st[3:6] = rng.normal(0, 0.3, 3)
    st[6:9] = rng.normal(0, 0.05, 3)
    st[9:12] = rng.normal(0, 0.3, 3)
    hips = np.array([[hx, hy, 0], [hx, -hy, 0], [-hx, hy, 0], [-hx, -hy, 0]], float)
    st[12:24] = (hips + np.c_[rng.normal(0, 0.03, (4, 2)), np.zeros(4)]).ravel()
    ref = np.zeros(24)
    ref[2] = h
    ref[3] = rng.uniform(-0.5, 0.5)
    ref[12:24] = st[12:24]
    return st, ref


def make_controller(robot, method, param, gait_adaptive, n, **over):
    mass, inertia, _, _ = ROBOTS[robot]
    config.robot, config.mass, config.inertia = robot, mass, np.array(inertia)
    mp = config.mpc_params
    mp.update(dict(type="sampling", sampling_method=method, control_parametrization=param,
                   num_parallel_computations=n, grf_max=mass * 9.81, grf_min=0, device="cpu",
                   optimize_step_freq=gait_adaptive, use_nonuniform_discretization=False, horizon=12, dt=0.02,
                   num_splines=2, mu=0.5))
    mp.update(over)
    name = "quadruped_pympc.controllers.sampling.centroidal_nmpc_jax" + ("_gait_adaptive" if gait_adaptive else "")
    for m in (name, "quadruped_pympc.controllers.sampling.centroidal_model_jax"):
        sys.modules.pop(m, None)
    mod = importlib.import_module(name)
    return mod.Sampling_MPC()


def eps_from_draws(method, n, p, draws, sigma):
    """Rebuild additional_random_parameters exactly as centroidal_nmpc_jax.py:647-677 / 806-812 / 951-958."""
    eps = np.zeros((n, p), np.float32)
    normals = [a for k, a in draws if k == "normal"]
    uniforms = [a for k, a in draws if k == "uniform"]
    if method == "random_sampling":
        t = int(n / 3)
        s = config.mpc_params["sigma_random_sampling"]
        eps[1:1 + t] = np.float32(s[0]) * normals[0]
        eps[1 + t:1 + 2 * t] = np.float32(s[1]) * normals[1]
        eps[1 + 2 * t:n] = uniforms[0]
    elif method == "mppi":
        eps[1:] = np.float32(config.mpc_params["sigma_mppi"]) * normals[0]
    else:
        eps[1:] = normals[0] * np.asarray(sigma, np.float32)
    return eps


def run_case(name, robot, method, param, gait_adaptive, n, gait, seed, optimize_swing=1, contact_override=None,
             iterations=1, best_scale=0.0, **over):
    rng = np.random.default_rng(seed)
    ctrl = make_controller(robot, method, param, gait_adaptive, n, **over)
    p = ctrl.num_control_parameters
    st, ref = synth_state(robot, rng)
    seq, timing, nominal = contact_sequence(gait, rng.uniform(), 12, 0.02)
    if contact_override is not None:
        seq = np.array(contact_override, float)
    best = (best_scale * rng.normal(size=p)).astype(np.float32)
    sigma = config.mpc_params["sigma_cem_mppi"]
    out = dict(robot=robot, method=method, parametrization=param, gait_adaptive=int(gait_adaptive), n=n,
               mass=config.mass, inertia=np.array(config.inertia), horizon=12, dt=0.02, mu=0.5,
               use_nonuniform_discretization=int(config.mpc_params["use_nonuniform_discretization"]),
               horizon_fine_grained=config.mpc_params["horizon_fine_grained"],
               dt_fine_grained=config.mpc_params["dt_fine_grained"],
               step_freq_available=np.array(config.mpc_params["step_freq_available"]),
               state=st, reference=ref, contact_sequence=seq, timing=timing, nominal_freq=nominal,
               optimize_swing=optimize_swing, iterations=iterations, best_in=best)
    for it in range(iterations):
        ctrl.with_newkey()
        standin.DRAWS.clear()
        if method == "cem_mppi":
            args = (st, ref, seq, best, ctrl.master_key, sigma)
            if gait_adaptive:  # direct call: the interface's 6-arg call is broken upstream for this variant
                args += (timing, nominal, optimize_swing)
            res = ctrl.jitted_compute_control(*args)
        else:
            res = ctrl.jitted_compute_control(st, ref, seq, best, ctrl.master_key, timing, nominal, optimize_swing)
        draws = list(standin.DRAWS)
        eps = eps_from_draws(method, n, p, draws, sigma)
        freqs = [a for k, a in draws if k == "choice"]
        tag = f"it{it}_"
        out[tag + "eps"] = eps
        if gait_adaptive:
            f = np.asarray(freqs[0], np.float32)
            if method == "cem_mppi":
                f = (f * np.float32(optimize_swing) + np.float32(nominal)).astype(np.float32)
            out[tag + "freqs"] = f
        out[tag + "sigma_in"] = np.broadcast_to(np.asarray(sigma, np.float32), (p,)).copy()
        keys = ("grf", "footholds", "predicted_state", "best_out", "best_cost", "best_freq", "costs", "sigma_out")
        if method == "cem_mppi" and gait_adaptive:  # ..._gait_adaptive.py:1123-1131 returns no best_freq
            keys = ("grf", "footholds", "predicted_state", "best_out", "best_cost", "costs", "sigma_out")
        for k, v in zip(keys, res):
            out[tag + k] = np.asarray(v, np.float32)
        best = out[tag + "best_out"]
        if method == "cem_mppi":
            sigma = out[tag + "sigma_out"]
    path = os.path.join(REPO, "tests", "golden", f"sampling_{name}.npz")
    np.savez_compressed(path, **out)
    print(f"{name:34s} N={n:4d} P={p:3d} best_cost={float(out[f'it{iterations-1}_best_cost']):.4f}")


FLIGHT = [[1, 1, 1, 0, 0, 0, 1, 1, 1, 1, 1, 1], [1, 1, 0, 0, 0, 1, 1, 1, 1, 1, 1, 1],
          [1, 1, 1, 0, 0, 1, 1, 1, 1, 1, 1, 1], [1, 0, 0, 0, 0, 0, 1, 1, 1, 1, 1, 1]]

CASES = [
    dict(name="c1_aliengo_random_zero", robot="aliengo", method="random_sampling", param="zero_order", gait_adaptive=False, n=96, gait="trot", seed=1),
    dict(name="aliengo_mppi_zero", robot="aliengo", method="mppi", param="zero_order", gait_adaptive=False, n=96, gait="trot", seed=2, best_scale=2.0),
    dict(name="aliengo_mppi_linear_nonuniform", robot="aliengo", method="mppi", param="linear_spline", gait_adaptive=False, n=96, gait="trot", seed=3, best_scale=2.0, use_nonuniform_discretization=True),
    dict(name="aliengo_random_cubic_flight", robot="aliengo", method="random_sampling", param="cubic_spline", gait_adaptive=False, n=96, gait="trot", seed=4, contact_override=FLIGHT, best_scale=1.0),
    dict(name="c2_go2_mppi_cubic_ga", robot="go2", method="mppi", param="cubic_spline", gait_adaptive=True, n=96, gait="bound", seed=5, best_scale=2.0),
    dict(name="go2_random_linear_ga_swing0", robot="go2", method="random_sampling", param="linear_spline", gait_adaptive=True, n=96, gait="trot", seed=6, optimize_swing=0, best_scale=1.0),
    dict(name="go2_mppi_zero_ga", robot="go2", method="mppi", param="zero_order", gait_adaptive=True, n=96, gait="trot", seed=7, best_scale=2.0),
    dict(name="c3_hyqreal_cem_cubic_2it", robot="hyqreal1", method="cem_mppi", param="cubic_spline", gait_adaptive=False, n=96, gait="pace", seed=8, iterations=2, best_scale=2.0),
    dict(name="hyqreal_cem_zero", robot="hyqreal1", method="cem_mppi", param="zero_order", gait_adaptive=False, n=64, gait="pace", seed=9),
    dict(name="go2_cem_cubic_ga", robot="go2", method="cem_mppi", param="cubic_spline", gait_adaptive=True, n=256, gait="trot", seed=10, best_scale=1.0),
]

if __name__ == "__main__":
    only = sys.argv[1:]
    for c in CASES:
        if not only or c["name"] in only:
            run_case(**c)
