"""Where the drop-in classes get their configuration.

The reference classes read the module `quadruped_pympc.config` at construction
(controllers/sampling/centroidal_nmpc_jax.py:15,27-33; helpers/visual_foothold_adaptation.py:4,56-57).  The drop-ins do
the same when that module is importable; otherwise (or when a caller wants several differently configured controllers
in one process, as the batched bench does) a config object with the same attribute names is passed explicitly.
`make_config()` builds such an object with config.py's defaults for the keys this path reads.
"""
import types

import numpy as np

_I_SMALL = np.array([[1.58460467e-01, 1.21660000e-04, -1.55444692e-02], [1.21660000e-04, 4.68645637e-01, -3.12000000e-05],
                     [-1.55444692e-02, -3.12000000e-05, 5.24474661e-01]])
_I_MID = np.array([[0.2310941359705289, -0.0014987128245817424, -0.021400468992761768],
                   [-0.0014987128245817424, 1.4485084687476608, 0.0004641447134275615],
                   [-0.021400468992761768, 0.0004641447134275615, 1.503217877350808]])
_I_HYQ = np.array([[4.55031444e+00, 2.75249434e-03, -5.11957307e-01], [2.75249434e-03, 2.02411774e+01, -7.38560592e-04],
                   [-5.11957307e-01, -7.38560592e-04, 2.14269772e+01]])
# config.py:20-66 (mass, inertia) per robot
ROBOTS = {"go1": (12.019, _I_SMALL), "go2": (15.019, _I_SMALL), "aliengo": (24.637, _I_MID), "b2": (83.49, _I_MID),
          "hyqreal1": (108.40, _I_HYQ), "hyqreal2": (126.69, _I_HYQ), "mini_cheetah": (12.5, _I_SMALL),
          "spot": (50.34, _I_MID)}

# config.py:209-243
TAMOLS_DEFAULTS = {
    "search_radius": 0.32, "search_resolution": 0.04, "patch_size": 3, "gradient_delta": 0.04,
    "weight_edge_avoidance": 10.0, "weight_roughness": 10, "weight_deviation": 2, "weight_kinematic": 2.0,
    "weight_nominal_kinematic": 0.0, "weight_reference_tracking": 10.0, "weight_stability": 20.0,
    "stability_margin": 0.06, "stability_hard": False, "stability_soft": True, "estimated_swing_time": 0.25,
    "l_min": {"go1": 0.15, "go2": 0.15, "aliengo": 0.1, "b2": 0.25, "hyqreal1": 0.25, "hyqreal2": 0.25,
              "mini_cheetah": 0.12, "spot": 0.20},
    "l_max": {"go1": 0.45, "go2": 0.45, "aliengo": 0.55, "b2": 0.75, "hyqreal1": 0.75, "hyqreal2": 0.75,
              "mini_cheetah": 0.40, "spot": 0.60},
    "slope_threshold": 0.7, "constraint_box_dx": 0.05, "constraint_box_dy": 0.05,
}


def resolve_config(config=None):
    """`config` if given, else the reference's own `quadruped_pympc.config` module."""
    if config is not None:
        return config
    try:
        import quadruped_pympc.config as cfg  # the reference package, when the user has it installed
        return cfg
    except Exception as e:  # noqa: BLE001
        raise ImportError("quadruped_pympc.config is not importable: pass config=make_config(...) "
                          "(quadruped_pympc_tamols_b200._config) to the constructor") from e


def make_config(robot="aliengo", hip_height=None, tamols_params=None, **mpc_overrides):
    """A stand-in for the `quadruped_pympc.config` module: attributes robot, mass, inertia, mpc_params,
    simulation_params with config.py's defaults for every key the sampling / TAMOLS path reads (config.py:70-243)."""
    mass, inertia = ROBOTS[robot]
    mpc = {
        "type": "sampling", "horizon": 12, "dt": 0.02, "grf_max": mass * 9.81, "grf_min": 0, "mu": 0.5,
        "use_nonuniform_discretization": False, "horizon_fine_grained": 2, "dt_fine_grained": 0.01,
        "optimize_step_freq": False, "step_freq_available": [1.4, 2.0, 2.4],
        "sampling_method": "random_sampling", "control_parametrization": "cubic_spline", "num_splines": 2,
        "num_parallel_computations": 10000, "num_sampling_iterations": 1, "device": "gpu",
        "sigma_cem_mppi": 3, "sigma_mppi": 3, "sigma_random_sampling": [0.2, 3, 10], "shift_solution": False,
    }
    unknown = set(mpc_overrides) - set(mpc)
    if unknown:
        raise KeyError(f"unknown mpc_params keys: {sorted(unknown)}")
    mpc.update(mpc_overrides)
    tp = dict(TAMOLS_DEFAULTS)
    tp["h_des"] = hip_height if hip_height is not None else 0.35  # config.py:226 uses the robot's hip_height
    tp.update(tamols_params or {})
    sim = {"visual_foothold_adaptation": "tamols", "tamols_params": tp}
    return types.SimpleNamespace(robot=robot, mass=mass, inertia=inertia.copy(), gravity_constant=9.81,
                                 mpc_params=mpc, simulation_params=sim)
