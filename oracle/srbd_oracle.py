"""ctypes front-end of oracle/srbd_oracle.c (TEST INFRASTRUCTURE, not product code).

The C file restates /root/reference/quadruped_pympc/controllers/sampling/centroidal_nmpc_jax*.py and
centroidal_model_jax.py in float32; see its header for the line-by-line map.  This module only
marshals numpy arrays.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_MAX_H = 64
PARAM = {"zero_order": 0, "linear_spline": 1, "cubic_spline": 2}
METHOD = {"random_sampling": 0, "mppi": 1, "cem_mppi": 2}


class OracleCfg(C.Structure):
    _fields_ = [
        ("horizon", C.c_int), ("parametrization", C.c_int), ("num_splines", C.c_int), ("gait_adaptive", C.c_int),
        ("mass", C.c_float), ("inertia", C.c_float * 9),
        ("mu", C.c_float), ("f_z_min", C.c_float), ("f_z_max", C.c_float),
        ("dts", C.c_float * ORACLE_MAX_H), ("q_diag", C.c_float * 24),
        ("duty_factor", C.c_float), ("mpc_dt", C.c_float),
        ("cem_elite", C.c_int), ("cem_sqrt", C.c_int), ("cem_lo", C.c_float), ("cem_hi", C.c_float),
    ]


class OracleResult(C.Structure):
    _fields_ = [("grf", C.c_float * 12), ("footholds", C.c_float * 12), ("predicted_state", C.c_float * 24),
                ("best_cost", C.c_float), ("best_freq", C.c_float), ("best_index", C.c_int)]


def build(force=False):
    """Compile the two oracle objects with oracle/Makefile (gcc only)."""
    out = os.path.join(_HERE, "_build", "libsrbd_oracle.so")
    if force or not os.path.exists(out) or os.path.getmtime(out) < os.path.getmtime(os.path.join(_HERE, "srbd_oracle.c")):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return out


_LIBS = {}


def _lib(fast=False):
    if fast not in _LIBS:
        build()
        lib = C.CDLL(os.path.join(_HERE, "_build", "libsrbd_oracle_fast.so" if fast else "libsrbd_oracle.so"))
        assert lib.oracle_sizeof_cfg() == C.sizeof(OracleCfg)
        assert lib.oracle_sizeof_result() == C.sizeof(OracleResult)
        _LIBS[fast] = lib
    return _LIBS[fast]


Q_DIAG = np.array([0, 0, 1500, 200, 200, 200, 500, 500, 0, 20, 20, 50] + [0] * 12, np.float32)


def make_cfg(*, mass, inertia, horizon=12, dt=0.02, parametrization="cubic_spline", num_splines=2, mu=0.5,
             grf_min=0.0, grf_max=None, gait_adaptive=False, use_nonuniform_discretization=False,
             horizon_fine_grained=2, dt_fine_grained=0.01, duty_factor=0.65):
    c = OracleCfg()
    c.horizon, c.parametrization, c.num_splines = horizon, PARAM[parametrization], num_splines
    c.gait_adaptive = int(gait_adaptive)
    c.mass = mass
    c.inertia[:] = [float(v) for v in np.asarray(inertia, np.float32).reshape(9)]
    c.mu, c.f_z_min = mu, grf_min
    c.f_z_max = mass * 9.81 if grf_max is None else grf_max
    dts = np.full(horizon, dt, np.float32)
    if use_nonuniform_discretization:  # centroidal_model_jax.py:42-53
        dts[:horizon_fine_grained] = dt_fine_grained
    c.dts[:horizon] = [float(v) for v in dts]
    c.q_diag[:] = [float(v) for v in Q_DIAG]
    c.duty_factor, c.mpc_dt = duty_factor, dt
    if gait_adaptive:  # centroidal_nmpc_jax_gait_adaptive.py:1115-1121
        c.cem_elite, c.cem_sqrt, c.cem_lo, c.cem_hi = 200, 0, 0.5, 3.0
    else:              # centroidal_nmpc_jax.py:1075-1081
        c.cem_elite, c.cem_sqrt, c.cem_lo, c.cem_hi = 10, 1, 0.2, 5.0
    return c


def num_params(cfg):
    return _lib().oracle_num_params(C.byref(cfg))


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _p(a):
    return a.ctypes.data_as(C.POINTER(C.c_float)) if a is not None else None


def rollout_costs(cfg, state, ref, contact_seq, params, timing=None, freqs=None, fast=False):
    params = _f32(params)
    n = params.shape[0]
    costs = np.empty(n, np.float32)
    st, rf = _f32(state), _f32(ref)
    cs = _f32(contact_seq) if contact_seq is not None else None
    tm = _f32(timing) if timing is not None else np.zeros(4, np.float32)
    fr = _f32(freqs) if freqs is not None else None
    _lib(fast).oracle_rollout_costs(C.byref(cfg), _p(st), _p(rf), _p(cs), _p(tm), _p(params), _p(fr), n, _p(costs))
    return costs


def solve(cfg, method, state, ref, contact_seq, best, eps, timing=None, freqs=None, fast=False):
    """One compute_control_* call.  Returns dict(grf, footholds, predicted_state, best, best_cost, best_freq,
    costs, sigma, best_index)."""
    eps = _f32(eps)
    n, p = eps.shape
    assert p == num_params(cfg)
    best = _f32(best).copy()
    costs = np.empty(n, np.float32)
    sigma = np.zeros(p, np.float32)
    res = OracleResult()
    st, rf, cs = _f32(state), _f32(ref), _f32(contact_seq)
    tm = _f32(timing) if timing is not None else np.zeros(4, np.float32)
    fr = _f32(freqs) if freqs is not None else None
    _lib(fast).oracle_solve(C.byref(cfg), METHOD[method], _p(st), _p(rf), _p(cs), _p(tm), _p(eps), _p(fr), n,
                            _p(best), _p(costs), _p(sigma), C.byref(res))
    return dict(grf=np.array(res.grf, np.float32), footholds=np.array(res.footholds, np.float32),
                predicted_state=np.array(res.predicted_state, np.float32), best=best,
                best_cost=np.float32(res.best_cost), best_freq=np.float32(res.best_freq), costs=costs,
                sigma=sigma if method == "cem_mppi" else None, best_index=int(res.best_index))


def set_threads(n, fast=True):
    """Use n OpenMP threads in the -O3 build (torchrun pins OMP_NUM_THREADS=1 per rank otherwise)."""
    _lib(fast).oracle_set_threads(int(n))


def max_threads(fast=True):
    return _lib(fast).oracle_max_threads()


def fill_noise(n, p, sigma, seed, fast=True):
    """Multithreaded CPU noise for bench.py's timed CPU arm (row 0 zero, sigma applied)."""
    eps = np.empty((n, p), np.float32)
    lib = _lib(fast)
    lib.oracle_fill_noise.argtypes = [C.POINTER(C.c_float), C.c_int, C.c_int, C.c_float, C.c_uint64]
    lib.oracle_fill_noise(_p(eps), n, p, float(sigma), int(seed))
    return eps
