"""ctypes binding of lib/libqpmpc_b200.so (C ABI: include/qpmpc.h).

This is the only place the Python host touches native code for the sampling path.  There is no CPU
fallback: if the shared library is missing or no sm_100 GPU is present, calls raise `QpmpcError`.
torch is used by callers only for device buffers (tensor.data_ptr()); nothing here needs it.
"""
import ctypes as C
import weakref
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("QPMPC_LIB") or os.path.join(_HERE, "lib", "libqpmpc_b200.so")
CSRC = os.path.join(_HERE, "csrc")

QPMPC_MAX_H = 32
QPMPC_MAX_FREQ = 8
PARAM = {"zero_order": 0, "linear_spline": 1, "cubic_spline": 2}
METHOD = {"random_sampling": 0, "mppi": 1, "cem_mppi": 2}


class QpmpcError(RuntimeError):
    pass


class QpmpcConfig(C.Structure):
    _fields_ = [
        ("horizon", C.c_int32), ("parametrization", C.c_int32), ("num_splines", C.c_int32),
        ("gait_adaptive", C.c_int32), ("sampling_method", C.c_int32), ("num_parallel_computations", C.c_int32),
        ("max_robots", C.c_int32), ("device", C.c_int32),
        ("mass", C.c_float), ("inertia", C.c_float * 9),
        ("mu", C.c_float), ("f_z_min", C.c_float), ("f_z_max", C.c_float),
        ("dts", C.c_float * QPMPC_MAX_H), ("q_diag", C.c_float * 12),
        ("duty_factor", C.c_float), ("mpc_dt", C.c_float),
        ("sigma_random_sampling", C.c_float * 3), ("sigma_mppi", C.c_float),
        ("num_step_freq", C.c_int32), ("step_freq_available", C.c_float * QPMPC_MAX_FREQ),
        ("cem_elite", C.c_int32), ("cem_sqrt", C.c_int32), ("cem_lo", C.c_float), ("cem_hi", C.c_float),
    ]


class QpmpcResult(C.Structure):
    _fields_ = [("grf", C.c_float * 12), ("footholds", C.c_float * 12), ("predicted_state", C.c_float * 24),
                ("best_cost", C.c_float), ("best_freq", C.c_float), ("best_index", C.c_int32), ("reserved", C.c_int32)]


RESULT_FLOATS = C.sizeof(QpmpcResult) // 4  # 52


class QpmpcBatch(C.Structure):
    _fields_ = [("num_robots", C.c_int32), ("robot_offset", C.c_int32),
                ("d_state", C.c_void_p), ("d_reference", C.c_void_p), ("d_contact_sequence", C.c_void_p),
                ("d_best_in", C.c_void_p), ("d_sigma_in", C.c_void_p), ("d_timing", C.c_void_p),
                ("d_nominal_freq", C.c_void_p), ("d_optimize_swing", C.c_void_p), ("d_eps", C.c_void_p),
                ("d_freqs", C.c_void_p), ("d_result", C.c_void_p), ("d_best_out", C.c_void_p),
                ("d_sigma_out", C.c_void_p), ("d_costs", C.c_void_p)]


MAX_H = 32  # QPMPC_MAX_H


class QpmpcGaitBatch(C.Structure):  # include/qpmpc.h: qpmpc_gait_batch
    _fields_ = [("num_robots", C.c_int32), ("horizon", C.c_int32), ("d_phase", C.c_void_p), ("d_init", C.c_void_p),
                ("d_phase_offset", C.c_void_p), ("d_step_freq", C.c_void_p), ("d_duty_factor", C.c_void_p),
                ("d_full_stance", C.c_void_p), ("step_dts", C.c_double * MAX_H), ("d_contact", C.c_void_p)]


QPMPC_MAX_OBS_SEGMENTS = 24
OBS_DTYPE = {"float32": 0, "float64": 1, "int32": 2, "uint8": 3}


class QpmpcObsSegment(C.Structure):  # include/qpmpc.h: qpmpc_obs_segment
    _fields_ = [("d_src", C.c_void_p), ("robot_stride", C.c_int64), ("dtype", C.c_int32), ("count", C.c_int32)]


class QpmpcObsLayout(C.Structure):  # include/qpmpc.h: qpmpc_obs_layout
    _fields_ = [("num_segments", C.c_int32), ("num_columns", C.c_int32), ("seg", QpmpcObsSegment * QPMPC_MAX_OBS_SEGMENTS)]


EXPORTS = ["qpmpc_create", "qpmpc_destroy", "qpmpc_num_params", "qpmpc_last_error", "qpmpc_version", "qpmpc_stream",
           "qpmpc_launch_count", "qpmpc_last_kernel", "qpmpc_solve", "qpmpc_solve_iterated", "qpmpc_copy_costs", "qpmpc_solve_batched", "qpmpc_rollout_costs", "qpmpc_generate_noise", "qpmpc_set_profiling", "qpmpc_profile_last",
           "qpmpc_measure_fp32_peak", "qpmpc_contact_sequences", "qpmpc_gait_run",
           "qpmpc_stance_torques", "qpmpc_pack_observables"]


def build(force=False):
    """Compile the sm_100a library in-tree with nvcc (csrc/Makefile)."""
    srcs = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(_HERE, "..", "include", "qpmpc.h")]
    stale = (not os.path.exists(LIB_PATH)) or any(
        os.path.getmtime(s) > os.path.getmtime(LIB_PATH) for s in srcs if os.path.exists(s))
    if force or stale:
        subprocess.check_call(["make", "-C", CSRC, "-s"] + (["-B"] if force else []))
    return LIB_PATH


_LIB = None


def lib():
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise QpmpcError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                             "(there is no CPU fallback for the sampling path)")
        L = C.CDLL(LIB_PATH)
        fp, vp = C.POINTER(C.c_float), C.c_void_p
        L.qpmpc_create.argtypes = [C.POINTER(QpmpcConfig), C.POINTER(vp)]
        L.qpmpc_destroy.argtypes = [vp]
        L.qpmpc_num_params.argtypes = [vp]
        L.qpmpc_last_error.restype = C.c_char_p
        L.qpmpc_version.restype = C.c_char_p
        L.qpmpc_stream.argtypes = [vp]
        L.qpmpc_stream.restype = vp
        L.qpmpc_launch_count.argtypes = [vp]
        L.qpmpc_launch_count.restype = C.c_int64
        L.qpmpc_last_kernel.argtypes = [vp]
        L.qpmpc_last_kernel.restype = C.c_char_p
        L.qpmpc_solve.argtypes = [vp, vp, vp, vp, vp, C.c_uint64, C.c_uint64, vp, C.c_float, C.c_int32, vp, vp, vp,
                                  vp, vp, vp, vp]  # host pointers passed as integers (cheapest ctypes marshalling)
        L.qpmpc_solve_iterated.argtypes = [vp, vp, vp, vp, vp, C.c_uint64, C.c_uint64, C.c_int32, vp, C.c_float, C.c_int32,
                                           vp, vp, vp, vp, vp]
        L.qpmpc_copy_costs.argtypes = [vp, vp]
        L.qpmpc_solve_batched.argtypes = [vp, C.POINTER(QpmpcBatch), C.c_uint64, C.c_uint64]
        L.qpmpc_rollout_costs.argtypes = [vp, vp, vp, vp, vp, vp, vp, C.c_int32, vp]
        L.qpmpc_generate_noise.argtypes = [vp, C.c_uint64, C.c_uint64, C.c_int32, vp, C.c_float, C.c_int32, vp, vp]
        L.qpmpc_set_profiling.argtypes = [vp, C.c_int32]
        L.qpmpc_profile_last.argtypes = [vp, fp, fp]
        L.qpmpc_measure_fp32_peak.argtypes = [C.c_int32, fp]
        L.qpmpc_contact_sequences.argtypes = [vp, C.POINTER(QpmpcGaitBatch)]
        L.qpmpc_gait_run.argtypes = [vp, C.POINTER(QpmpcGaitBatch), C.c_double]
        L.qpmpc_stance_torques.argtypes = [vp, C.c_int32, vp, vp, vp]
        L.qpmpc_pack_observables.argtypes = [vp, C.POINTER(QpmpcObsLayout), C.c_int32, C.c_double, vp]
        _LIB = L
    return _LIB


def check(rc):
    if rc != 0:
        raise QpmpcError(f"qpmpc error {rc}: {lib().qpmpc_last_error().decode()}")


Q_DIAG = (0.0, 0.0, 1500.0, 200.0, 200.0, 200.0, 500.0, 500.0, 0.0, 20.0, 20.0, 50.0)  # centroidal_nmpc_jax.py:118-130


def make_config(*, mass, inertia, horizon=12, dt=0.02, parametrization="cubic_spline", num_splines=2, mu=0.5,
                grf_min=0.0, grf_max=None, sampling_method="random_sampling", num_parallel_computations=10000,
                gait_adaptive=False, use_nonuniform_discretization=False, horizon_fine_grained=2,
                dt_fine_grained=0.01, duty_factor=0.65, sigma_random_sampling=(0.2, 3, 10), sigma_mppi=3,
                step_freq_available=(1.4, 2.0, 2.4), max_robots=1, device=0):
    """Fill a qpmpc_config from the config.py keys Sampling_MPC.__init__ reads (centroidal_nmpc_jax.py:23-177)."""
    if parametrization not in PARAM:
        raise QpmpcError(f"unknown control_parametrization {parametrization!r}")
    if sampling_method not in METHOD:
        raise QpmpcError(f"unknown sampling_method {sampling_method!r}")
    if horizon > QPMPC_MAX_H:
        raise QpmpcError(f"horizon {horizon} > {QPMPC_MAX_H}")
    c = QpmpcConfig()
    c.horizon, c.parametrization, c.num_splines = int(horizon), PARAM[parametrization], int(num_splines)
    c.gait_adaptive, c.sampling_method = int(bool(gait_adaptive)), METHOD[sampling_method]
    c.num_parallel_computations, c.max_robots, c.device = int(num_parallel_computations), int(max_robots), int(device)
    c.mass = float(mass)
    c.inertia[:] = [float(v) for v in np.asarray(inertia, np.float32).reshape(9)]
    c.mu, c.f_z_min = float(mu), float(grf_min)
    c.f_z_max = float(mass * 9.81 if grf_max is None else grf_max)
    dts = np.full(horizon, dt, np.float32)
    if use_nonuniform_discretization:  # centroidal_model_jax.py:42-53
        dts[:horizon_fine_grained] = dt_fine_grained
    c.dts[:horizon] = [float(v) for v in dts]
    c.q_diag[:] = Q_DIAG
    c.duty_factor, c.mpc_dt = float(duty_factor), float(dt)
    c.sigma_random_sampling[:] = [float(v) for v in sigma_random_sampling]
    c.sigma_mppi = float(sigma_mppi)
    sfa = list(step_freq_available)[:QPMPC_MAX_FREQ]
    c.num_step_freq = len(sfa)
    c.step_freq_available[:len(sfa)] = [float(v) for v in sfa]
    if gait_adaptive:  # centroidal_nmpc_jax_gait_adaptive.py:1115-1121
        c.cem_elite, c.cem_sqrt, c.cem_lo, c.cem_hi = 200, 0, 0.5, 3.0
    else:              # centroidal_nmpc_jax.py:1075-1081
        c.cem_elite, c.cem_sqrt, c.cem_lo, c.cem_hi = 10, 1, 0.2, 5.0
    return c


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _fp(a):
    return a.ctypes.data if a is not None else None


class DeviceCosts:
    """costs[N] of a solve, still on the device -- the analogue of the jax device array upstream returns as the 7th
    value of jitted_compute_control and never reads in the control loop (srbd_controller_interface.py:131-168).
    Behaves like a read-only float32 numpy array: the values are fetched (one 4 N byte copy) the first time they are
    looked at, or just before the handle's next solve overwrites them."""

    def __init__(self, handle, n):
        self._handle, self._n, self._host = handle, n, None

    def _fetch(self):
        if self._host is None:
            h = self._handle
            out = np.empty(self._n, np.float32)
            if h is not None and getattr(h, "_h", None) is not None and h._h.value:
                check(lib().qpmpc_copy_costs(h._h, out.ctypes.data))
            else:
                raise QpmpcError("the controller was closed before its costs were read")
            self._host, self._handle = out, None
        return self._host

    shape = property(lambda self: (self._n,))
    dtype = np.dtype(np.float32)
    ndim = 1
    size = property(lambda self: self._n)

    def __array__(self, dtype=None, copy=None):
        a = self._fetch()
        return a if dtype is None else a.astype(dtype)

    def __len__(self):
        return self._n

    def __getitem__(self, k):
        return self._fetch()[k]

    def __iter__(self):
        return iter(self._fetch())

    def __repr__(self):
        return f"DeviceCosts({self._fetch()!r})" if self._host is not None else f"DeviceCosts(<{self._n} float32 on the device>)"


class Handle:
    """Owns one qpmpc_handle (device scratch + stream).  Thin: every method is one C call."""

    def __init__(self, cfg: QpmpcConfig):
        self.cfg = cfg
        self._h = C.c_void_p()
        check(lib().qpmpc_create(C.byref(cfg), C.byref(self._h)))
        self.P = lib().qpmpc_num_params(self._h)
        self.N = cfg.num_parallel_computations
        self.H = cfg.horizon

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            pending = getattr(self, "_lazy_costs", None)
            c_ = pending() if pending is not None else None
            if c_ is not None and c_._host is None:  # costs somebody still holds: fetch before the buffers go away
                try:
                    c_._fetch()
                except Exception:  # noqa: BLE001  (interpreter shutdown)
                    pass
            self._lazy_costs = None
            lib().qpmpc_destroy(self._h)
            self._h = C.c_void_p()

    __del__ = close

    @property
    def stream(self):
        return lib().qpmpc_stream(self._h)

    @property
    def launch_count(self):
        return lib().qpmpc_launch_count(self._h)

    @property
    def last_kernel(self):
        """"pair" | "quad" | "thread": the rollout kernel the most recent launch used."""
        return lib().qpmpc_last_kernel(self._h).decode()

    def _io(self):
        """Persistent host staging of solve(): float32 arrays whose addresses are taken ONCE (ndarray.ctypes.data costs
        about a microsecond per access, a dozen of them per call were a third of the end-to-end latency)."""
        io = getattr(self, "_io_cache", None)
        if io is None:
            P, H = self.P, self.H
            arr = dict(st=np.zeros(24, np.float32), rf=np.zeros(24, np.float32), cs=np.zeros(4 * H, np.float32),
                       bi=np.zeros(P, np.float32), tm=np.zeros(4, np.float32), sg=np.zeros(P, np.float32),
                       res=np.zeros(RESULT_FLOATS, np.float32), bo=np.zeros(P, np.float32), so=np.zeros(P, np.float32))
            ptr = {k: v.ctypes.data for k, v in arr.items()}
            arr["cs2"] = arr["cs"].reshape(4, H)  # a view: the [4, H] shape callers pass
            io = self._io_cache = (arr, ptr, lib().qpmpc_solve, lib().qpmpc_solve_iterated)
        return io

    def solve(self, state, reference, contact_sequence, best_in, seed, offset, timing=None, nominal_freq=0.0,
              optimize_swing=0, sigma_in=None, d_eps=None, d_freqs=None, want_costs=True, iterations=1):
        """One jitted_compute_control call; host numpy in, dict of host numpy out.  iterations > 1: that many chained
        solves on the device (keys (seed, offset), (seed, offset + 1), ...; mean and sigma carried on the device,
        srbd_controller_interface.py:131-170), outputs of the last one."""
        arr, ptr, f_solve, f_iter = self._io()
        pending = getattr(self, "_lazy_costs", None)
        if pending is not None:  # somebody still holds the previous solve's costs: fetch them before they are overwritten
            c_ = pending()
            if c_ is not None and c_._host is None:
                c_._fetch()
            self._lazy_costs = None
        try:  # the shapes the callers use: [24], [24], [4, H], [P]
            arr["st"][:] = state
            arr["rf"][:] = reference
            arr["cs2"][:] = contact_sequence
            arr["bi"][:] = best_in
        except (ValueError, TypeError):  # anything else of the right size
            cs, bi = np.asarray(contact_sequence), np.asarray(best_in)
            if cs.size != 4 * self.H or bi.size != self.P:
                raise QpmpcError(f"contact_sequence must be [4,{self.H}] and best [{self.P}]") from None
            arr["st"][:] = np.asarray(state).reshape(24)
            arr["rf"][:] = np.asarray(reference).reshape(24)
            arr["cs"][:] = cs.reshape(-1)
            arr["bi"][:] = bi.reshape(-1)
        p_tm = p_sg = p_so = None
        if timing is not None:
            arr["tm"][:] = timing
            p_tm = ptr["tm"]
        if sigma_in is not None:
            arr["sg"][:] = sigma_in  # scalar or [P]
            p_sg, p_so = ptr["sg"], ptr["so"]
        costs = np.empty(self.N, np.float32) if want_costs is True else None
        p_costs = costs.ctypes.data if costs is not None else None
        if iterations == 1:
            check(f_solve(self._h, ptr["st"], ptr["rf"], ptr["cs"], ptr["bi"], int(seed), int(offset), p_tm,
                          float(nominal_freq), int(optimize_swing), p_sg, d_eps, d_freqs, ptr["res"], ptr["bo"], p_so,
                          p_costs))
        else:
            if d_eps is not None or d_freqs is not None:
                raise QpmpcError("chained iterations draw their own noise (no d_eps / d_freqs)")
            check(f_iter(self._h, ptr["st"], ptr["rf"], ptr["cs"], ptr["bi"], int(seed), int(offset), int(iterations),
                         p_tm, float(nominal_freq), int(optimize_swing), p_sg, ptr["res"], ptr["bo"], p_so, p_costs))
        if want_costs == "lazy":
            costs = DeviceCosts(self, self.N)
            self._lazy_costs = weakref.ref(costs)
        res = arr["res"].copy()  # fresh arrays: the caller keeps them across calls
        return dict(grf=res[0:12], footholds=res[12:24], predicted_state=res[24:48], best=arr["bo"].copy(),
                    best_cost=res[48], best_freq=res[49], costs=costs,
                    sigma=arr["so"].copy() if sigma_in is not None else None,
                    best_index=int(res[50:51].view(np.int32)[0]))

    def set_profiling(self, on=True):
        check(lib().qpmpc_set_profiling(self._h, int(on)))

    def profile_last(self):
        """(rollout_ms, finalize_ms) of the last solve (needs set_profiling(True))."""
        a, b = C.c_float(), C.c_float()
        check(lib().qpmpc_profile_last(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def solve_batched(self, batch: QpmpcBatch, seed, offset):
        check(lib().qpmpc_solve_batched(self._h, C.byref(batch), int(seed), int(offset)))

    def contact_sequences(self, gait_batch: QpmpcGaitBatch):
        check(lib().qpmpc_contact_sequences(self._h, C.byref(gait_batch)))

    def gait_run(self, gait_batch: QpmpcGaitBatch, dt):
        check(lib().qpmpc_gait_run(self._h, C.byref(gait_batch), float(dt)))

    def stance_torques(self, num_robots, d_result, d_feet_jac, d_tau):
        """tau[R,4,3] = -J^T f from the device result records (wb_interface.py:368-371); device addresses."""
        check(lib().qpmpc_stance_torques(self._h, int(num_robots), d_result, d_feet_jac, d_tau))

    def pack_observables(self, layout: QpmpcObsLayout, num_robots, time, d_rows):
        """One float64 dataset row per robot for this step (MatLogger.record_step, simulation.py:241-304); device addresses."""
        check(lib().qpmpc_pack_observables(self._h, C.byref(layout), int(num_robots), float(time), d_rows))

    def rollout_costs(self, state, reference, contact_sequence, d_params, n, d_costs, timing=None, d_freqs=None):
        st, rf, cs = _f32(state).reshape(24), _f32(reference).reshape(24), _f32(contact_sequence)
        tm = _f32(timing).reshape(4) if timing is not None else None
        check(lib().qpmpc_rollout_costs(self._h, _fp(st), _fp(rf), _fp(cs), _fp(tm), d_params, d_freqs, int(n), d_costs))

    def generate_noise(self, seed, offset, d_eps_out, d_freqs_out=None, robot=0, sigma_in=None, nominal_freq=0.0,
                       optimize_swing=0):
        sg = None
        if sigma_in is not None:
            sg = _f32(np.broadcast_to(np.asarray(sigma_in, np.float32), (self.P,)))
        check(lib().qpmpc_generate_noise(self._h, int(seed), int(offset), int(robot), _fp(sg), float(nominal_freq),
                                         int(optimize_swing), d_eps_out, d_freqs_out))


def measure_fp32_peak(device=0):
    """Measured FP32 FMA throughput of `device` in TFLOP/s (FMA = 2 flops)."""
    v = C.c_float()
    check(lib().qpmpc_measure_fp32_peak(int(device), C.byref(v)))
    return v.value
