"""ctypes binding of the TAMOLS entry points of lib/libqpmpc_b200.so (C ABI: include/tamols.h).

The only place the Python host touches native code for the TAMOLS path.  No CPU fallback: if the shared library is
missing or no sm_100 GPU is present, calls raise `TamolsError`.
"""
import ctypes as C
import math

import numpy as np

from . import _cabi

FLAG_HAS_VELOCITY = 1
FLAG_STABILITY = 2
F64, F32 = 0, 1
LEGS = ("FL", "FR", "RL", "RR")
DIAGONAL = {"FL": "RR", "FR": "RL", "RL": "FR", "RR": "FL"}  # visual_foothold_adaptation.py:640-645


class TamolsError(RuntimeError):
    pass


class TamolsParams(C.Structure):
    _fields_ = [(n, C.c_double) for n in (
        "gradient_delta", "slope_threshold", "weight_edge_avoidance", "weight_roughness", "weight_deviation",
        "weight_nominal_kinematic", "weight_reference_tracking", "weight_stability", "stability_margin",
        "estimated_swing_time", "h_des", "l_min", "l_max", "stance_duration", "sensor_offset", "foot_clearance",
        "collision_margin")] + [("collision_alphas", C.c_double * 5), ("constraint_box_dx", C.c_double),
                                ("constraint_box_dy", C.c_double)]


class TamolsLegJob(C.Structure):
    _fields_ = [("map", C.c_int32), ("r0", C.c_int32), ("c0", C.c_int32), ("nr", C.c_int32), ("nc", C.c_int32),
                ("flags", C.c_int32), ("origin", C.c_double * 2), ("row_axis", C.c_double * 2),
                ("col_axis", C.c_double * 2), ("seed", C.c_double * 3), ("hip", C.c_double * 3),
                ("diag_foot", C.c_double * 2), ("base_xy", C.c_double * 2), ("vel", C.c_double * 3)]


class TamolsResult(C.Structure):
    _fields_ = [("best_index", C.c_int32), ("num_finite", C.c_int32), ("best_score", C.c_double),
                ("best_xyz", C.c_double * 3), ("vertex1", C.c_double * 3), ("vertex2", C.c_double * 3)]


JOB_DTYPE = np.dtype([("map", "<i4"), ("r0", "<i4"), ("c0", "<i4"), ("nr", "<i4"), ("nc", "<i4"), ("flags", "<i4"),
                      ("origin", "<f8", 2), ("row_axis", "<f8", 2), ("col_axis", "<f8", 2), ("seed", "<f8", 3),
                      ("hip", "<f8", 3), ("diag_foot", "<f8", 2), ("base_xy", "<f8", 2), ("vel", "<f8", 3)])
RESULT_DTYPE = np.dtype([("best_index", "<i4"), ("num_finite", "<i4"), ("best_score", "<f8"), ("best_xyz", "<f8", 3),
                         ("vertex1", "<f8", 3), ("vertex2", "<f8", 3)])
assert JOB_DTYPE.itemsize == C.sizeof(TamolsLegJob) and RESULT_DTYPE.itemsize == C.sizeof(TamolsResult)
# include/tamols.h: tamols_sensor_leg (the data pointer travels as an integer)
SENSOR_LEG_DTYPE = np.dtype([("data", "<u8"), ("rows", "<i4"), ("cols", "<i4"), ("flags", "<i4"), ("reserved", "<i4"),
                             ("seed", "<f8", 3), ("hip", "<f8", 3), ("diag_foot", "<f8", 2), ("base_xy", "<f8", 2),
                             ("vel", "<f8", 3)])
assert SENSOR_LEG_DTYPE.itemsize == 8 + 16 + 8 * 13

class TamolsSeedBatch(C.Structure):  # include/tamols.h: tamols_seed_batch
    _fields_ = [("num_robots", C.c_int32), ("hist_len", C.c_int32),
                ("d_base_position", C.c_void_p), ("d_base_euler", C.c_void_p), ("d_base_lin_vel", C.c_void_p),
                ("d_ref_lin_vel", C.c_void_p), ("d_hips", C.c_void_p), ("d_feet", C.c_void_p), ("d_contact", C.c_void_p),
                ("d_com_height", C.c_void_p), ("d_stance_time", C.c_void_p), ("d_lift_off_z", C.c_void_p),
                ("d_com_offset_b", C.c_void_p), ("d_vel_hist", C.c_void_p), ("d_hist_count", C.c_void_p),
                ("hip_height", C.c_double), ("hip_offset", C.c_double), ("gravity", C.c_double),
                ("d_ref_feet", C.c_void_p), ("d_jobs", C.c_void_p),
                ("map_rows", C.c_int32), ("map_cols", C.c_int32), ("win_rows", C.c_int32), ("win_cols", C.c_int32),
                ("map_origin", C.c_double * 2), ("map_row_axis", C.c_double * 2), ("map_col_axis", C.c_double * 2)]


STATIC_MAX_JOBS, STATIC_MAX_PARTIALS = 4096, 65536  # include/tamols.h TAMOLS_STATIC_MAX_*

EXPORTS = ["tamols_create", "tamols_destroy", "tamols_set_params", "tamols_last_error", "tamols_stream",
           "tamols_launch_count", "tamols_last_kernel", "tamols_score_device", "tamols_score_device_static", "tamols_score_host", "tamols_seed_jobs", "tamols_fit_grid", "tamols_score_sensors"]


def make_params(tamols_params=None, robot_name="aliengo"):
    """tamols_params dict -> struct with the reference's .get(key, default) defaults
    (visual_foothold_adaptation.py:214-215, 298-305, 377-378, 438-439, 537, 636, 671)."""
    tp = dict(tamols_params or {})
    p = TamolsParams()
    p.gradient_delta = float(tp.get("gradient_delta", 0.04))
    p.slope_threshold = float(tp.get("slope_threshold", 0.7))
    p.weight_edge_avoidance = float(tp.get("weight_edge_avoidance", 15.0))
    p.weight_roughness = float(tp.get("weight_roughness", 10.0))
    p.weight_deviation = float(tp.get("weight_deviation", 1.0))
    p.weight_nominal_kinematic = float(tp.get("weight_nominal_kinematic", 20.0))
    p.weight_reference_tracking = float(tp.get("weight_reference_tracking", 2.0))
    p.weight_stability = float(tp.get("weight_stability", 10.0))
    p.stability_margin = float(tp.get("stability_margin", 0.06))
    p.estimated_swing_time = float(tp.get("estimated_swing_time", 0.25))
    p.h_des = float(tp.get("h_des", 0.25))
    p.l_min = float(tp.get("l_min", {}).get(robot_name, 0.15))
    p.l_max = float(tp.get("l_max", {}).get(robot_name, 0.45))
    p.stance_duration, p.sensor_offset, p.foot_clearance, p.collision_margin = 0.3, 0.02, 0.005, 0.02
    p.collision_alphas[:] = [float(a) for a in np.linspace(0.2, 0.8, 5)]
    p.constraint_box_dx = float(tp.get("constraint_box_dx", 0.05))
    p.constraint_box_dy = float(tp.get("constraint_box_dy", 0.05))
    return p


_FIT_OUT = None


def fit_grid(xy):
    """Regular-grid description (origin, row_axis, col_axis) of sensor coordinates xy[R, C, 2]; raises when the
    points are not a regular orthogonal grid (the analytic nearest-cell lookup needs one).  One C call
    (tamols_fit_grid, host only): the numpy version of this check cost more than the scoring launch."""
    global _FIT_OUT
    xy = np.ascontiguousarray(xy, np.float64)
    if _FIT_OUT is None:
        out = np.zeros((3, 2))
        base = out.ctypes.data
        _FIT_OUT = (out, base, base + 16, base + 32, lib().tamols_fit_grid)
    out, p0, p1, p2, f = _FIT_OUT
    check(f(xy.ctypes.data, xy.shape[0], xy.shape[1], p0, p1, p2))
    o = out.copy()
    return o[0], o[1], o[2]


_BOUND = False


def lib():
    global _BOUND
    L = _cabi.lib()
    if not _BOUND:
        vp = C.c_void_p
        L.tamols_create.argtypes = [C.POINTER(TamolsParams), C.c_int32, C.c_int32, C.POINTER(vp)]
        L.tamols_destroy.argtypes = [vp]
        L.tamols_set_params.argtypes = [vp, C.POINTER(TamolsParams)]
        L.tamols_last_error.restype = C.c_char_p
        L.tamols_stream.argtypes = [vp]
        L.tamols_stream.restype = vp
        L.tamols_launch_count.argtypes = [vp]
        L.tamols_launch_count.restype = C.c_int64
        L.tamols_last_kernel.argtypes = [vp]
        L.tamols_last_kernel.restype = C.c_char_p
        L.tamols_score_device.argtypes = [vp, vp, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, vp, vp,
                                          C.c_int32, C.c_int32, C.c_int32, vp, vp, C.c_int64]
        L.tamols_score_device_static.argtypes = L.tamols_score_device.argtypes
        L.tamols_score_host.argtypes = [vp, vp, C.c_int32, C.c_int32, C.c_int32, vp, vp, C.c_int32, vp, vp, C.c_int64]
        L.tamols_seed_jobs.argtypes = [vp, C.POINTER(TamolsSeedBatch)]
        L.tamols_fit_grid.argtypes = [vp, C.c_int32, C.c_int32, vp, vp, vp]
        L.tamols_score_sensors.argtypes = [vp, vp, C.c_int32, vp]
        _BOUND = True
    return L


def check(rc):
    if rc != 0:
        raise TamolsError(f"tamols error {rc}: {lib().tamols_last_error().decode()}")


class Handle:
    """Owns one tamols_handle (device scratch + stream).  Thin: every method is one C call."""

    def __init__(self, params: TamolsParams, device=0, max_jobs=4):
        self._h = C.c_void_p()
        self.params = params
        self.device = int(device)
        check(lib().tamols_create(C.byref(params), int(device), int(max_jobs), C.byref(self._h)))

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            lib().tamols_destroy(self._h)
            self._h = C.c_void_p()

    __del__ = close

    @property
    def stream(self):
        return lib().tamols_stream(self._h)

    @property
    def launch_count(self):
        return lib().tamols_launch_count(self._h)

    @property
    def last_kernel(self):
        return lib().tamols_last_kernel(self._h).decode()

    def set_params(self, params: TamolsParams):
        self.params = params
        check(lib().tamols_set_params(self._h, C.byref(params)))

    def score_host(self, heights, jobs, xy=None, want_scores=False):
        """heights f64 [M, R, C]; jobs structured array (JOB_DTYPE); xy optional f64 [M, R, C, 2].
        Returns (results structured array, scores [n_jobs, max cells] or None)."""
        shape = np.shape(heights)
        n = len(jobs)
        st = getattr(self, "_stage", None)
        if st is None or st[0] != (shape, n):
            # persistent host staging with addresses taken once (ndarray.ctypes.data costs ~1 us per access)
            M, R, Cc = shape
            # (records travel as bytes: numpy copies structured arrays field by field, 5 us for four jobs)
            arr = (np.zeros((M, R, Cc)), np.zeros((M, R, Cc, 2)), np.zeros(n * JOB_DTYPE.itemsize, np.uint8),
                   np.zeros(n * RESULT_DTYPE.itemsize, np.uint8))
            st = self._stage = ((shape, n), arr, tuple(a.ctypes.data for a in arr), lib().tamols_score_host)
        _, (hm, xys, jb_u8, res_u8), (p_h, p_xy, p_j, p_r), f = st
        M, R, Cc = shape
        hm[...] = heights
        jb = np.ascontiguousarray(jobs, JOB_DTYPE)
        jb_u8[:] = jb.view(np.uint8)
        if xy is not None:
            xys[...] = xy
        scores, stride, p_s = None, 0, None
        if want_scores:
            stride = int(np.max(jb["nr"].astype(np.int64) * jb["nc"]))
            scores = np.empty((n, stride), np.float64)
            p_s = scores.ctypes.data
        check(f(self._h, p_h, M, R, Cc, p_xy if xy is not None else None, p_j, n, p_r, p_s, stride))
        return res_u8.copy().view(RESULT_DTYPE), scores

    def score_sensors(self, legs):
        """legs: structured array (SENSOR_LEG_DTYPE) whose `data` fields are the addresses of float64 [R, C, 1, 3]
        arrays the caller keeps alive.  Returns the results structured array."""
        n = len(legs)
        st = getattr(self, "_sensor_stage", None)
        if st is None or st[0] != n:
            res = np.zeros(n * RESULT_DTYPE.itemsize, np.uint8)
            st = self._sensor_stage = (n, res, res.ctypes.data, lib().tamols_score_sensors)
        _, res, p_res, f = st
        check(f(self._h, legs.ctypes.data, n, p_res))
        return res.copy().view(RESULT_DTYPE)

    def seed_jobs(self, batch: TamolsSeedBatch):
        """Seed footholds (and job records) of a batch of robots; asynchronous on self.stream."""
        check(lib().tamols_seed_jobs(self._h, C.byref(batch)))

    def score_device(self, d_heights, height_dtype, num_maps, rows, cols, pitch, d_jobs, n_jobs, max_nr, max_nc,
                     d_results, d_xy=None, d_scores=None, scores_stride=0):
        """All pointers are device addresses (ints); asynchronous on self.stream."""
        check(lib().tamols_score_device(self._h, d_heights, int(height_dtype), int(num_maps), int(rows), int(cols),
                                        int(pitch), d_xy, d_jobs, int(n_jobs), int(max_nr), int(max_nc), d_results,
                                        d_scores, int(scores_stride)))

    def score_device_static(self, d_heights, height_dtype, num_maps, rows, cols, pitch, d_jobs, n_jobs, max_nr, max_nc,
                            d_results, d_xy=None, d_scores=None, scores_stride=0):
        """score_device with the kernel's scratch in the library's static device array: no allocation inside the call
        (include/tamols.h: limits STATIC_MAX_JOBS / STATIC_MAX_PARTIALS, calls that overlap on the GPU race)."""
        check(lib().tamols_score_device_static(self._h, d_heights, int(height_dtype), int(num_maps), int(rows), int(cols),
                                               int(pitch), d_xy, d_jobs, int(n_jobs), int(max_nr), int(max_nc),
                                               d_results, d_scores, int(scores_stride)))
