"""ctypes front-end of oracle/tamols_oracle.c (TEST INFRASTRUCTURE, not product code).

The C file restates /root/reference/quadruped_pympc/helpers/visual_foothold_adaptation.py:153-714 in float64; this
module only marshals numpy arrays.  The struct layouts are those of include/tamols.h (shared with the product binding).
"""
import ctypes as C
import os
import subprocess

import numpy as np

from quadruped_pympc_tamols_b200._tamols_cabi import JOB_DTYPE, RESULT_DTYPE, TamolsParams  # struct layouts only

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIBS = {}


def build(force=False):
    out = os.path.join(_HERE, "_build", "libtamols_oracle.so")
    src = os.path.join(_HERE, "tamols_oracle.c")
    if force or not os.path.exists(out) or os.path.getmtime(out) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return out


def _lib(fast=False):
    if fast not in _LIBS:
        build()
        L = C.CDLL(os.path.join(_HERE, "_build", "libtamols_oracle_fast.so" if fast else "libtamols_oracle.so"))
        vp = C.c_void_p
        L.tamols_oracle_score.argtypes = [C.POINTER(TamolsParams), vp, C.c_int32, C.c_int32, C.c_int32, C.c_int32, vp,
                                          vp, C.c_int32, vp, vp, C.c_int64]
        _LIBS[fast] = L
    return _LIBS[fast]


def max_threads(fast=True):
    return int(_lib(fast).tamols_oracle_max_threads())


def score(params: TamolsParams, heights, jobs, xy=None, want_scores=True, fast=False):
    """heights f64 [M, R, C]; jobs JOB_DTYPE array -> (results RESULT_DTYPE array, scores [n, max cells] | None)."""
    hm = np.ascontiguousarray(heights, np.float64)
    M, R, Cc = hm.shape
    jb = np.ascontiguousarray(jobs, JOB_DTYPE)
    n = jb.shape[0]
    res = np.zeros(n, RESULT_DTYPE)
    xyp = None
    if xy is not None:
        xy = np.ascontiguousarray(xy, np.float64)
        xyp = xy.ctypes.data
    scores, stride = None, 0
    if want_scores:
        stride = int(np.max(jb["nr"].astype(np.int64) * jb["nc"]))
        scores = np.full((n, stride), np.inf)
    _lib(fast).tamols_oracle_score(C.byref(params), hm.ctypes.data, M, R, Cc, Cc, xyp, jb.ctypes.data, n,
                                   res.ctypes.data, scores.ctypes.data if scores is not None else None, stride)
    return res, scores
