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
