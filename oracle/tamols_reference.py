"""Runs the reference's OWN TAMOLS code (/root/reference/quadruped_pympc/helpers/visual_foothold_adaptation.py,
unmodified) on leg jobs in this repo's format.  TEST INFRASTRUCTURE; works only where /root/reference exists (the build
container): used by oracle/make_golden_tamols.py to create tests/golden/tamols_*.npz and by the container-only
cross-check test.  gym_quadruped / mujoco are absent, so the stubs of oracle/jax_numpy_standin.py stand in for
LegsAttr, the robot config and mujoco.viewer.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
REF = os.environ.get("QPMPC_REFERENCE", "/root/reference")
LEGS = ("FL", "FR", "RL", "RR")


def available():
    return os.path.exists(os.path.join(REF, "quadruped_pympc", "helpers", "visual_foothold_adaptation.py"))


_MOD = None


def _load():
    """Import the reference module under stubs without letting this repo's drop-in package shadow it."""
    global _MOD
    if _MOD is None:
        sys.path.insert(0, HERE)
        import jax_numpy_standin as standin
        standin.install_robot_stubs()
        saved = {k: v for k, v in sys.modules.items() if k == "quadruped_pympc" or k.startswith("quadruped_pympc.")}
        for k in saved:
            del sys.modules[k]
        sys.path.insert(0, REF)
        try:
            import quadruped_pympc.config as cfg
            import quadruped_pympc.helpers.visual_foothold_adaptation as vfa_mod
            assert os.path.abspath(vfa_mod.__file__).startswith(os.path.abspath(REF)), vfa_mod.__file__
        finally:
            sys.path.remove(REF)
        _MOD = (cfg, vfa_mod, standin.LegsAttr)
    return _MOD


class SensorMap:
    """Stand-in for gym_quadruped.sensors.heightmap.HeightMap: `.data[R, C, 1, 3]` and `.get_height(p)`."""

    def __init__(self, xy, z):
        self.data = np.concatenate([xy, z[..., None]], axis=-1)[:, :, None, :].copy() if xy is not None else None

    def get_height(self, p):
        d = self.data[:, :, 0, :]
        k = np.argmin((d[..., 0] - p[0]) ** 2 + (d[..., 1] - p[1]) ** 2)
        return float(d.reshape(-1, 3)[k, 2])


def run(robot, tamols_params, heights, xy, jobs, contact=None, with_scores=True):
    """One compute_adaptation call on 4 leg jobs (legs FL, FR, RL, RR in order).
    heights [M, R, C], xy [M, R, C, 2]; each job's window of its map is that leg's sensor map.
    Returns dict(best_xyz[4,3], vertex1/2[4,3] (nan when unset), scores[4, max cells], best_index[4], ret)."""
    cfg, vfa_mod, LegsAttr = _load()
    cfg.robot = robot
    cfg.simulation_params["tamols_params"] = dict(tamols_params)
    vfa = vfa_mod.VisualFootholdAdaptation(legs_order=LEGS, adaptation_strategy="tamols")
    assert len(jobs) == 4
    maps, seeds, hips = {}, {}, {}
    for l, name in enumerate(LEGS):
        j = jobs[l]
        sl = (int(j["map"]), slice(int(j["r0"]), int(j["r0"] + j["nr"])), slice(int(j["c0"]), int(j["c0"] + j["nc"])))
        maps[name] = SensorMap(xy[sl], heights[sl])
        seeds[name] = np.array(j["seed"], np.float64)
        hips[name] = np.array(j["hip"], np.float64)
    j0 = jobs[0]
    has_vel = bool(j0["flags"] & 1)
    vel = np.array(j0["vel"], np.float64) if has_vel else None
    stab = [bool(j["flags"] & 2) for j in jobs]
    base, feet = None, None
    if any(stab):
        k = stab.index(True)
        base = np.array([jobs[k]["base_xy"][0], jobs[k]["base_xy"][1], 0.0])
        fp = {}
        for l, name in enumerate(LEGS):  # diagonal foot of leg l is the current foot of its partner
            partner = {"FL": "RR", "FR": "RL", "RL": "FR", "RR": "FL"}[name]
            if stab[l]:
                fp[partner] = np.array([jobs[l]["diag_foot"][0], jobs[l]["diag_foot"][1], 0.0])
        for name in LEGS:
            fp.setdefault(name, np.zeros(3))
        feet = LegsAttr(**fp)
        if contact is None:
            contact = np.array([0 if s else 1 for s in stab])
    ref = LegsAttr(**{n: seeds[n].copy() for n in LEGS})
    ret = vfa.compute_adaptation(LEGS, ref, LegsAttr(**hips), LegsAttr(**maps), vel, np.zeros(3), np.zeros(3),
                                 base_position=base, current_contact=contact, current_feet_pos=feet)
    out = dict(ret=ret, best_xyz=np.stack([np.asarray(ref[n], np.float64) for n in LEGS]))
    v1, v2 = np.full((4, 3), np.nan), np.full((4, 3), np.nan)
    for l, n in enumerate(LEGS):
        if vfa.footholds_constraints[n] is not None:
            v1[l], v2[l] = vfa.footholds_constraints[n]
    out["vertex1"], out["vertex2"] = v1, v2
    if with_scores:  # the loop of :188-206 once more, recording every score
        stride = int(max(int(j["nr"]) * int(j["nc"]) for j in jobs))
        scores = np.full((4, stride), np.inf)
        best = np.full(4, -1, np.int64)
        for l, n in enumerate(LEGS):
            fast = vfa_mod.FastHeightMap(maps[n])
            bs = float("inf")
            for i, cxy in enumerate(vfa._generate_candidates(fast, seeds[n])):
                cand = np.array([cxy[0], cxy[1], 0.0])
                cand[2] = fast.get_height(cand) + 0.005
                s = vfa._compute_tamols_score(cand, seeds[n], hips[n], fast, l, n, LEGS)
                scores[l, i] = s
                if s < bs:
                    bs, best[l] = s, i
        out["scores"], out["best_index"] = scores, best
    return out
