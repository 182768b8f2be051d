"""Generate tests/golden/tamols_*.npz by running the reference's OWN TAMOLS source
(/root/reference/quadruped_pympc/helpers/visual_foothold_adaptation.py, unmodified, scipy cKDTree and all) through
oracle/tamols_reference.py.

TEST INFRASTRUCTURE.  Run in the build container only (needs /root/reference):
    python oracle/make_golden_tamols.py
Each fixture stores the inputs in this repo's job format (heights, sensor xy, leg jobs, tamols_params as the struct
bytes) and the reference's outputs: every candidate's score, the winning index, the adapted foothold and the
constraint box.
"""
import ctypes as C
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
sys.path.insert(0, REPO)

import synthetic as sy  # noqa: E402
from oracle import tamols_reference as tr  # noqa: E402
from quadruped_pympc_tamols_b200 import _tamols_cabi as tc  # noqa: E402

OUT = os.path.join(REPO, "tests", "golden")


def save(name, robot, tp, heights, xy, jobs, contact=None):
    ref = tr.run(robot, tp, heights, xy, jobs, contact=contact)
    p = tc.make_params(tp, robot)
    np.savez_compressed(
        os.path.join(OUT, f"tamols_{name}.npz"), robot=robot, params=np.frombuffer(bytes(p), np.uint8),
        heights=heights, xy=xy, jobs=jobs, ref_scores=ref["scores"], ref_best_index=ref["best_index"],
        ref_best_xyz=ref["best_xyz"], ref_vertex1=ref["vertex1"], ref_vertex2=ref["vertex2"])
    fin = np.isfinite(ref["scores"]).sum(1)
    print(f"tamols_{name}: best {ref['best_index'].tolist()} finite {fin.tolist()}")


def multi_tile():
    """One 96 x 96 candidate window per leg on a 128 x 128 @ 0.02 m map: 36 tiles of 8 x 32 or 9 tiles of 32 x 32 per
    job in k_tamols_score, so the per-tile partials, the arrival ticket and the merge of the last block are pinned on
    the reference's own loop (visual_foothold_adaptation.py:188-206), not only on the oracle.  36,864 candidates
    through the reference's Python loop: about a minute."""
    rng = np.random.default_rng(500)
    sc = sy.sweep_scene("go2", rng, "stepping_stones_tall", window=96, map_rows=128, map_cols=128, res=0.02, seed=5)
    xy = sy.grid_xy(sc["origin"], sc["row_axis"], sc["col_axis"], 128, 128)[None]
    save("c4_go2_stepping_stones_tall_w96", "go2", sy.tamols_params_for("go2"), sc["heights"], xy, sc["jobs"])


def half_cell_delta():
    """gradient_delta = half the cell size on the sensor's (yaw-rotated, offset) 13 x 7 @ 0.04 m maps: the nine patch
    points sit half a cell from the candidate, so the kernel's tie detection sends the block down the per-point
    look-up path (Stencil.exact).  On such grids no query is exactly equidistant from two cells and the analytic
    nearest cell is the KD-tree's.  (On an AXIS-ALIGNED grid with this delta the queries are exactly equidistant; the
    reference's answer then depends on float64 round-off of the two squared distances and, when those are equal too --
    61 % of the cases measured -- on scipy's tree traversal: not reproducible by any rule)"""
    rng = np.random.default_rng(600)
    sc = sy.robot_scene("go2", rng, "noise", speed=0.4, seed=21, map_offset_sigma=0.05)
    save("go2_noise_halfcell_delta", "go2", sy.tamols_params_for("go2", gradient_delta=0.02), sc["heights"], sc["xy"],
         sc["jobs"], sc["contact"])


def main():
    os.makedirs(OUT, exist_ok=True)
    if "--multi-tile-only" in sys.argv:
        return multi_tile()
    if "--half-cell-only" in sys.argv:
        return half_cell_delta()
    # C3-T: hyqreal on stepping stones, 13 x 7 @ 0.04 m per-leg maps, trot apex
    for k in range(3):
        rng = np.random.default_rng(100 + k)
        sc = sy.robot_scene("hyqreal1", rng, "stepping_stones", seed=k, map_offset_sigma=0.03 * k)
        save(f"c3t_hyqreal_stones_{k}", "hyqreal1", sy.tamols_params_for("hyqreal1"), sc["heights"], sc["xy"],
             sc["jobs"], sc["contact"])
    cases = [("aliengo", "stepping_stones_tall", 0.6, 0.06), ("go2", "stepping_stones_tall", 0.5, 0.05),
             ("aliengo", "stairs", 0.4, 0.05), ("go2", "noise", 0.3, 0.04), ("aliengo", "flat", 0.3, 0.0),
             ("hyqreal1", "stairs", 0.8, 0.08)]
    for k, (robot, kind, speed, off) in enumerate(cases):
        rng = np.random.default_rng(200 + k)
        sc = sy.robot_scene(robot, rng, kind, speed=speed, seed=k, map_offset_sigma=off)
        save(f"{robot}_{kind}", robot, sy.tamols_params_for(robot), sc["heights"], sc["xy"], sc["jobs"], sc["contact"])
    # non-default weights incl. the nominal-kinematic term (0 in config.py) and the .get() defaults (empty dict)
    rng = np.random.default_rng(300)
    sc = sy.robot_scene("go2", rng, "noise", speed=0.5, seed=7, map_offset_sigma=0.05)
    save("go2_noise_defaults", "go2", {}, sc["heights"], sc["xy"], sc["jobs"], sc["contact"])
    tp = sy.tamols_params_for("aliengo", weight_nominal_kinematic=20.0, weight_edge_avoidance=15.0, gradient_delta=0.08)
    rng = np.random.default_rng(301)
    sc = sy.robot_scene("aliengo", rng, "stepping_stones_tall", speed=0.5, seed=8, map_offset_sigma=0.05)
    save("aliengo_stones_nominal_delta8", "aliengo", tp, sc["heights"], sc["xy"], sc["jobs"], sc["contact"])
    # edge cases: backwards and (almost) zero velocity, all legs in stance (stability term off), all four in swing
    rng = np.random.default_rng(302)
    sc = sy.robot_scene("aliengo", rng, "stairs", speed=-0.4, seed=9, map_offset_sigma=0.05, swing_pattern=(0, 0, 0, 0))
    save("aliengo_stairs_backwards_allswing", "aliengo", sy.tamols_params_for("aliengo"), sc["heights"], sc["xy"],
         sc["jobs"], sc["contact"])
    rng = np.random.default_rng(303)
    sc = sy.robot_scene("go2", rng, "stepping_stones_tall", speed=0.0, seed=10, map_offset_sigma=0.05,
                        swing_pattern=(1, 1, 1, 1))
    sc["jobs"]["vel"] = [1e-3, 2e-3, 0.0]
    save("go2_stones_still_allstance", "go2", sy.tamols_params_for("go2"), sc["heights"], sc["xy"], sc["jobs"],
         sc["contact"])
    # unreachable maps (hip 1 m above): every candidate rejected -> fallback branch (:223-228)
    rng = np.random.default_rng(304)
    sc = sy.robot_scene("go2", rng, "flat", seed=11)
    sc["jobs"]["hip"][:, 2] += 1.0
    save("go2_unreachable", "go2", sy.tamols_params_for("go2"), sc["heights"], sc["xy"], sc["jobs"], sc["contact"])
    # C4-style windows: one axis-aligned 96 x 96 @ 0.02 m map, four 32 x 32 candidate windows (gradient_delta = 2 cells)
    for k, kind in enumerate(("flat", "stepping_stones_tall", "noise", "stairs")):
        rng = np.random.default_rng(400 + k)
        sc = sy.sweep_scene("go2", rng, kind, window=32, map_rows=96, map_cols=96, res=0.02, seed=k)
        xy = sy.grid_xy(sc["origin"], sc["row_axis"], sc["col_axis"], 96, 96)[None]
        save(f"c4_go2_{kind}_w32", "go2", sy.tamols_params_for("go2"), sc["heights"], xy, sc["jobs"])
    multi_tile()
    half_cell_delta()


if __name__ == "__main__":
    main()
