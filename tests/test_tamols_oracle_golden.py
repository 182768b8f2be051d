"""CPU: the C oracle (oracle/tamols_oracle.c) against fixtures produced by executing the reference's own TAMOLS source
(visual_foothold_adaptation.py with scipy cKDTree and np.linalg.lstsq) -- oracle/make_golden_tamols.py."""
import os

import numpy as np
import pytest

from oracle import tamols_oracle as to
from tests.util import check_tamols_against_reference, load_tamols_fixture, tamols_fixtures


@pytest.mark.parametrize("path", tamols_fixtures(), ids=lambda p: os.path.basename(p)[7:-4])
@pytest.mark.parametrize("explicit_xy", [True, False], ids=["sensor_xy", "grid_xy"])
def test_oracle_matches_reference_fixture(path, explicit_xy):
    p, heights, xy, jobs, g = load_tamols_fixture(path)
    res, scores = to.score(p, heights, jobs, xy=xy if explicit_xy else None)
    check_tamols_against_reference(res, scores, g)


def test_fast_build_is_bit_identical():
    """The -O3 -fopenmp object timed as the CPU baseline computes the same scores as the strict checker."""
    p, heights, xy, jobs, g = load_tamols_fixture(tamols_fixtures()[0])
    a = to.score(p, heights, jobs, xy=xy, fast=False)
    b = to.score(p, heights, jobs, xy=xy, fast=True)
    assert np.array_equal(a[1], b[1]) and np.array_equal(a[0]["best_index"], b[0]["best_index"])


def test_known_answer_flat_map_picks_cell_nearest_seed():
    """flat map, zero velocity -> argmin = reachable cell minimising the deviation term."""
    import synthetic as sy
    from quadruped_pympc_tamols_b200 import _tamols_cabi as tc
    rng = np.random.default_rng(5)
    sc = sy.robot_scene("aliengo", rng, "flat", speed=0.0, map_offset_sigma=0.05, swing_pattern=(1, 1, 1, 1))
    sc["jobs"]["vel"] = 0.0
    p = tc.make_params(sy.tamols_params_for("aliengo"), "aliengo")
    res, scores = to.score(p, sc["heights"], sc["jobs"], xy=sc["xy"])
    for l in range(4):
        d = np.sum((sc["xy"][l].reshape(-1, 2) - sc["seeds"][l, :2]) ** 2, axis=1)
        assert res["best_index"][l] == int(np.argmin(d))


def test_known_answer_step_edge_is_avoided():
    """with a tall step under the seed the chosen cell is not adjacent to the discontinuity."""
    import synthetic as sy
    from quadruped_pympc_tamols_b200 import _tamols_cabi as tc
    rows, cols, res_m = 13, 7, 0.04
    o, ra, ca = np.array([0.0, 0.0]), np.array([res_m, 0.0]), np.array([0.0, res_m])
    xy = sy.grid_xy(o, ra, ca, rows, cols)[None]
    heights = np.where(xy[..., 0] >= 6.5 * res_m, 0.12, 0.0)  # step between rows 6 and 7
    hip = np.array([6 * res_m, 3 * res_m, 0.35])
    seed = np.array([6 * res_m, 3 * res_m, 0.0])  # directly on the lower lip of the step
    job = sy.make_job(0, o, ra, ca, 0, 0, rows, cols, seed, hip, None, None, np.zeros(3), None, leg=0)
    p = tc.make_params(sy.tamols_params_for("aliengo"), "aliengo")
    res, _ = to.score(p, heights, np.array([job]), xy=xy)
    r = int(res["best_index"][0]) // cols
    assert r <= 5 or r >= 8, r
