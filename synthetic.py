"""Synthetic inputs shared by tests/, bench.py and the golden-fixture generators.

Nothing here is product code or oracle code: it only fabricates terrains, per-leg height maps and TAMOLS leg jobs with
numpy so that every consumer (reference harness, CPU oracle, CUDA path) sees bit-identical float64 inputs.
"""
import numpy as np

from quadruped_pympc_tamols_b200._tamols_cabi import DIAGONAL, FLAG_HAS_VELOCITY, FLAG_STABILITY, JOB_DTYPE, LEGS

# robot -> (hip height [assumed, gym_quadruped is not in the tree], hip half-length, hip half-width)
ROBOT_GEOM = {"aliengo": (0.35, 0.24, 0.15), "go2": (0.30, 0.19, 0.13), "hyqreal1": (0.55, 0.44, 0.30)}

# config.py:209-243 verbatim (h_des is the robot hip height there)
TAMOLS_PARAMS = {
    "search_radius": 0.32, "search_resolution": 0.04, "patch_size": 3, "gradient_delta": 0.04,
    "weight_edge_avoidance": 10.0, "weight_roughness": 10, "weight_deviation": 2, "weight_kinematic": 2.0,
    "weight_nominal_kinematic": 0.0, "weight_reference_tracking": 10.0, "weight_stability": 20.0,
    "stability_margin": 0.06, "stability_hard": False, "stability_soft": True, "estimated_swing_time": 0.25,
    "h_des": 0.35,
    "l_min": {"go1": 0.15, "go2": 0.15, "aliengo": 0.1, "b2": 0.25, "hyqreal1": 0.25, "hyqreal2": 0.25,
              "mini_cheetah": 0.12, "spot": 0.20},
    "l_max": {"go1": 0.45, "go2": 0.45, "aliengo": 0.55, "b2": 0.75, "hyqreal1": 0.75, "hyqreal2": 0.75,
              "mini_cheetah": 0.40, "spot": 0.60},
    "slope_threshold": 0.7, "constraint_box_dx": 0.05, "constraint_box_dy": 0.05,
}


def tamols_params_for(robot, **overrides):
    tp = dict(TAMOLS_PARAMS)
    tp["h_des"] = ROBOT_GEOM[robot][0]
    tp.update(overrides)
    return tp


# ------------------------------------------------------------------------------------------------ terrains
def terrain_height(kind, x, y, seed=0):
    """Height field z(x, y) of the named synthetic terrain (vectorised, float64)."""
    x, y = np.asarray(x, np.float64), np.asarray(y, np.float64)
    if kind == "flat":
        return np.zeros(np.broadcast(x, y).shape)
    if kind == "stepping_stones":
        # docs/USING_STEPPING_STONES.md:87-90: radius 0.15 m, spacing 0.40 m, height 0.05 m, 3 per row, odd rows offset
        row = np.floor(x / 0.40 + 0.5)
        cx = row * 0.40
        shift = np.where(np.mod(row, 2) == 1, 0.20, 0.0)
        z = np.zeros(np.broadcast(x, y).shape)
        for k in (-1, 0, 1):
            cy = k * 0.40 + shift
            z = np.where((x - cx) ** 2 + (y - cy) ** 2 <= 0.15 ** 2, 0.05, z)
        return z
    if kind == "stepping_stones_tall":
        return terrain_height("stepping_stones", x, y, seed) * 3.0
    if kind == "stairs":
        return 0.15 * np.floor(x / 0.30)
    if kind == "noise":
        rng = np.random.default_rng(1234 + seed)
        z = np.zeros(np.broadcast(x, y).shape)
        for _ in range(12):
            kx, ky = rng.uniform(-12, 12, 2)
            z = z + rng.normal(0, 1) * np.sin(kx * x + ky * y + rng.uniform(0, 2 * np.pi))
        return z * (0.03 / np.sqrt(6.0))
    raise ValueError(kind)


def grid_xy(origin, row_axis, col_axis, rows, cols):
    """xy[r, c] = (origin + r * row_axis) + c * col_axis, the order the CUDA path and the oracle use."""
    r = np.arange(rows, dtype=np.float64)[:, None, None]
    c = np.arange(cols, dtype=np.float64)[None, :, None]
    return (np.asarray(origin, np.float64) + r * np.asarray(row_axis, np.float64)) + c * np.asarray(col_axis, np.float64)


def leg_map(kind, center_xy, yaw, rows=13, cols=7, res=0.04, seed=0):
    """A per-leg sensor map like simulation/simulation.py:493-509 (13 x 7 @ 0.04 m, centred on the nominal foothold,
    rotated by the base yaw).  Returns (origin, row_axis, col_axis, xy[R, C, 2], z[R, C])."""
    cy, sy = np.cos(yaw), np.sin(yaw)
    row_axis = np.array([cy, sy]) * res
    col_axis = np.array([-sy, cy]) * res
    origin = np.asarray(center_xy, np.float64) - ((rows - 1) / 2) * row_axis - ((cols - 1) / 2) * col_axis
    xy = grid_xy(origin, row_axis, col_axis, rows, cols)
    z = terrain_height(kind, xy[..., 0], xy[..., 1], seed)
    return origin, row_axis, col_axis, xy, z


# ------------------------------------------------------------------------------------------------ robot scenes
def robot_scene(robot, rng, kind, swing_pattern=None, speed=0.3, rows=13, cols=7, res=0.04, yaw=None, seed=0,
                map_offset_sigma=0.0):
    """One robot at a trot swing apex on terrain `kind`: per-leg maps + everything compute_adaptation receives.
    Returns a dict of float64 arrays: heights[4, R, C], xy[4, R, C, 2], jobs (JOB_DTYPE[4]), plus the raw inputs."""
    hip_h, hx, hy = ROBOT_GEOM[robot]
    yaw = rng.uniform(-0.3, 0.3) if yaw is None else yaw
    base = np.array([rng.uniform(-1.0, 1.0), rng.uniform(-0.2, 0.2), hip_h])
    cy, sy = np.cos(yaw), np.sin(yaw)
    Rz = np.array([[cy, -sy], [sy, cy]])
    hips_b = np.array([[hx, hy], [hx, -hy], [-hx, hy], [-hx, -hy]])
    hips = np.c_[base[:2] + hips_b @ Rz.T, np.full(4, hip_h)]
    vel = np.array([speed * cy, speed * sy, 0.0]) + np.r_[rng.normal(0, 0.02, 2), 0.0]
    seeds = hips.copy()
    seeds[:, :2] += rng.normal(0, 0.05, (4, 2)) + 0.1 * vel[:2]
    seeds[:, 2] = terrain_height(kind, seeds[:, 0], seeds[:, 1], seed)
    feet = hips.copy()
    feet[:, :2] += rng.normal(0, 0.04, (4, 2))
    feet[:, 2] = terrain_height(kind, feet[:, 0], feet[:, 1], seed)
    if swing_pattern is None:
        swing_pattern = [(0, 1, 1, 0), (1, 0, 0, 1)][int(rng.integers(0, 2))]  # trot apex: one diagonal pair in swing
    contact = np.array(swing_pattern, np.int64)
    heights = np.zeros((4, rows, cols))
    xy = np.zeros((4, rows, cols, 2))
    jobs = np.zeros(4, JOB_DTYPE)
    for l in range(4):
        centre = seeds[l, :2] + rng.normal(0, map_offset_sigma, 2) if map_offset_sigma > 0 else seeds[l, :2]
        o, ra, ca, xy[l], heights[l] = leg_map(kind, centre, yaw, rows, cols, res, seed)
        jobs[l] = make_job(l, o, ra, ca, 0, 0, rows, cols, seeds[l], hips[l], feet, base, vel, contact, leg=l)
    return dict(robot=robot, kind=kind, yaw=yaw, base=base, hips=hips, vel=vel, seeds=seeds, feet=feet, contact=contact,
                heights=heights, xy=xy, jobs=jobs)


def make_job(map_index, origin, row_axis, col_axis, r0, c0, nr, nc, seed, hip, feet, base, vel, contact, leg):
    """Pack one leg's inputs the way VisualFootholdAdaptation.compute_adaptation sees them
    (visual_foothold_adaptation.py:154-167, 631-675)."""
    j = np.zeros((), JOB_DTYPE)
    j["map"], j["r0"], j["c0"], j["nr"], j["nc"] = map_index, r0, c0, nr, nc
    flags = 0
    if vel is not None:
        flags |= FLAG_HAS_VELOCITY
        j["vel"] = vel
    if base is not None and feet is not None and not (contact is not None and leg < len(contact) and contact[leg] == 1):
        flags |= FLAG_STABILITY
        j["base_xy"] = base[:2]
        j["diag_foot"] = feet[LEGS.index(DIAGONAL[LEGS[leg]])][:2]
    j["flags"] = flags
    j["origin"], j["row_axis"], j["col_axis"] = origin, row_axis, col_axis
    j["seed"], j["hip"] = seed, hip
    return j


def sweep_scene(robot, rng, kind, window, map_rows=2048, map_cols=2048, res=0.02, seed=0, robots=1):
    """one axis-aligned global map and, per robot, 4 leg jobs whose candidate windows (window x window
    cells) are cut around each seed.  Returns heights[1, rows, cols] and jobs[robots * 4]."""
    hip_h, hx, hy = ROBOT_GEOM[robot]
    origin = np.array([-(map_rows - 1) / 2 * res, -(map_cols - 1) / 2 * res])
    row_axis, col_axis = np.array([res, 0.0]), np.array([0.0, res])
    xy = grid_xy(origin, row_axis, col_axis, map_rows, map_cols)
    heights = terrain_height(kind, xy[..., 0], xy[..., 1], seed)[None]
    half_extent = (min(map_rows, map_cols) / 2 - window) * res
    jobs = np.zeros(robots * 4, JOB_DTYPE)
    hips_b = np.array([[hx, hy], [hx, -hy], [-hx, hy], [-hx, -hy]])
    for k in range(robots):
        base = np.array([rng.uniform(-1, 1) * max(half_extent - hx, 0), rng.uniform(-1, 1) * max(half_extent - hy, 0), 0])
        base[2] = hip_h + float(terrain_height(kind, base[0], base[1], seed))
        hips = np.c_[base[:2] + hips_b, np.full(4, base[2])]
        vel = np.array([0.3, 0.0, 0.0]) + np.r_[rng.normal(0, 0.02, 2), 0.0]
        seeds = hips.copy()
        seeds[:, :2] += rng.normal(0, 0.05, (4, 2)) + 0.1 * vel[:2]
        seeds[:, 2] = terrain_height(kind, seeds[:, 0], seeds[:, 1], seed)
        feet = hips.copy()
        feet[:, :2] += rng.normal(0, 0.04, (4, 2))
        contact = np.array([(0, 1, 1, 0), (1, 0, 0, 1)][int(rng.integers(0, 2))])
        for l in range(4):
            rc = np.rint((seeds[l, 0] - origin[0]) / res).astype(int)
            cc = np.rint((seeds[l, 1] - origin[1]) / res).astype(int)
            r0 = int(np.clip(rc - window // 2, 0, map_rows - window))
            c0 = int(np.clip(cc - window // 2, 0, map_cols - window))
            jobs[4 * k + l] = make_job(0, origin, row_axis, col_axis, r0, c0, window, window, seeds[l], hips[l], feet,
                                       base, vel, contact, leg=l)
    return dict(heights=heights, jobs=jobs, origin=origin, row_axis=row_axis, col_axis=col_axis, res=res)
