"""CPU restatement of the reference's seed-foothold heuristic and of the job records built from the seeds, for a batch
of robots.  TEST INFRASTRUCTURE (tests/, smoke, bench CPU legs only); the product path is the CUDA kernel behind
tamols_seed_jobs (include/tamols.h).

(a) seed_footholds: quadruped_pympc/helpers/foothold_reference_generator.py:84-157 (compute_footholds_reference), float64,
    the moving-average deque (:95-96) as an explicit list per robot.  Pinned by tests/golden/seed_footholds.npz, produced
    by oracle/make_golden_seed.py from the reference's own class.
(b) jobs_for_global_map: the tamols_leg_job records VisualFootholdAdaptation.compute_adaptation would assemble
    (helpers/visual_foothold_adaptation.py:154-166, 631-664) when every leg's candidate window is cut from ONE global
    regular height map around the seed.  The reference has no such function (its windows are MuJoCo ray-casts,
    wb_interface.py:233-234): this part is "parity unpinned" against the reference and pinned only against this
    restatement.
"""
import numpy as np


def seed_footholds(base_position, base_euler, base_lin_vel, ref_lin_vel, hips, com_height, stance_time, lift_off_z,
                   histories, hip_height, hip_offset=0.1, gravity=9.81, hist_len=20, com_offset_b=None):
    """[R,4,3] reference footholds; `histories` (list of R lists of 2-vectors) is appended to and trimmed in place."""
    R = len(base_position)
    out = np.zeros((R, 4, 3))
    for r in range(R):
        yaw = base_euler[r][2]
        W2H = np.array([np.cos(yaw), np.sin(yaw), -np.sin(yaw), np.cos(yaw)]).reshape(2, 2)
        v_h = W2H @ np.asarray(base_lin_vel[r][:2], np.float64)
        ref_h = W2H @ np.asarray(ref_lin_vel[r][:2], np.float64)
        histories[r].append(v_h)
        del histories[r][:-hist_len]
        mvg = np.mean(list(histories[r]), axis=0)
        delta = np.clip((stance_time[r] / 2.0) * ref_h, -hip_height * 1.5, hip_height * 1.5)
        err = np.sqrt(com_height[r] / gravity) * (mvg - ref_h)
        err = np.where(err > 0.05, 0.05, err)
        err = np.where(err < -0.05, -0.05, err)
        off = delta + err
        cw = np.zeros(3)
        if com_offset_b is not None:
            cr, sr = np.cos(base_euler[r][0]), np.sin(base_euler[r][0])
            cp, sp = np.cos(base_euler[r][1]), np.sin(base_euler[r][1])
            cy, sy = np.cos(yaw), np.sin(yaw)
            Rm = np.array([[cp * cy, sr * sp * cy - cr * sy, cr * sp * cy + sr * sy],
                           [cp * sy, sr * sp * sy + cr * cy, cr * sp * sy - sr * cy],
                           [-sp, sr * cp, cr * cp]])
            cw = Rm @ np.asarray(com_offset_b[r], np.float64)
        for l in range(4):
            f = W2H @ (np.asarray(hips[r][l][:2], np.float64) - np.asarray(base_position[r][:2], np.float64))
            f[1] += -hip_offset if l & 1 else hip_offset
            f = f + off
            w = W2H.T @ f + np.asarray(base_position[r][:2], np.float64) + cw[:2]
            out[r, l] = [w[0], w[1], lift_off_z[r][l]]
    return out


def jobs_for_global_map(job_dtype, seeds, hips, feet, contact, base_position, base_lin_vel, map_shape, origin, row_axis,
                        col_axis, win, flag_velocity=1, flag_stability=2):
    """tamols_leg_job[4R] (numpy structured array of `job_dtype`) for windows of `win` = (rows, cols) cells centred on
    the cell nearest to each seed, clamped to the map."""
    R = len(seeds)
    jobs = np.zeros(4 * R, job_dtype)
    origin, row_axis, col_axis = (np.asarray(a, np.float64) for a in (origin, row_axis, col_axis))
    nr, nc = min(win[0], map_shape[0]), min(win[1], map_shape[1])
    for r in range(R):
        for l in range(4):
            j = jobs[4 * r + l]
            q = seeds[r, l, :2] - origin
            ri = int(np.rint((q[0] * row_axis[0] + q[1] * row_axis[1]) / (row_axis[0] * row_axis[0] + row_axis[1] * row_axis[1])))
            ci = int(np.rint((q[0] * col_axis[0] + q[1] * col_axis[1]) / (col_axis[0] * col_axis[0] + col_axis[1] * col_axis[1])))
            j["map"], j["nr"], j["nc"] = 0, nr, nc
            j["r0"] = min(max(ri - nr // 2, 0), map_shape[0] - nr)
            j["c0"] = min(max(ci - nc // 2, 0), map_shape[1] - nc)
            j["origin"], j["row_axis"], j["col_axis"] = origin, row_axis, col_axis
            j["seed"], j["hip"], j["vel"] = seeds[r, l], hips[r][l], base_lin_vel[r]
            flags = flag_velocity
            stance = contact is not None and contact[r][l] == 1
            if feet is not None and not stance:
                flags |= flag_stability
                j["base_xy"] = base_position[r][:2]
                j["diag_foot"] = feet[r][3 - l][:2]
            j["flags"] = flags
    return jobs
