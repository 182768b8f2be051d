"""Generates tests/golden/seed_footholds.npz from the reference's OWN FootholdReferenceGenerator
(/root/reference/quadruped_pympc/helpers/foothold_reference_generator.py, unmodified, imported under the gym_quadruped /
mujoco stubs of oracle/jax_numpy_standin.py).  Build container only.  Each robot is one generator object driven for T
control steps (so that the 20-entry velocity deque fills and rolls over)."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("QPMPC_REFERENCE", "/root/reference")


def main():
    sys.path.insert(0, HERE)
    import jax_numpy_standin as standin
    standin.install_robot_stubs()
    for k in [k for k in sys.modules if k == "quadruped_pympc" or k.startswith("quadruped_pympc.")]:
        del sys.modules[k]
    sys.path.insert(0, REF)
    import quadruped_pympc.helpers.foothold_reference_generator as mod
    assert os.path.abspath(mod.__file__).startswith(os.path.abspath(REF))
    LegsAttr = standin.LegsAttr
    rng = np.random.default_rng(77)
    R, T = 12, 30
    hip_height = 0.33
    rec = dict(base_position=np.zeros((T, R, 3)), base_euler=np.zeros((T, R, 3)), base_lin_vel=np.zeros((T, R, 3)),
               ref_lin_vel=np.zeros((T, R, 3)), hips=np.zeros((T, R, 4, 3)), com_height=np.zeros((T, R)),
               stance_time=np.zeros((T, R)), lift_off_z=np.zeros((T, R, 4)), com_offset_b=np.zeros((R, 3)),
               ref_feet=np.zeros((T, R, 4, 3)))
    for r in range(R):
        lift = LegsAttr(*[np.array([0.0, 0.0, rng.normal(0, 0.05)]) for _ in range(4)])
        g = mod.FootholdReferenceGenerator(stance_time=0.4, lift_off_positions=lift, hip_height=hip_height)
        if r % 3 == 2:
            g.com_pos_offset_b = rng.normal(0, 0.02, 3)
        rec["com_offset_b"][r] = g.com_pos_offset_b
        pos = rng.normal(0, 2.0, 3)
        pos[2] = 0.3 + rng.normal(0, 0.02)
        for t in range(T):
            eul = np.array([rng.normal(0, 0.05), rng.normal(0, 0.05), rng.uniform(-3.1, 3.1)])
            vel = rng.normal(0, 0.6, 3)
            ref = np.array([rng.normal(0, 1.2), rng.normal(0, 0.3), 0.0]) * (3.0 if r == 5 else 1.0)  # r = 5 hits the clip
            hips = np.array([[0.2, 0.1, 0.0], [0.2, -0.1, 0.0], [-0.2, 0.1, 0.0], [-0.2, -0.1, 0.0]]) + rng.normal(0, 0.01, (4, 3))
            c, s = np.cos(eul[2]), np.sin(eul[2])
            hips_w = hips @ np.array([[c, s, 0], [-s, c, 0], [0, 0, 1.0]]) + pos
            g.stance_time = float(rng.uniform(0.2, 0.5))
            for l, name in enumerate(("FL", "FR", "RL", "RR")):
                g.lift_off_positions[name][2] = rng.normal(0, 0.05)
                rec["lift_off_z"][t, r, l] = g.lift_off_positions[name][2]
            com_h = float(rng.uniform(0.25, 0.6))
            out = g.compute_footholds_reference(base_position=pos, base_ori_euler_xyz=eul, base_xy_lin_vel=vel[:2],
                                                ref_base_xy_lin_vel=ref[:2],
                                                hips_position=LegsAttr(*[hips_w[l] for l in range(4)]),
                                                com_height_nominal=com_h)
            rec["base_position"][t, r], rec["base_euler"][t, r] = pos, eul
            rec["base_lin_vel"][t, r], rec["ref_lin_vel"][t, r] = vel, ref
            rec["hips"][t, r], rec["com_height"][t, r], rec["stance_time"][t, r] = hips_w, com_h, g.stance_time
            rec["ref_feet"][t, r] = np.array([out[n] for n in ("FL", "FR", "RL", "RR")])
            pos = pos + vel * 0.02
    out = os.path.join(os.path.dirname(HERE), "tests", "golden", "seed_footholds.npz")
    np.savez_compressed(out, hip_height=hip_height, hip_offset=0.1, gravity=9.81, hist_len=20, **rec)
    print("wrote", out, rec["ref_feet"].shape)


if __name__ == "__main__":
    main()
