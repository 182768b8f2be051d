"""TEST INFRASTRUCTURE -- CPU restatement of the reference's stance torque mapping.  Only tests/ may
import this; the product path (qpmpc_stance_torques, include/qpmpc.h) never does.

Reference: WBInterface.compute_stance_and_swing_torque, interfaces/wb_interface.py:368-371

    tau.FL = -np.matmul(feet_jac.FL[:, legs_qvel_idx.FL].T, nmpc_GRFs.FL)      (and FR, RL, RR)

The reference statement IS a numpy expression; `stance_torques` below evaluates that very expression per robot and leg
(same matmul, same float32 -> float64 promotion of the jax force array), so the oracle is pinned by construction:
there is nothing to restate beyond the indexing of a [R,4,3,3] batch.  The surrounding method needs MuJoCo objects and
cannot be imported here.
"""
import numpy as np


def stance_torques(feet_jac, grfs):
    """feet_jac [R,4,3,3] float64 (each leg's feet_jac[leg][:, legs_qvel_idx[leg]]), grfs [R,12] float32 (nmpc_GRFs of
    each robot, legs FL FR RL RR) -> tau [R,4,3] float64."""
    feet_jac, grfs = np.asarray(feet_jac, np.float64), np.asarray(grfs, np.float32)
    R = feet_jac.shape[0]
    tau = np.zeros((R, 4, 3))
    for r in range(R):
        for leg in range(4):
            tau[r, leg] = -np.matmul(feet_jac[r, leg].T, grfs[r, 3 * leg:3 * leg + 3])  # wb_interface.py:368-371
    return tau
