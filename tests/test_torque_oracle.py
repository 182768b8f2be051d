"""CPU: the stance torque oracle is the reference's numpy expression (interfaces/wb_interface.py:368-371); checked here
against an independent einsum and on the hand-computable cases."""
import numpy as np

from oracle import torque_oracle as to


def test_oracle_equals_einsum_and_hand_cases():
    rng = np.random.default_rng(0)
    jac, grf = rng.normal(size=(7, 4, 3, 3)), rng.normal(size=(7, 12)).astype(np.float32)
    tau = to.stance_torques(jac, grf)
    ref = -np.einsum("rlij,rli->rlj", jac, grf.astype(np.float64).reshape(7, 4, 3))
    assert tau.shape == (7, 4, 3) and np.allclose(tau, ref, rtol=1e-14, atol=1e-14)
    eye = np.tile(np.eye(3), (1, 4, 1, 1))
    f = np.arange(12, dtype=np.float32)[None]
    assert np.array_equal(to.stance_torques(eye, f).ravel(), -f.ravel().astype(np.float64))
    # a pure z force on a leg whose Jacobian maps joint 2 to z only
    J = np.zeros((1, 4, 3, 3))
    J[0, 2, 2, 1] = 0.25
    f = np.zeros((1, 12), np.float32)
    f[0, 8] = 100.0
    tau = to.stance_torques(J, f)
    assert tau[0, 2, 1] == -25.0 and np.count_nonzero(tau) == 1
