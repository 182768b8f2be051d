"""GPU: the batched stance torque mapping (qpmpc_stance_torques through the C ABI) against the
reference's own numpy expression (oracle/torque_oracle.py, interfaces/wb_interface.py:368-371)."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
from oracle import torque_oracle as to  # noqa: E402
from quadruped_pympc_tamols_b200 import _cabi  # noqa: E402
from tests.test_sampling_parity_gpu import ROBOTS, synth_problem  # noqa: E402

pytestmark = pytest.mark.gpu
TAU_RTOL = 1e-13  # float64; numpy's matmul may fuse or reorder the three products of a row


def _dev(a, dt):
    return torch.as_tensor(np.ascontiguousarray(a), dtype=dt, device="cuda")


@pytest.mark.parametrize("R", [1, 5, 1000])
def test_stance_torques_from_result_records(R):
    """Random Jacobians and force records (incl. zeros, huge and negative entries) for R robots."""
    rng = np.random.default_rng(R)
    mass, inertia, _, _ = ROBOTS["go2"]
    h = _cabi.Handle(_cabi.make_config(mass=mass, inertia=np.array(inertia), sampling_method="mppi",
                                       num_parallel_computations=256, max_robots=R))
    res = rng.normal(size=(R, _cabi.RESULT_FLOATS)).astype(np.float32)
    res[:, :12] *= 10.0 ** rng.integers(-3, 4, size=(R, 12))
    res[0, :3] = 0.0
    jac = rng.normal(size=(R, 4, 3, 3))
    d_res, d_jac = _dev(res, torch.float32), _dev(jac, torch.float64)
    d_tau = torch.full((R, 4, 3), np.nan, dtype=torch.float64, device="cuda")
    h.stance_torques(R, d_res.data_ptr(), d_jac.data_ptr(), d_tau.data_ptr())
    torch.cuda.synchronize()
    tau, ref = d_tau.cpu().numpy(), to.stance_torques(jac, res[:, :12])
    scale = np.abs(jac).transpose(0, 1, 3, 2) @ np.abs(res[:, :12].astype(np.float64)).reshape(R, 4, 3, 1)
    assert np.all(np.abs(tau - ref) <= TAU_RTOL * scale[..., 0] + 1e-300)
    assert np.all(tau[0, 0] == 0.0)
    h.close()


def test_solve_then_torques_stays_on_the_device():
    """states -> qpmpc_solve_batched -> qpmpc_stance_torques on one stream, nothing on the host in between."""
    rng = np.random.default_rng(11)
    mass, inertia, _, _ = ROBOTS["aliengo"]
    R, n = 6, 512
    h = _cabi.Handle(_cabi.make_config(mass=mass, inertia=np.array(inertia), sampling_method="mppi",
                                       parametrization="zero_order", num_parallel_computations=n, max_robots=R))
    probs = [synth_problem("aliengo", "trot", rng) for _ in range(R)]
    d_state, d_ref = _dev(np.stack([p[0] for p in probs]), torch.float32), _dev(np.stack([p[1] for p in probs]), torch.float32)
    d_seq = _dev(np.stack([p[2] for p in probs]), torch.float32)
    d_best = torch.zeros((R, h.P), dtype=torch.float32, device="cuda")
    d_res = torch.zeros((R, _cabi.RESULT_FLOATS), dtype=torch.float32, device="cuda")
    d_bo = torch.zeros((R, h.P), dtype=torch.float32, device="cuda")
    jac = rng.normal(size=(R, 4, 3, 3)) * 0.3
    d_jac = _dev(jac, torch.float64)
    d_tau = torch.zeros((R, 4, 3), dtype=torch.float64, device="cuda")
    b = _cabi.QpmpcBatch(num_robots=R, d_state=d_state.data_ptr(), d_reference=d_ref.data_ptr(),
                         d_contact_sequence=d_seq.data_ptr(), d_best_in=d_best.data_ptr(), d_result=d_res.data_ptr(),
                         d_best_out=d_bo.data_ptr())
    h.solve_batched(b, 7, 0)
    h.stance_torques(R, d_res.data_ptr(), d_jac.data_ptr(), d_tau.data_ptr())
    torch.cuda.synchronize()
    grf = d_res.cpu().numpy()[:, :12]
    assert np.any(grf != 0)
    ref = to.stance_torques(jac, grf)
    assert np.allclose(d_tau.cpu().numpy(), ref, rtol=1e-12, atol=1e-12)
    h.close()


def test_stance_torques_rejects_null_pointers():
    mass, inertia, _, _ = ROBOTS["go2"]
    h = _cabi.Handle(_cabi.make_config(mass=mass, inertia=np.array(inertia), sampling_method="mppi",
                                       num_parallel_computations=256))
    with pytest.raises(Exception):
        h.stance_torques(1, 0, 0, 0)
    with pytest.raises(Exception):
        h.stance_torques(0, 1, 1, 1)
    h.close()
