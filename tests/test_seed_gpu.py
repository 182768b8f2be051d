"""GPU: seed footholds and TAMOLS job records for a batch of robots (tamols_seed_jobs through the C ABI
and the FootholdReferenceGenerator drop-in) against the fixture of the reference's own class, against the oracle, and
as a device-resident pipeline states -> seeds -> jobs -> scores compared with host-built jobs.
Bar: float64, 1e-12 absolute on the seeds (cos/sin of the two math libraries may differ in the last bit); job records
and the scored footholds bit-exact given equal seeds."""
import numpy as np
import pytest

import synthetic as sy
from oracle import seed_oracle as so  # checker only
from quadruped_pympc_tamols_b200 import _tamols_cabi as tc
from quadruped_pympc_tamols_b200.helpers.foothold_reference_generator import FootholdReferenceGenerator
from tests.util import GOLDEN

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def handle():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    h = tc.Handle(tc.make_params(sy.tamols_params_for("go2"), "go2"), max_jobs=4 * 4096)
    yield h
    h.close()


def test_seed_footholds_match_reference_fixture(handle):
    import torch
    g = np.load(GOLDEN + "/seed_footholds.npz")
    T, R = g["ref_feet"].shape[:2]
    gen = FootholdReferenceGenerator(0.4, np.zeros((4, 3)), int(g["hist_len"]), float(g["hip_height"]), num_robots=R,
                                     handle=handle, gravity_constant=float(g["gravity"]))
    gen.com_pos_offset_b = torch.as_tensor(g["com_offset_b"], device="cuda")
    for t in range(T):
        gen.stance_time = torch.as_tensor(g["stance_time"][t], device="cuda")
        gen.lift_off_positions[:, :, 2] = torch.as_tensor(g["lift_off_z"][t], device="cuda")
        out = gen.compute_footholds_reference(g["base_position"][t], g["base_euler"][t], g["base_lin_vel"][t],
                                              g["ref_lin_vel"][t], g["hips"][t], g["com_height"][t])
        torch.cuda.synchronize()
        got = out.cpu().numpy()
        assert np.max(np.abs(got - g["ref_feet"][t])) <= 1e-12, t
        assert np.array_equal(got[:, :, 2], g["lift_off_z"][t])


def test_pipeline_states_to_footholds_stays_on_the_device(handle):
    """2048 robots on one 2048 x 2048 map: the kernel's job records equal the oracle's built from the kernel's seeds
    (bit-exact), and scoring them on the device gives the same footholds as scoring host-built jobs."""
    import torch
    rng = np.random.default_rng(8)
    R, w = 2048, 32
    rows = cols = 2048
    res = 0.04
    origin, row_axis, col_axis = np.array([-40.96, -40.96]), np.array([res, 0.0]), np.array([0.0, res])
    xy = sy.grid_xy(origin, row_axis, col_axis, rows, cols)
    heights = sy.terrain_height("noise", xy[..., 0], xy[..., 1], seed=3)
    pos = np.c_[rng.uniform(-30, 30, (R, 2)), np.full(R, 0.3)]
    eul = np.c_[rng.normal(0, 0.03, (R, 2)), rng.uniform(-3.1, 3.1, R)]
    vel, ref = rng.normal(0, 0.4, (R, 3)), np.c_[rng.normal(0.4, 0.3, R), rng.normal(0, 0.1, R), np.zeros(R)]
    local = np.array([[0.19, 0.13, -0.3], [0.19, -0.13, -0.3], [-0.19, 0.13, -0.3], [-0.19, -0.13, -0.3]])
    c, s = np.cos(eul[:, 2]), np.sin(eul[:, 2])
    rot = np.stack([np.stack([c, -s], -1), np.stack([s, c], -1)], -2)  # [R,2,2] body -> world (yaw only)
    hips = np.concatenate([np.einsum("rij,lj->rli", rot, local[:, :2]) + pos[:, None, :2],
                           np.broadcast_to(local[:, 2] + 0.3 + 0.3, (R, 4))[..., None]], -1)
    feet = hips + rng.normal(0, 0.03, (R, 4, 3))
    feet[..., 2] = 0.0
    contact = (rng.random((R, 4)) < 0.5).astype(np.uint8)
    gen = FootholdReferenceGenerator(rng.uniform(0.2, 0.5, R), np.zeros((R, 4, 3)), 20, 0.3, num_robots=R, handle=handle)
    d_jobs = torch.zeros(4 * R * tc.JOB_DTYPE.itemsize, dtype=torch.uint8, device="cuda")
    hist = [[] for _ in range(R)]
    for step in range(3):  # a few control steps so that the history matters
        seeds = gen.compute_footholds_reference(pos, eul, vel, ref, hips, 0.3, feet_position=feet, current_contact=contact,
                                                global_map=(rows, cols, origin, row_axis, col_axis), window=(w, w),
                                                jobs_out=d_jobs)
        torch.cuda.synchronize()
        want = so.seed_footholds(pos, eul, vel, ref, hips, np.full(R, 0.3), gen.stance_time.cpu().numpy(),
                                 np.zeros((R, 4)), hist, 0.3)
        got = seeds.cpu().numpy()
        assert np.max(np.abs(got - want)) <= 1e-12
    jobs = np.frombuffer(d_jobs.cpu().numpy().tobytes(), tc.JOB_DTYPE)
    want_jobs = so.jobs_for_global_map(tc.JOB_DTYPE, got, hips, feet, contact, pos, vel, (rows, cols), origin, row_axis,
                                       col_axis, (w, w))
    for name in tc.JOB_DTYPE.names:
        assert np.array_equal(jobs[name], want_jobs[name]), name
    assert jobs["r0"].min() >= 0 and (jobs["r0"] + jobs["nr"]).max() <= rows
    # score: device-built jobs on the device against host-built jobs through the host call on a sub-batch
    d_h = torch.as_tensor(heights, dtype=torch.float64, device="cuda").contiguous()
    d_res = torch.zeros(4 * R * tc.RESULT_DTYPE.itemsize, dtype=torch.uint8, device="cuda")
    handle.score_device(d_h.data_ptr(), tc.F64, 1, rows, cols, cols, d_jobs.data_ptr(), 4 * R, w, w, d_res.data_ptr())
    torch.cuda.ExternalStream(handle.stream).synchronize()
    res_dev = np.frombuffer(d_res.cpu().numpy().tobytes(), tc.RESULT_DTYPE)
    sub = slice(0, 4 * 64)
    res_host, _ = handle.score_host(heights[None], want_jobs[sub])
    for name in tc.RESULT_DTYPE.names:
        assert np.array_equal(res_dev[name][sub], res_host[name]), name
    assert (res_dev["best_index"] >= 0).mean() > 0.5  # most legs find a foothold on this terrain


def test_bad_arguments_fail_loudly(handle):
    with pytest.raises(tc.TamolsError):
        FootholdReferenceGenerator(0.4, np.zeros((4, 3)), hip_height=0.3)  # no handle: there is no CPU path
    with pytest.raises(tc.TamolsError, match="missing device pointer"):
        handle.seed_jobs(tc.TamolsSeedBatch(num_robots=1, hist_len=20))
