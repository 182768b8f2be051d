"""GPU: dataset rows packed on the device (qpmpc_pack_observables through the C ABI and the drop-in
BatchedMatLogger) against the table the reference's own MatLogger class produced (tests/golden/observable_rows.npz), and
the stance torques against the fixture of the reference statements (tests/golden/stance_torque.npz)."""
import os

import numpy as np
import pytest

torch = pytest.importorskip("torch")
from quadruped_pympc_tamols_b200 import _cabi  # noqa: E402
from quadruped_pympc_tamols_b200.helpers.dataset_logger import BatchedMatLogger  # noqa: E402
from tests.test_observables_oracle import GOLDEN, load_rows  # noqa: E402
from tests.test_sampling_parity_gpu import ROBOTS  # noqa: E402

pytestmark = pytest.mark.gpu


def _handle(R):
    mass, inertia, _, _ = ROBOTS["go2"]
    return _cabi.Handle(_cabi.make_config(mass=mass, inertia=np.array(inertia), sampling_method="mppi",
                                          num_parallel_computations=256, max_robots=R))


def test_rows_equal_the_reference_matlogger(tmp_path):
    g, names, values = load_rows()
    T, R = values[0].shape[:2]
    h = _handle(R)
    ns = len(g["state_names"])
    lg = BatchedMatLogger([str(n) for n in g["state_names"]], tuple(str(n) for n in g["obs_names"]),
                          str(tmp_path / "log.mat"), num_robots=R, handle=h, max_steps=T)
    for t in range(T):
        dev = {n: torch.as_tensor(np.ascontiguousarray(v[t]), device="cuda") for n, v in zip(names, values)}  # dtype kept (GRFs float32)
        lg.record_step(float(g["src_time"][t]), {n: dev[n] for n in names[:ns]}, {n: dev[n] for n in names[ns:]})
    assert lg.header == [str(x) for x in g["header"]]
    assert np.array_equal(lg.table(), g["data"])  # bit-exact: gathers and casts to float64
    files = lg.flush()
    from scipy.io import loadmat
    assert len(files) == R
    m = loadmat(files[1])
    assert np.array_equal(m["data"], g["data"][1]) and [str(x[0]) for x in m["header"].ravel()] == lg.header
    lg2 = BatchedMatLogger([str(n) for n in g["state_names"]], tuple(str(n) for n in g["obs_names"]),
                           str(tmp_path / "batch.npz"), num_robots=R, handle=h, max_steps=1)
    dev = {n: torch.as_tensor(np.ascontiguousarray(v[0]), device="cuda") for n, v in zip(names, values)}
    lg2.record_step(0.0, {n: dev[n] for n in names[:ns]}, {n: dev[n] for n in names[ns:]})
    with pytest.raises(_cabi.QpmpcError):
        lg2.record_step(1.0, {n: dev[n] for n in names[:ns]}, {n: dev[n] for n in names[ns:]})  # max_steps reached
    z = np.load(lg2.flush()[0])
    assert z["data"].shape == (R, 1, len(lg.header))
    h.close()


def test_rows_straight_from_the_result_records():
    """nmpc_GRFs as a strided view into the qpmpc_result records of a batched solve: no copy, no host."""
    R = 7
    h = _handle(R)
    res = torch.as_tensor(np.random.default_rng(0).normal(size=(R, _cabi.RESULT_FLOATS)).astype(np.float32), device="cuda")
    ref = torch.as_tensor(np.random.default_rng(1).normal(size=(R, 24)), device="cuda")  # float64 reference rows
    phase = torch.as_tensor(np.random.default_rng(2).uniform(size=(R, 4)), device="cuda")
    lg = BatchedMatLogger([], ("ref_base_height", "nmpc_GRFs", "phase_signal"), "/tmp/_unused.npz", num_robots=R, handle=h)
    lg.record_step(0.5, {}, {"ref_base_height": ref[:, 2], "nmpc_GRFs": res[:, :12].view(R, 4, 3), "phase_signal": phase})
    tab = lg.table()
    assert lg.header == ["time", "ref_base_height"] + [f"nmpc_GRFs_{i}" for i in range(12)] + [f"phase_signal_{i}" for i in range(4)]
    want = np.concatenate([np.full((R, 1), 0.5), ref.cpu().numpy()[:, 2:3], res.cpu().numpy()[:, :12].astype(np.float64),
                           phase.cpu().numpy()], axis=1)
    assert np.array_equal(tab[:, 0], want)
    h.close()


def test_stance_torques_equal_the_reference_statements():
    g = np.load(os.path.join(GOLDEN, "stance_torque.npz"))
    R = g["grf"].shape[0]
    h = _handle(R)
    res = np.zeros((R, _cabi.RESULT_FLOATS), np.float32)
    res[:, :12] = g["grf"]
    d_res = torch.as_tensor(res, device="cuda")
    d_jac = torch.as_tensor(np.ascontiguousarray(g["feet_jac"]), device="cuda")
    d_tau = torch.zeros((R, 4, 3), dtype=torch.float64, device="cuda")
    h.stance_torques(R, d_res.data_ptr(), d_jac.data_ptr(), d_tau.data_ptr())
    torch.cuda.synchronize()
    scale = np.abs(g["feet_jac"]).transpose(0, 1, 3, 2) @ np.abs(g["grf"].astype(np.float64)).reshape(R, 4, 3, 1)
    assert np.all(np.abs(d_tau.cpu().numpy() - g["tau"]) <= 4 * np.finfo(np.float64).eps * scale[..., 0])
    h.close()


def test_pack_observables_rejects_bad_layouts():
    h = _handle(2)
    lay = _cabi.QpmpcObsLayout()
    lay.num_segments, lay.num_columns = 1, 5
    x = torch.zeros((2, 3), device="cuda")
    lay.seg[0] = _cabi.QpmpcObsSegment(x.data_ptr(), 3, 0, 3)
    out = torch.zeros((2, 4), dtype=torch.float64, device="cuda")
    with pytest.raises(_cabi.QpmpcError):
        h.pack_observables(lay, 2, 0.0, out.data_ptr())  # num_columns != 1 + 3
    lay.num_columns = 4
    lay.seg[0].dtype = 9
    with pytest.raises(_cabi.QpmpcError):
        h.pack_observables(lay, 2, 0.0, out.data_ptr())
    h.close()
