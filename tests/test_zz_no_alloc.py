"""GPU: tamols_score_device_static (include/tamols.h) -- the TAMOLS scoring launch with its scratch (arrival tickets +
per-tile partials) in the library's file-scope `static __device__` array, for callers that cannot call cudaMalloc.

Checked against the existing entry point tamols_score_device through the C ABI on the same device buffers.  Both run
the same kernel, so the bar is the one the TAMOLS tests use for two runs of the CUDA path: records and per-candidate
scores identical (tests/test_tamols_parity_gpu.py), on two calls in a row that share the array, with a call of a
different shape in between so that the second one finds foreign contents in it.

This file sorts last on purpose (a failure here hides nothing after it) and passes
both alone in a fresh process and after the other tests (the array is per process and per GPU, the tickets are cleared
by every call).
"""
import numpy as np
import pytest

import synthetic as sy

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def tc():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    from quadruped_pympc_tamols_b200 import _tamols_cabi
    return _tamols_cabi


class Scene:
    """Device buffers of one sweep scene, shared by the two entry points."""

    def __init__(self, tc, robot, kind, window, map_size, robots, seed, dtype="f64"):
        import torch
        rng = np.random.default_rng(seed)
        sc = sy.sweep_scene(robot, rng, kind, window=window, map_rows=map_size, map_cols=map_size, res=0.02, seed=seed,
                            robots=robots)
        self.tc, self.jobs = tc, sc["jobs"]
        self.p = tc.make_params(sy.tamols_params_for(robot), robot)
        self.M, self.R, self.C = sc["heights"].shape
        self.dtype = tc.F64 if dtype == "f64" else tc.F32
        self.d_h = torch.from_numpy(sc["heights"].astype(np.float64 if dtype == "f64" else np.float32)).cuda()
        self.d_jobs = torch.from_numpy(np.frombuffer(self.jobs.tobytes(), np.uint8).copy()).cuda()
        self.n = len(self.jobs)
        self.max_nr, self.max_nc = int(self.jobs["nr"].max()), int(self.jobs["nc"].max())
        self.stride = self.max_nr * self.max_nc

    def run(self, h, static):
        import torch
        d_res = torch.zeros(self.n * self.tc.RESULT_DTYPE.itemsize, dtype=torch.uint8, device="cuda")
        d_scores = torch.full((self.n, self.stride), float("nan"), dtype=torch.float64, device="cuda")
        torch.cuda.synchronize()
        f = h.score_device_static if static else h.score_device
        f(self.d_h.data_ptr(), self.dtype, self.M, self.R, self.C, self.C, self.d_jobs.data_ptr(), self.n, self.max_nr,
          self.max_nc, d_res.data_ptr(), d_scores=d_scores.data_ptr(), scores_stride=self.stride)
        torch.cuda.synchronize()  # device-wide: also covers the handle's own (non-blocking) stream
        return np.frombuffer(d_res.cpu().numpy().tobytes(), self.tc.RESULT_DTYPE).copy(), d_scores.cpu().numpy()


def same(a, b):
    return a[0].tobytes() == b[0].tobytes() and np.array_equal(a[1], b[1], equal_nan=True)


@pytest.mark.parametrize("tile_rows", ["auto", "8", "32"])
def test_static_scratch_equals_allocating_entry_point(tc, tile_rows, monkeypatch):
    if tile_rows == "auto":
        monkeypatch.delenv("TAMOLS_TILE_ROWS", raising=False)
    else:
        monkeypatch.setenv("TAMOLS_TILE_ROWS", tile_rows)  # read once per handle
    big = Scene(tc, "go2", "stepping_stones_tall", window=96, map_size=512, robots=8, seed=3)   # 32 jobs, many tiles each
    small = Scene(tc, "aliengo", "noise", window=40, map_size=128, robots=3, seed=5, dtype="f32")  # 12 jobs, fewer tiles
    h_ref, h = tc.Handle(big.p, max_jobs=big.n), tc.Handle(big.p, max_jobs=big.n)
    want_big = big.run(h_ref, static=False)
    assert np.isfinite(want_big[1]).any() and (want_big[0]["best_index"] >= 0).any()
    # two calls in a row through the static array
    first, second = big.run(h, static=True), big.run(h, static=True)
    assert same(first, want_big) and same(second, want_big)
    # a call of another shape, dtype and parameter set leaves foreign partials behind; then the big one again
    h_ref.set_params(small.p)
    h.set_params(small.p)
    want_small = small.run(h_ref, static=False)
    assert same(small.run(h, static=True), want_small)
    h.set_params(big.p)
    assert same(big.run(h, static=True), want_big)
    # the array belongs to the library, not to the handle: a second handle continues on it
    h2 = tc.Handle(small.p, max_jobs=1)  # max_jobs bounds the handle's own buffers only
    assert same(small.run(h2, static=True), want_small)
    assert h.launch_count == 4 and h2.launch_count == 1
    for x in (h_ref, h, h2):
        x.close()


def test_static_entry_point_allocates_nothing(tc):
    """Device memory in use does not move across the call, on a FRESH handle whose own partial buffer is still empty
    (the allocating entry point grows it on that same input)."""
    import torch
    sc = Scene(tc, "go2", "stairs", window=96, map_size=512, robots=8, seed=7)
    warm = tc.Handle(sc.p, max_jobs=sc.n)
    want = sc.run(warm, static=False)   # also loads the kernels (lazy module loading allocates on first use)
    sc.run(warm, static=True)
    h = tc.Handle(sc.p, max_jobs=sc.n)
    d_res = torch.zeros(sc.n * tc.RESULT_DTYPE.itemsize, dtype=torch.uint8, device="cuda")
    d_scores = torch.zeros((sc.n, sc.stride), dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    h.score_device_static(sc.d_h.data_ptr(), sc.dtype, sc.M, sc.R, sc.C, sc.C, sc.d_jobs.data_ptr(), sc.n, sc.max_nr,
                          sc.max_nc, d_res.data_ptr(), d_scores=d_scores.data_ptr(), scores_stride=sc.stride)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0, (free0, free1)
    got = np.frombuffer(d_res.cpu().numpy().tobytes(), tc.RESULT_DTYPE).copy(), d_scores.cpu().numpy()
    assert same(got, want)
    warm.close()
    h.close()


def test_static_entry_point_rejects_sizes_beyond_the_array(tc):
    sc = Scene(tc, "go2", "flat", window=32, map_size=512, robots=1, seed=1)
    h = tc.Handle(sc.p, max_jobs=sc.n)
    args = (sc.d_h.data_ptr(), sc.dtype, sc.M, sc.R, sc.C, sc.C, sc.d_jobs.data_ptr())
    import torch
    d_res = torch.zeros(sc.n * tc.RESULT_DTYPE.itemsize, dtype=torch.uint8, device="cuda")
    with pytest.raises(tc.TamolsError, match="TAMOLS_STATIC_MAX_JOBS"):
        h.score_device_static(*args, tc.STATIC_MAX_JOBS + 1, 32, 32, d_res.data_ptr())
    # 4096 jobs x (256 / 32) x (256 / 32) tiles = 262,144 partials > 65,536
    with pytest.raises(tc.TamolsError, match="TAMOLS_STATIC_MAX_PARTIALS"):
        h.score_device_static(*args, tc.STATIC_MAX_JOBS, 256, 256, d_res.data_ptr())
    assert h.launch_count == 0
    # and the handle is still usable afterwards
    want = sc.run(h, static=False)
    assert same(sc.run(h, static=True), want)
    h.close()
