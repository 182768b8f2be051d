"""GPU: the CUDA TAMOLS path (through the C ABI of include/tamols.h) against (a) the fixtures produced by the
reference's own code and (b) the CPU oracle on larger seeded inputs.

Parity bar (BASELINE.json north_star): selected cell index bit-exact except for documented cost ties under 1e-6.  The
scores themselves are float64 and are required to agree with the reference to 1e-9 and with the oracle (same operation
order, IEEE sqrt/div on both sides) to 1e-12.
"""
import os

import numpy as np
import pytest

import synthetic as sy
from tests.util import check_tamols_against_reference, load_tamols_fixture, tamols_fixtures

pytestmark = pytest.mark.gpu


@pytest.fixture(params=["auto", "8", "32", "window"], autouse=True)
def tile_rows(request, monkeypatch):
    """Both tile shapes of k_tamols_score and the one-block-per-job k_tamols_window at every size (the library otherwise
    picks by launch size; the environment is read once per handle)."""
    monkeypatch.delenv("TAMOLS_TILE_ROWS", raising=False)
    monkeypatch.delenv("TAMOLS_KERNEL", raising=False)
    if request.param == "window":
        monkeypatch.setenv("TAMOLS_KERNEL", "window")
    elif request.param != "auto":
        monkeypatch.setenv("TAMOLS_KERNEL", "tile")
        monkeypatch.setenv("TAMOLS_TILE_ROWS", request.param)
    return request.param


@pytest.fixture(scope="module")
def tc():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    from quadruped_pympc_tamols_b200 import _tamols_cabi
    return _tamols_cabi


def _against_oracle(res, scores, ores, oscores, tol=1e-12):
    fin = np.isfinite(oscores)
    assert np.array_equal(fin, np.isfinite(scores))
    if fin.any():
        assert np.max(np.abs(scores[fin] - oscores[fin])) <= tol
    for k in range(len(ores)):
        a, b = int(res["best_index"][k]), int(ores["best_index"][k])
        if a != b:  # only a sub-1e-6 tie may move the argmin
            assert a >= 0 and b >= 0 and abs(oscores[k, a] - oscores[k, b]) < 1e-6, (k, a, b)
        else:
            assert np.array_equal(res["best_xyz"][k], ores["best_xyz"][k])
        assert int(res["num_finite"][k]) == int(ores["num_finite"][k])


@pytest.mark.parametrize("path", tamols_fixtures(), ids=lambda p: os.path.basename(p)[7:-4])
@pytest.mark.parametrize("explicit_xy", [True, False], ids=["sensor_xy", "grid_xy"])
def test_cuda_matches_reference_fixture(tc, path, explicit_xy):
    p, heights, xy, jobs, g = load_tamols_fixture(path)
    h = tc.Handle(p, max_jobs=len(jobs))
    res, scores = h.score_host(heights, jobs, xy=xy if explicit_xy else None, want_scores=True)
    check_tamols_against_reference(res, scores, g)
    # a second call on the same handle (ticket counters were reset by the kernel) gives the same answer
    res2, scores2 = h.score_host(heights, jobs, xy=xy if explicit_xy else None, want_scores=True)
    assert np.array_equal(res, res2) and np.array_equal(scores, scores2)
    h.close()


def _device_run(tc, p, heights, jobs, dtype="f64", pitch=None, want_scores=True):
    """tamols_score_device on torch-allocated device buffers."""
    import torch
    M, R, Cc = heights.shape
    pitch = pitch or Cc
    host = np.zeros((M, R, pitch), np.float64 if dtype == "f64" else np.float32)
    host[:, :, :Cc] = heights
    d_h = torch.from_numpy(host).cuda()
    d_jobs = torch.from_numpy(np.frombuffer(jobs.tobytes(), np.uint8).copy()).cuda()
    n = len(jobs)
    d_res = torch.zeros(n * tc.RESULT_DTYPE.itemsize, dtype=torch.uint8, device="cuda")
    max_nr, max_nc = int(jobs["nr"].max()), int(jobs["nc"].max())
    stride = max_nr * max_nc
    d_scores = torch.full((n, stride), float("inf"), dtype=torch.float64, device="cuda") if want_scores else None
    h = tc.Handle(p, max_jobs=n)
    torch.cuda.synchronize()
    h.score_device(d_h.data_ptr(), tc.F64 if dtype == "f64" else tc.F32, M, R, Cc, pitch, d_jobs.data_ptr(), n, max_nr,
                   max_nc, d_res.data_ptr(), d_scores=d_scores.data_ptr() if want_scores else None, scores_stride=stride)
    torch.cuda.synchronize()  # device-wide: also covers the handle's own (non-blocking) stream
    launches = h.launch_count
    h.close()
    res = np.frombuffer(d_res.cpu().numpy().tobytes(), tc.RESULT_DTYPE).copy()
    return res, (d_scores.cpu().numpy() if want_scores else None), launches


@pytest.mark.parametrize("kind", ["flat", "stepping_stones_tall", "noise", "stairs"])
@pytest.mark.parametrize("window", [32, 96])
def test_cuda_matches_oracle_on_sweep_windows(tc, kind, window):
    """C4-style: windows cut from one 512 x 512 @ 0.02 m map, 8 robots x 4 legs, device API (TMA staging)."""
    from oracle import tamols_oracle as to
    rng = np.random.default_rng(11)
    sc = sy.sweep_scene("go2", rng, kind, window=window, map_rows=512, map_cols=512, res=0.02, seed=3, robots=8)
    p = tc.make_params(sy.tamols_params_for("go2"), "go2")
    res, scores, launches = _device_run(tc, p, sc["heights"], sc["jobs"])
    assert launches == 1
    ores, oscores = to.score(p, sc["heights"], sc["jobs"], fast=True)
    _against_oracle(res, scores, ores, oscores)
    # size-independent property: the reported winner is the first minimum of the reported scores
    for k in range(len(res)):
        fin = np.isfinite(scores[k])
        assert res["best_index"][k] == (int(np.argmin(scores[k])) if fin.any() else -1)


def test_cuda_multi_tile_merge_matches_reference_run(tc, tile_rows):
    """The 96 x 96-window fixture produced by the reference's own loop (oracle/make_golden_tamols.py multi_tile):
    36 (8-row) or 9 (32-row) tiles per job, so partials, arrival ticket and last-block merge are pinned on the
    reference itself; device API, f64 and TMA / plain staging, and the winners sit in different tiles than tile 0."""
    path = [f for f in tamols_fixtures() if f.endswith("_w96.npz")]
    assert len(path) == 1
    p, heights, xy, jobs, g = load_tamols_fixture(path[0])
    assert int(jobs["nr"].max()) == 96 and int(jobs["nc"].max()) == 96
    th = 32 if tile_rows == "32" else 8  # auto: 4 jobs x 9 tiles < SM count -> 8-row tiles ("window": no tiles, 2 bands)
    for pitch in (128, 129):
        res, scores, launches = _device_run(tc, p, heights, jobs, pitch=pitch)
        assert launches == 1
        check_tamols_against_reference(res, scores, g)
        tiles = {(int(b) // 96 // th, int(b) % 96 // 32) for b in res["best_index"]}
        assert (0, 0) not in tiles and len(tiles) >= 1


@pytest.mark.parametrize("kind", ["noise", "stepping_stones_tall", "flat", "stairs"])
def test_cuda_scores_have_the_oracles_bits(tc, kind):
    """div_exact (csrc/tamols.cu: divisions by 2 delta, 6 delta^2 and 9 through the precomputed reciprocal with two
    residual steps), the zero-guarded sqrt and the squared reachability thresholds are claimed bit-identical to IEEE
    div / sqrt: every per-candidate score must then carry the bits of the strict (-ffp-contract=off) C oracle, which
    divides and takes roots with the host's IEEE instructions.  'noise' exercises the residual steps (non-zero
    numerators everywhere), the flat terrains the zero path, the window border the reachability thresholds."""
    from oracle import tamols_oracle as to
    rng = np.random.default_rng(21)
    sc = sy.sweep_scene("go2", rng, kind, window=64, map_rows=256, map_cols=256, res=0.02, seed=6, robots=6)
    p = tc.make_params(sy.tamols_params_for("go2"), "go2")
    res, scores, _ = _device_run(tc, p, sc["heights"], sc["jobs"])
    ores, oscores = to.score(p, sc["heights"], sc["jobs"], fast=False)
    assert np.isfinite(oscores).sum() > 1000
    assert np.array_equal(scores.view(np.uint64), oscores.view(np.uint64))
    assert np.array_equal(res["best_index"], ores["best_index"]) and np.array_equal(res["num_finite"], ores["num_finite"])
    assert np.array_equal(res["best_score"].view(np.uint64), ores["best_score"].view(np.uint64))


@pytest.mark.parametrize("kind", ["noise", "stepping_stones_tall"])
def test_cuda_result_only_launch_and_reach_box(tc, kind, tile_rows):
    """A launch without the per-candidate score output (what a batched caller runs) takes k_tamols_window's plain pass 1,
    which skips rows and 32-cell segments outside the bounding box of the reach circles: the result records must still
    carry the oracle's bits.  Hips are moved so that the box is clipped by every window border, misses the window
    entirely (no finite candidate: best_index -1) or covers it whole; one robot stands still (no velocity circle)."""
    from oracle import tamols_oracle as to
    rng = np.random.default_rng(33)
    sc = sy.sweep_scene("go2", rng, kind, window=64, map_rows=256, map_cols=256, res=0.02, seed=9, robots=6)
    jobs = sc["jobs"].copy()
    shifts = [(0.0, 0.0), (0.55, 0.0), (-0.55, 0.0), (0.0, 0.6), (0.0, -0.6), (0.5, -0.5), (3.0, 0.0), (0.0, -3.0)]
    for k in range(len(jobs)):
        dx, dy = shifts[k % len(shifts)]
        jobs["hip"][k, 0] += dx
        jobs["hip"][k, 1] += dy
        jobs["seed"][k, 0] += dx
        jobs["seed"][k, 1] += dy
    jobs["vel"][20:] = [2.0, -1.0, 0.0]   # hip + vel * 0.3 s is 0.67 m away: the two circles barely meet or do not
    jobs["flags"][:4] &= ~1                # TAMOLS_FLAG_HAS_VELOCITY off: one circle only
    p = tc.make_params(sy.tamols_params_for("go2"), "go2")
    res, _, launches = _device_run(tc, p, sc["heights"], jobs, want_scores=False)
    assert launches == 1
    ores, _ = to.score(p, sc["heights"], jobs, fast=False)
    assert np.array_equal(res["best_index"], ores["best_index"]) and np.array_equal(res["num_finite"], ores["num_finite"])
    assert np.array_equal(res["best_score"].view(np.uint64), ores["best_score"].view(np.uint64))
    assert np.array_equal(res["best_xyz"], ores["best_xyz"])
    assert (res["best_index"] == -1).any() and (res["num_finite"] > 300).any()  # both ends of the range are present


def test_cuda_plain_staging_equals_tma_staging(tc):
    """Odd row pitch (no tensor map possible) takes the plain-load kernel: identical results."""
    from oracle import tamols_oracle as to
    rng = np.random.default_rng(5)
    sc = sy.sweep_scene("aliengo", rng, "noise", window=40, map_rows=127, map_cols=127, res=0.02, seed=1, robots=3)
    p = tc.make_params(sy.tamols_params_for("aliengo"), "aliengo")
    a = _device_run(tc, p, sc["heights"], sc["jobs"], pitch=127)
    b = _device_run(tc, p, sc["heights"], sc["jobs"], pitch=128)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    ores, oscores = to.score(p, sc["heights"], sc["jobs"], fast=True)
    _against_oracle(a[0], a[1], ores, oscores)


def test_cuda_f32_height_storage(tc):
    """f32 height storage (halves the map bytes): equals the oracle run on the f32-rounded map."""
    from oracle import tamols_oracle as to
    rng = np.random.default_rng(6)
    sc = sy.sweep_scene("go2", rng, "noise", window=64, map_rows=256, map_cols=256, res=0.02, seed=2, robots=4)
    p = tc.make_params(sy.tamols_params_for("go2"), "go2")
    res, scores, _ = _device_run(tc, p, sc["heights"], sc["jobs"], dtype="f32")
    rounded = sc["heights"].astype(np.float32).astype(np.float64)
    ores, oscores = to.score(p, rounded, sc["jobs"], fast=True)
    _against_oracle(res, scores, ores, oscores)


def test_cuda_full_size_window_properties(tc):
    """BASELINE.json configs[3] size: 4 legs x 256 x 256 = 64k candidates on a 2048 x 2048 map.  The oracle still
    finishes in seconds at this size, so compare fully, and check idempotence across launches."""
    from oracle import tamols_oracle as to
    rng = np.random.default_rng(8)
    sc = sy.sweep_scene("go2", rng, "stepping_stones_tall", window=256, map_rows=2048, map_cols=2048, res=0.02, seed=4)
    p = tc.make_params(sy.tamols_params_for("go2"), "go2")
    res, scores, _ = _device_run(tc, p, sc["heights"], sc["jobs"])
    res2, scores2, _ = _device_run(tc, p, sc["heights"], sc["jobs"])
    assert np.array_equal(res, res2) and np.array_equal(scores, scores2)
    ores, oscores = to.score(p, sc["heights"], sc["jobs"], fast=True)
    _against_oracle(res, scores, ores, oscores)


def test_bad_arguments_fail_loudly(tc):
    p = tc.make_params({}, "go2")
    h = tc.Handle(p, max_jobs=2)
    jobs = np.zeros(3, tc.JOB_DTYPE)
    with pytest.raises(tc.TamolsError, match="max_jobs"):
        h.score_host(np.zeros((1, 8, 8)), jobs)
    jobs = np.zeros(1, tc.JOB_DTYPE)
    jobs["nr"], jobs["nc"] = 9, 8
    with pytest.raises(tc.TamolsError, match="window"):
        h.score_host(np.zeros((1, 8, 8)), jobs)
    h.close()


def test_host_call_mapped_results_equal_copied_results(tc, monkeypatch):
    """tamols_score_host without scores publishes its results through a mapped block + done flag; with
    TAMOLS_ZEROCOPY=0 (or when scores are requested) through a D2H copy: identical bytes, also on repeated calls."""
    rng = np.random.default_rng(2)
    scenes = [sy.robot_scene("hyqreal1", rng, "stepping_stones") for _ in range(3)]
    p = tc.make_params(sy.tamols_params_for("hyqreal1"), "hyqreal1")
    monkeypatch.delenv("TAMOLS_ZEROCOPY", raising=False)
    h = tc.Handle(p, max_jobs=4)
    monkeypatch.setenv("TAMOLS_ZEROCOPY", "0")  # read once per handle (tamols_create)
    h0 = tc.Handle(p, max_jobs=4)
    for s in scenes:
        a, _ = h.score_host(s["heights"], s["jobs"], xy=s["xy"])
        b, _ = h.score_host(s["heights"], s["jobs"], xy=s["xy"], want_scores=True)
        c, _ = h0.score_host(s["heights"], s["jobs"], xy=s["xy"])
        assert a.tobytes() == b.tobytes() == c.tobytes()
    h.close()
    h0.close()
