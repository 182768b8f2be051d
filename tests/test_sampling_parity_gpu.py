"""GPU parity tests of the sampling-MPC path: the CUDA kernels, called through the C ABI (include/qpmpc.h), against
(a) the committed fixtures made from the reference's own source and (b) the C oracle on seeded inputs.

Tolerances (BASELINE.json north_star): rollout costs 1e-4 relative (FP32), updated control sequence 1e-3 absolute.
"""
import os
import zlib

import numpy as np
import pytest

torch = pytest.importorskip("torch")

from oracle import srbd_oracle as so  # noqa: E402  (checker only)
from quadruped_pympc_tamols_b200 import _cabi  # noqa: E402
from tests.util import (fixture_kwargs, oracle_tail, record_parity, rel_err, sampling_fixtures,  # noqa: E402
                        softmin_tolerance, softmin_update_f64)

pytestmark = pytest.mark.gpu

COST_RTOL = 1e-4
BEST_ATOL = 1e-3

ROBOTS = {
    "aliengo": (24.637, [[0.2310941359705289, -0.0014987128245817424, -0.021400468992761768],
                         [-0.0014987128245817424, 1.4485084687476608, 0.0004641447134275615],
                         [-0.021400468992761768, 0.0004641447134275615, 1.503217877350808]], 0.33, (0.24, 0.15)),
    "go2": (15.019, [[1.58460467e-01, 1.21660000e-04, -1.55444692e-02],
                     [1.21660000e-04, 4.68645637e-01, -3.12000000e-05],
                     [-1.55444692e-02, -3.12000000e-05, 5.24474661e-01]], 0.27, (0.19, 0.13)),
    "hyqreal1": (108.40, [[4.55031444e+00, 2.75249434e-03, -5.11957307e-01],
                          [2.75249434e-03, 2.02411774e+01, -7.38560592e-04],
                          [-5.11957307e-01, -7.38560592e-04, 2.14269772e+01]], 0.55, (0.44, 0.30)),
}
GAITS = {"trot": ([0.5, 1.0, 1.0, 0.5], 1.4, 0.65), "pace": ([0.8, 0.3, 0.8, 0.3], 1.4, 0.7),
         "bound": ([0.5, 0.5, 0.0, 0.0], 1.8, 0.65)}


@pytest.fixture(params=["auto", "quad", "thread", "quad-2launch", "thread-2launch"])
def kernel(request, monkeypatch):
    """Every rollout kernel at every size: QPMPC_KERNEL pins the choice the library otherwise makes from the launch size
    (quad = 4 lanes per rollout, latency regime; thread = one thread per rollout, throughput regime)."""
    name, _, two = request.param.partition("-")
    if name == "auto":
        monkeypatch.delenv("QPMPC_KERNEL", raising=False)
    else:
        monkeypatch.setenv("QPMPC_KERNEL", name)
    if two:  # QPMPC_FUSED=0: rollout kernel + stand-alone k_finalize instead of the fused last-block finalize
        monkeypatch.setenv("QPMPC_FUSED", "0")
    else:
        monkeypatch.delenv("QPMPC_FUSED", raising=False)
    return request.param


def synth_problem(robot, gait, rng, horizon=12, dt=0.02):
    """SURVEY 8(d) C1 state / reference distribution + a host gait clock contact sequence."""
    _, _, h, (hx, hy) = ROBOTS[robot]
    st = np.zeros(24)
    st[0:3] = [0.0, 0.0, h + rng.normal(0, 0.02)]
    st[3:6] = rng.normal(0, 0.3, 3)
    st[6:9] = rng.normal(0, 0.05, 3)
    st[9:12] = rng.normal(0, 0.3, 3)
    hips = np.array([[hx, hy, 0], [hx, -hy, 0], [-hx, hy, 0], [-hx, -hy, 0]], float)
    st[12:24] = (hips + np.c_[rng.normal(0, 0.03, (4, 2)), np.zeros(4)]).ravel()
    ref = np.zeros(24)
    ref[2] = h
    ref[3] = rng.uniform(-0.5, 0.5)
    ref[12:24] = st[12:24]
    offs, freq, duty = GAITS[gait]
    t = (np.array(offs) + rng.uniform()) % 1.0
    timing = t.copy()
    seq = np.zeros((4, horizon))
    for n in range(horizon):
        seq[:, n] = (t < duty).astype(float)
        t = (t + dt * freq) % 1.0
    return st, ref, seq, timing, freq


def dev(a):
    return torch.as_tensor(np.ascontiguousarray(a, np.float32), device="cuda")


def compare(r, o, method, tag="", best_atol=BEST_ATOL, eps=None, best_in=None, tail=None):
    """Every output of one solve against the oracle.  With `eps` (the noise both sides consumed) the update is checked
    twice: tightly (1e-4) against the float64 softmin of the CUDA path's OWN costs, which pins the update arithmetic,
    and against the oracle's update with the conditioning-aware tolerance of tests/util.py:softmin_tolerance."""
    assert rel_err(r["costs"], o["costs"]) < COST_RTOL, tag
    if eps is not None and method != "random_sampling":
        assert np.max(np.abs(r["best"] - softmin_update_f64(best_in, r["costs"], eps))) < 1e-4, tag
        best_atol = softmin_tolerance(r["costs"], o["costs"], eps, base=best_atol)
    assert r["best_index"] == o["best_index"] or abs(o["costs"][r["best_index"]] - o["best_cost"]) <= COST_RTOL * abs(o["best_cost"]), tag
    assert abs(float(r["best_cost"]) - float(o["best_cost"])) <= COST_RTOL * abs(float(o["best_cost"])), tag
    assert np.max(np.abs(r["best"] - o["best"])) < best_atol, tag
    assert np.max(np.abs(r["grf"] - o["grf"])) < best_atol * 10, tag
    # the prediction inherits the update's conditioning (it is one Euler step under the first GRFs): 1e-4 while the
    # plain 1e-3 bar applies to the update, scaled with the widened tolerance otherwise; `tail` (when given) is the
    # oracle's GRF / prediction evaluated at the CUDA path's OWN mean and pins that arithmetic tightly
    assert rel_err(r["predicted_state"], o["predicted_state"], 1e-2) < 1e-4 * best_atol / BEST_ATOL, tag
    if tail is not None:
        assert np.max(np.abs(r["grf"] - tail[0])) < 1e-4 * max(1.0, float(np.max(np.abs(tail[0])))), tag
        assert rel_err(r["predicted_state"], tail[1], 1e-2) < 1e-5, tag
    assert abs(float(r["best_freq"]) - float(o["best_freq"])) < 1e-6, tag
    assert np.all(r["footholds"] == 0)
    if method == "cem_mppi":
        assert rel_err(r["sigma"], o["sigma"]) < 1e-3, tag


@pytest.mark.parametrize("path", sampling_fixtures(), ids=lambda p: os.path.basename(p)[9:-4])
def test_cuda_matches_reference_fixture(path, kernel):
    g = np.load(path)
    kw = fixture_kwargs(g)
    method = str(g["method"])
    n = int(g["n"])
    h = _cabi.Handle(_cabi.make_config(sampling_method=method, num_parallel_computations=n,
                                       step_freq_available=tuple(g["step_freq_available"]), **kw))
    best = g["best_in"]
    for it in range(int(g["iterations"])):
        t = f"it{it}_"
        eps = dev(g[t + "eps"])
        freqs = dev(g[t + "freqs"]) if t + "freqs" in g else None
        r = h.solve(g["state"], g["reference"], g["contact_sequence"], best, seed=0, offset=it, timing=g["timing"],
                    nominal_freq=float(g["nominal_freq"]), optimize_swing=int(g["optimize_swing"]),
                    sigma_in=g[t + "sigma_in"] if method == "cem_mppi" else None, d_eps=eps.data_ptr(),
                    d_freqs=freqs.data_ptr() if freqs is not None else None)
        assert rel_err(r["costs"], g[t + "costs"]) < COST_RTOL
        assert np.max(np.abs(r["best"] - g[t + "best_out"])) < BEST_ATOL
        assert np.max(np.abs(r["grf"] - g[t + "grf"])) < BEST_ATOL * 10
        assert rel_err(r["predicted_state"], g[t + "predicted_state"], 1e-2) < 1e-4
        assert abs(float(r["best_cost"]) - float(g[t + "best_cost"])) <= COST_RTOL * abs(float(g[t + "best_cost"]))
        if t + "best_freq" in g:
            assert abs(float(r["best_freq"]) - float(g[t + "best_freq"])) < 1e-6
        if method == "cem_mppi":
            assert rel_err(r["sigma"], g[t + "sigma_out"]) < 1e-3
        best = g[t + "best_out"]
    h.close()


CASES = [  # (robot, gait, method, parametrization, gait_adaptive, N)  -- BASELINE.json configs C1..C3 at full size first
    ("aliengo", "trot", "random_sampling", "zero_order", False, 10000),
    ("go2", "bound", "mppi", "cubic_spline", True, 10000),
    ("hyqreal1", "pace", "cem_mppi", "cubic_spline", False, 10000),
    ("aliengo", "trot", "mppi", "zero_order", False, 4097),
    ("aliengo", "trot", "mppi", "linear_spline", False, 1000),
    ("go2", "trot", "random_sampling", "linear_spline", True, 3001),
    ("go2", "trot", "mppi", "zero_order", True, 2048),
    ("go2", "trot", "cem_mppi", "cubic_spline", True, 1500),
    ("hyqreal1", "pace", "cem_mppi", "zero_order", False, 777),
    ("aliengo", "trot", "random_sampling", "cubic_spline", False, 31),
]


@pytest.mark.parametrize("robot,gait,method,param,ga,n", CASES)
def test_philox_solve_matches_oracle_on_same_noise(robot, gait, method, param, ga, n, kernel):
    """In-kernel Philox path: dump the noise the solve draws for (seed, offset), feed it to the oracle, compare
    every output; two iterations so CEM's sigma carry and the MPPI best carry are exercised."""
    rng = np.random.default_rng(zlib.crc32(repr((robot, gait, method, param, ga, n)).encode()))
    mass, inertia, _, _ = ROBOTS[robot]
    kw = dict(mass=mass, inertia=np.array(inertia), parametrization=param, gait_adaptive=ga)
    h = _cabi.Handle(_cabi.make_config(sampling_method=method, num_parallel_computations=n, **kw))
    ocfg = so.make_cfg(**kw)
    P = h.P
    st, ref, seq, timing, nominal = synth_problem(robot, gait, rng)
    best = (2.0 * rng.normal(size=P)).astype(np.float32)
    sigma = np.full(P, 3.0, np.float32)
    d_eps = torch.empty((n, P), dtype=torch.float32, device="cuda")
    d_freqs = torch.empty(n, dtype=torch.float32, device="cuda")
    for it in range(2):
        oswing = 1 - it
        h.generate_noise(42, it, d_eps.data_ptr(), d_freqs.data_ptr(), sigma_in=sigma if method == "cem_mppi" else None,
                         nominal_freq=nominal, optimize_swing=oswing)
        eps = d_eps.cpu().numpy()
        freqs = d_freqs.cpu().numpy() if ga else None
        assert np.all(eps[0] == 0)
        r = h.solve(st, ref, seq, best, seed=42, offset=it, timing=timing, nominal_freq=nominal, optimize_swing=oswing,
                    sigma_in=sigma if method == "cem_mppi" else None)
        o = so.solve(ocfg, method, st, ref, seq, best, eps, timing=timing, freqs=freqs, fast=False)
        if n == 10000:  # the BASELINE configs C1..C3 at full size go into the terminal summary (tests/conftest.py)
            record_parity(f"{robot} {gait} {method} {param}{' ga' if ga else ''} N={n} it{it} [{kernel}:{h.last_kernel}]", method, r, o, eps)
        compare(r, o, method, tag=f"it{it}", eps=eps, best_in=best,
                tail=oracle_tail(so, ocfg, st, ref, seq, r["best"], timing=timing))
        # the supplied-noise path must agree with the in-kernel path bit for bit on the costs.  (The packed pair kernel
        # folds the parameter addition into the Box-Muller product, best + r cos = one FMA: its parameters carry one
        # rounding fewer than best + eps of the dumped noise, <= 1 ulp, so its costs agree to round-off instead.)
        r2 = h.solve(st, ref, seq, best, seed=42, offset=it, timing=timing, nominal_freq=nominal, optimize_swing=oswing,
                     sigma_in=sigma if method == "cem_mppi" else None, d_eps=d_eps.data_ptr(),
                     d_freqs=d_freqs.data_ptr() if ga else None)
        if h.last_kernel == "pair":
            assert rel_err(r["costs"], r2["costs"]) < 5e-6
            assert np.max(np.abs(r["best"] - r2["best"])) < softmin_tolerance(r["costs"], r2["costs"], eps, base=1e-5)
        else:
            assert np.array_equal(r["costs"], r2["costs"])
            assert np.max(np.abs(r["best"] - r2["best"])) < 1e-5
        best = r["best"]
        if method == "cem_mppi":
            sigma = r["sigma"]
    h.close()


@pytest.mark.parametrize("case", ["ties_at_the_threshold", "all_costs_saturated", "few_blocks"])
def test_cem_elite_selection_ties_and_fallback(case, kernel):
    """The CEM elites are argsort(costs)[:10] with a STABLE sort (centroidal_nmpc_jax.py:1075): ties go to the lower
    index.  finalize_robot first bounds the 10th smallest cost by the 10th smallest block minimum and ranks the few
    costs below that bound; it falls back to the radix select when there are fewer blocks than elites or too many
    candidates.  Supplied noise builds the hard cases: the row of rank 6 copied into 40 rows of 40 different blocks (the
    common cost straddles rank 10), every cost saturated at 1e6 (NaN state: the elites are rows 0..9, through the fallback's tie
    bisection), and a solve of 300 rollouts (fewer blocks than elites)."""
    rng = np.random.default_rng({"ties_at_the_threshold": 1, "all_costs_saturated": 2, "few_blocks": 3}[case])
    mass, inertia, _, _ = ROBOTS["hyqreal1"]
    n = 300 if case == "few_blocks" else 10000
    kw = dict(mass=mass, inertia=np.array(inertia), parametrization="cubic_spline", gait_adaptive=False)
    h = _cabi.Handle(_cabi.make_config(sampling_method="cem_mppi", num_parallel_computations=n, **kw))
    ocfg = so.make_cfg(**kw)
    P = h.P
    st, ref, seq, timing, nominal = synth_problem("hyqreal1", "pace", rng)
    best = np.zeros(P, np.float32)
    sigma = np.full(P, 3.0, np.float32)
    eps = (rng.normal(size=(n, P)) * sigma).astype(np.float32)
    eps[0] = 0
    if case == "ties_at_the_threshold":
        # the row of rank 6 copied into 40 more rows, in 40 different rollout blocks: 41 equal costs on ranks 6..46
        o0 = so.solve(ocfg, "cem_mppi", st, ref, seq, best, eps, timing=timing, fast=False)
        src = int(np.argsort(o0["costs"], kind="stable")[6])
        dst = np.arange(40) * 211 + 7
        dst = dst[dst != src]
        eps[dst] = eps[src]
    if case == "all_costs_saturated":
        st = st.copy()
        st[2] = np.nan
    d_eps = dev(eps)
    r = h.solve(st, ref, seq, best, seed=0, offset=0, timing=timing, nominal_freq=nominal, sigma_in=sigma, d_eps=d_eps.data_ptr())
    o = so.solve(ocfg, "cem_mppi", st, ref, seq, best, eps, timing=timing, fast=False)
    order = np.argsort(o["costs"], kind="stable")[:10]
    if case == "ties_at_the_threshold":
        tied = np.flatnonzero(o["costs"] == o["costs"][src])
        below = int(np.sum(o["costs"] < o["costs"][src]))
        assert len(tied) >= 40 and below < 10 < below + len(tied), (len(tied), below)  # the tie group straddles rank 10
        assert np.array_equal(r["costs"][tied], np.full(len(tied), r["costs"][src]))  # ... on the GPU as well
    if case == "all_costs_saturated":
        assert np.all(r["costs"] == 1e6) and np.all(o["costs"] == 1e6) and np.array_equal(order, np.arange(10))
    else:
        assert rel_err(r["costs"], o["costs"]) < COST_RTOL
    # sigma from exactly the oracle's elite rows (float64 restatement of :1075-1081: cov of the rows, sqrt, clip)
    rows = eps[order].astype(np.float64)
    want = np.clip(np.sqrt(rows.var(axis=0, ddof=1) + 1e-8), 0.2, 5.0)
    assert rel_err(r["sigma"], want) < 1e-4
    assert rel_err(r["sigma"], o["sigma"]) < 1e-3
    h.close()


def test_noise_statistics_and_structure():
    mass, inertia, _, _ = ROBOTS["aliengo"]
    n = 9001
    for method in ("random_sampling", "mppi", "cem_mppi"):
        h = _cabi.Handle(_cabi.make_config(mass=mass, inertia=inertia, sampling_method=method, parametrization="zero_order",
                                           num_parallel_computations=n))
        d_eps = torch.empty((n, h.P), dtype=torch.float32, device="cuda")
        sig = np.linspace(0.5, 4.0, h.P).astype(np.float32)
        h.generate_noise(7, 3, d_eps.data_ptr(), sigma_in=sig if method == "cem_mppi" else None)
        e = d_eps.cpu().numpy().astype(np.float64)
        assert np.all(e[0] == 0)
        if method == "random_sampling":  # centroidal_nmpc_jax.py:647-677
            t = n // 3
            a, b, u = e[1:1 + t], e[1 + t:1 + 2 * t], e[1 + 2 * t:]
            assert np.allclose(a / 0.2, b / 3.0, rtol=1e-5, atol=1e-6)  # same key -> same normals upstream
            assert abs(a.std() / 0.2 - 1) < 0.01 and abs(a.mean()) < 0.002
            assert u.min() >= -10 and u.max() < 10 and abs(u.mean()) < 0.05 and abs(u.std() - 20 / np.sqrt(12)) < 0.05
        elif method == "mppi":
            z = e[1:] / 3.0
            assert abs(z.mean()) < 0.005 and abs(z.std() - 1) < 0.005
            assert abs(np.mean(z ** 4) - 3) < 0.05  # kurtosis of a normal
            assert abs(np.corrcoef(z[:, 0], z[:, 1])[0, 1]) < 0.05
        else:
            z = e[1:] / sig
            assert abs(z.std() - 1) < 0.005
        d2 = torch.empty_like(d_eps)
        h.generate_noise(7, 4, d2.data_ptr(), sigma_in=sig if method == "cem_mppi" else None)
        assert not torch.equal(d_eps[1:], d2[1:])  # with_newkey (offset + 1) gives a fresh stream
        h.close()


def test_noise_stream_independence():
    """Stream v3 (csrc/philox.cuh) seeds the blocks 2 m and 2 m + 1 of a row from the two halves of ONE Philox4x32-7
    output and runs a 64-bit generator for twelve words from each: the values of a row must still be independent of
    each other (also across the two blocks of a pair), consecutive rows independent, and the marginal normal."""
    from scipy import stats
    mass, inertia, _, _ = ROBOTS["aliengo"]
    n = 20001
    h = _cabi.Handle(_cabi.make_config(mass=mass, inertia=inertia, sampling_method="mppi", parametrization="zero_order",
                                       num_parallel_computations=n))
    d_eps = torch.empty((n, h.P), dtype=torch.float32, device="cuda")
    h.generate_noise(11, 5, d_eps.data_ptr())
    z = d_eps.cpu().numpy().astype(np.float64)[1:] / 3.0
    h.close()
    m = z.shape[0]
    corr = np.corrcoef(z.T)
    off = np.abs(corr - np.eye(h.P)).max()
    assert off < 5.5 / np.sqrt(m), off  # 10,296 pairs: the expected maximum is 4.3 / sqrt(m)
    lag = max(abs(np.corrcoef(z[:-1, p], z[1:, p])[0, 1]) for p in range(h.P))
    assert lag < 5.0 / np.sqrt(m), lag
    sq = np.abs(np.corrcoef((z ** 2).T) - np.eye(h.P)).max()  # dependence the linear correlation would miss
    assert sq < 5.5 / np.sqrt(m), sq
    d_all = stats.kstest(z.ravel(), "norm").statistic
    assert d_all < 1.95 / np.sqrt(z.size), d_all  # the 0.1 % critical value of the pooled 2.9 M values
    worst = min(stats.kstest(z[:, p], "norm").pvalue for p in range(h.P))
    assert worst > 1e-5, worst  # 144 columns: the smallest p-value of 144 uniform draws is ~0.007


@pytest.mark.parametrize("method", ["mppi", "cem_mppi", "random_sampling"])
@pytest.mark.parametrize("param", ["zero_order", "cubic_spline"])
def test_packed_generator_equals_scalar_generator(monkeypatch, method, param):
    """The two noise generators of csrc/noise.cuh -- scalar (any kernel, finalize, dump) and packed (two rows per thread,
    k_rollout_pair) -- produce the same values: QPMPC_DUMP=pair makes qpmpc_generate_noise overwrite the dump with the
    packed generator's rows.  Bit for bit, except that row 0 may be -0.0 where the scalar one writes +0.0."""
    mass, inertia, _, _ = ROBOTS["aliengo"]
    n = 4097
    outs = []
    for mode in ("", "pair"):
        monkeypatch.setenv("QPMPC_DUMP", mode)
        h = _cabi.Handle(_cabi.make_config(mass=mass, inertia=inertia, sampling_method=method, parametrization=param,
                                           num_parallel_computations=n))
        d = torch.empty((n, h.P), dtype=torch.float32, device="cuda")
        sig = np.linspace(0.5, 4.0, h.P).astype(np.float32)
        h.generate_noise(42, 5, d.data_ptr(), sigma_in=sig if method == "cem_mppi" else None, robot=3)
        outs.append(d.cpu().numpy())
        h.close()
    a, b = outs
    assert np.all(a[0] == 0) and np.all(b[0] == 0)
    assert np.array_equal(a[1:].view(np.uint32), b[1:].view(np.uint32))
    assert np.std(a[1:]) > 0.1


@pytest.mark.parametrize("param,ga", [("zero_order", False), ("linear_spline", True), ("cubic_spline", False), ("cubic_spline", True)])
def test_rollout_costs_matches_oracle(param, ga, kernel):
    """jit_vectorized_rollout: explicit parameter matrix -> raw costs, incl. a diverging (non-finite) rollout."""
    rng = np.random.default_rng(11)
    mass, inertia, _, _ = ROBOTS["go2"]
    kw = dict(mass=mass, inertia=np.array(inertia), parametrization=param, gait_adaptive=ga)
    n = 5000
    h = _cabi.Handle(_cabi.make_config(sampling_method="mppi", num_parallel_computations=n, **kw))
    ocfg = so.make_cfg(**kw)
    st, ref, seq, timing, _ = synth_problem("go2", "trot", rng)
    params = (20 * rng.normal(size=(n, h.P))).astype(np.float32)
    freqs = rng.choice([1.4, 2.0, 2.4], size=n).astype(np.float32)
    d_params, d_freqs = dev(params), dev(freqs)
    d_costs = torch.empty(n, dtype=torch.float32, device="cuda")
    h.rollout_costs(st, ref, seq, d_params.data_ptr(), n, d_costs.data_ptr(), timing=timing,
                    d_freqs=d_freqs.data_ptr() if ga else None)
    c = d_costs.cpu().numpy()
    o = so.rollout_costs(ocfg, st, ref, seq, params, timing=timing, freqs=freqs if ga else None)
    assert rel_err(c, o) < COST_RTOL
    h.close()


def test_known_answers_on_gpu():
    """(i) rest -> fz = mg/4; (ii) swing leg force exactly 0; (iii) flight -> free fall; (vi) MPPI with
    equal costs -> mean of eps; (vii) sample 0 reproduces the previous best."""
    mass, inertia, _, _ = ROBOTS["aliengo"]
    n = 256
    h = _cabi.Handle(_cabi.make_config(mass=mass, inertia=inertia, sampling_method="mppi", parametrization="zero_order",
                                       num_parallel_computations=n))
    st = np.zeros(24)
    st[2] = 0.33
    st[12:24] = np.array([[0.24, 0.15, 0], [0.24, -0.15, 0], [-0.24, 0.15, 0], [-0.24, -0.15, 0]]).ravel()
    zero = dev(np.zeros((n, h.P)))
    r = h.solve(st, st.copy(), np.ones((4, 12)), np.zeros(h.P), 0, 0, d_eps=zero.data_ptr())
    assert np.allclose(r["grf"].reshape(4, 3)[:, 2], mass * 9.81 / 4, rtol=1e-6) and np.all(r["costs"] < 1e-6)
    seq = np.ones((4, 12))
    seq[1] = 0
    r = h.solve(st, st.copy(), seq, np.ones(h.P), 0, 0, d_eps=zero.data_ptr())
    assert np.all(r["grf"][3:6] == 0) and np.allclose(r["grf"].reshape(4, 3)[[0, 2, 3], 2], mass * 9.81 / 3 + 1, rtol=1e-6)
    eps = np.random.default_rng(0).normal(size=(n, h.P)).astype(np.float32)
    eps[0] = 0
    r = h.solve(st, st.copy(), np.zeros((4, 12)), np.zeros(h.P), 0, 0, d_eps=dev(eps).data_ptr())
    assert np.all(r["grf"] == 0) and abs(r["predicted_state"][5] + 9.81 * 0.02) < 1e-6
    assert np.allclose(r["costs"], r["costs"][0]) and np.allclose(r["best"], eps.mean(0), atol=1e-5)
    best = np.random.default_rng(1).normal(size=h.P).astype(np.float32)
    r1 = h.solve(st, st.copy(), np.ones((4, 12)), best, 5, 0)
    r2 = h.solve(st, st.copy(), np.ones((4, 12)), best, 6, 9)
    assert r1["costs"][0] == r2["costs"][0] and not np.array_equal(r1["costs"][1:], r2["costs"][1:])
    h.close()


def test_batched_robots_match_single_solves(kernel):
    """qpmpc_solve_batched: R independent robots in one launch == R single solves with the same Philox subsequence
    (robot r draws from subsequence r) and == the oracle on the dumped noise."""
    rng = np.random.default_rng(3)
    mass, inertia, _, _ = ROBOTS["aliengo"]
    R, n = 5, 1024
    kw = dict(mass=mass, inertia=np.array(inertia), parametrization="zero_order")
    h = _cabi.Handle(_cabi.make_config(sampling_method="mppi", num_parallel_computations=n, max_robots=R, **kw))
    ocfg = so.make_cfg(**kw)
    P = h.P
    probs = [synth_problem("aliengo", "trot", rng) for _ in range(R)]
    bests = (rng.normal(size=(R, P))).astype(np.float32)
    d_state, d_ref = dev(np.stack([p[0] for p in probs])), dev(np.stack([p[1] for p in probs]))
    d_seq, d_best = dev(np.stack([p[2] for p in probs])), dev(bests)
    d_res = torch.zeros((R, _cabi.RESULT_FLOATS), dtype=torch.float32, device="cuda")
    d_bo = torch.zeros((R, P), dtype=torch.float32, device="cuda")
    d_costs = torch.zeros((R, n), dtype=torch.float32, device="cuda")
    b = _cabi.QpmpcBatch(num_robots=R, d_state=d_state.data_ptr(), d_reference=d_ref.data_ptr(),
                         d_contact_sequence=d_seq.data_ptr(), d_best_in=d_best.data_ptr(), d_result=d_res.data_ptr(),
                         d_best_out=d_bo.data_ptr(), d_costs=d_costs.data_ptr())
    h.solve_batched(b, 99, 1)
    torch.cuda.synchronize()
    res, bo, costs = d_res.cpu().numpy(), d_bo.cpu().numpy(), d_costs.cpu().numpy()
    d_eps = torch.empty((n, P), dtype=torch.float32, device="cuda")
    for r in range(R):
        h.generate_noise(99, 1, d_eps.data_ptr(), robot=r)
        eps = d_eps.cpu().numpy()
        o = so.solve(ocfg, "mppi", probs[r][0], probs[r][1], probs[r][2], bests[r], eps)
        assert rel_err(costs[r], o["costs"]) < COST_RTOL
        assert np.max(np.abs(bo[r] - softmin_update_f64(bests[r], costs[r], eps))) < 1e-4
        tol = softmin_tolerance(costs[r], o["costs"], eps, base=BEST_ATOL)
        assert np.max(np.abs(bo[r] - o["best"])) < tol
        assert np.max(np.abs(res[r, :12] - o["grf"])) < tol * 10
        assert rel_err(res[r, 24:48], o["predicted_state"], 1e-2) < 1e-4 * tol / BEST_ATOL
        grf_t, pred_t = oracle_tail(so, ocfg, probs[r][0], probs[r][1], probs[r][2], bo[r])
        assert np.max(np.abs(res[r, :12] - grf_t)) < 1e-4 * max(1.0, float(np.max(np.abs(grf_t))))
        assert rel_err(res[r, 24:48], pred_t, 1e-2) < 1e-5
        assert abs(res[r, 48] - o["best_cost"]) <= COST_RTOL * abs(o["best_cost"])
    assert not np.array_equal(costs[0, 1:], costs[1, 1:])
    h.close()


def test_batched_results_do_not_depend_on_sharding():
    """robot r draws from Philox subsequence robot_offset + r, so 6 robots in one launch equal
    two launches of 3 robots (what two ranks would run) bit for bit."""
    rng = np.random.default_rng(9)
    mass, inertia, _, _ = ROBOTS["go2"]
    R, n = 6, 512
    h = _cabi.Handle(_cabi.make_config(mass=mass, inertia=np.array(inertia), sampling_method="mppi",
                                       parametrization="cubic_spline", num_parallel_computations=n, max_robots=R))
    P = h.P
    probs = [synth_problem("go2", "trot", rng) for _ in range(R)]
    d_state, d_ref = dev(np.stack([p[0] for p in probs])), dev(np.stack([p[1] for p in probs]))
    d_seq, d_best = dev(np.stack([p[2] for p in probs])), dev(rng.normal(size=(R, P)))

    def run(lo, hi):
        k = hi - lo
        d_res = torch.zeros((k, _cabi.RESULT_FLOATS), dtype=torch.float32, device="cuda")
        d_bo = torch.zeros((k, P), dtype=torch.float32, device="cuda")
        d_costs = torch.zeros((k, n), dtype=torch.float32, device="cuda")
        b = _cabi.QpmpcBatch(num_robots=k, robot_offset=lo, d_state=d_state[lo:hi].data_ptr(),
                             d_reference=d_ref[lo:hi].data_ptr(), d_contact_sequence=d_seq[lo:hi].data_ptr(),
                             d_best_in=d_best[lo:hi].data_ptr(), d_result=d_res.data_ptr(), d_best_out=d_bo.data_ptr(),
                             d_costs=d_costs.data_ptr())
        h.solve_batched(b, 5, 2)
        torch.cuda.synchronize()
        return d_res.cpu().numpy(), d_bo.cpu().numpy(), d_costs.cpu().numpy()

    full = run(0, R)
    halves = [run(0, 3), run(3, 6)]
    for k in range(3):
        assert np.array_equal(full[k], np.concatenate([halves[0][k], halves[1][k]]))
    assert not np.array_equal(full[2][0, 1:], full[2][3, 1:])
    h.close()


@pytest.mark.parametrize("method,want_costs", [("mppi", False), ("cem_mppi", True), ("random_sampling", False)])
def test_zero_copy_solve_equals_copy_path(monkeypatch, method, want_costs):
    """qpmpc_solve's default I/O (inputs in the kernel parameter block, results as {word, sequence} pairs in a mapped
    host block), inputs by H2D copy, inputs mapped, and the classic H2D / D2H copies + sync: bit-identical."""
    rng = np.random.default_rng(21)
    mass, inertia, _, _ = ROBOTS["go2"]
    st, ref, seq, timing, nominal = synth_problem("go2", "bound", rng)
    outs = []
    for zc, inputs in (("", ""), ("", "copy"), ("", "mapped"), ("0", "")):
        monkeypatch.setenv("QPMPC_ZEROCOPY", zc)
        monkeypatch.setenv("QPMPC_INPUTS", inputs)
        h = _cabi.Handle(_cabi.make_config(mass=mass, inertia=np.array(inertia), sampling_method=method,
                                           parametrization="cubic_spline", gait_adaptive=(method != "cem_mppi"),
                                           num_parallel_computations=10000))
        best = np.zeros(h.P, np.float32)
        sig = np.full(h.P, 3.0, np.float32) if method == "cem_mppi" else None
        for it in range(3):  # repeated calls reuse the block and bump the sequence number
            r = h.solve(st, ref, seq, best, 3, it, timing=timing, nominal_freq=nominal, optimize_swing=1, sigma_in=sig,
                        want_costs=want_costs)
            best = r["best"].copy()
            if sig is not None:
                sig = r["sigma"].copy()
        outs.append({k: np.array(v) for k, v in r.items() if v is not None})
        h.close()
    for other in outs[1:]:
        assert outs[0].keys() == other.keys()
        for k in outs[0]:
            assert np.array_equal(outs[0][k], other[k]), k


@pytest.mark.parametrize("method", ["mppi", "random_sampling"])
def test_quad_block_sizes_agree(monkeypatch, method):
    """The three quad-kernel block sizes (32, 72, 128 rollouts per block) split the same 10,000 rollouts differently:
    costs, argmin and best cost are bit-identical, the MPPI mean agrees to summation-order round-off."""
    rng = np.random.default_rng(5)
    mass, inertia, _, _ = ROBOTS["go2"]
    st, ref, seq, timing, nominal = synth_problem("go2", "trot", rng)
    outs = []
    for qb in ("32", "72", "128"):
        monkeypatch.setenv("QPMPC_KERNEL", "quad")
        monkeypatch.setenv("QPMPC_QB", qb)
        h = _cabi.Handle(_cabi.make_config(mass=mass, inertia=np.array(inertia), sampling_method=method,
                                           parametrization="linear_spline", gait_adaptive=True,
                                           num_parallel_computations=10000))
        outs.append(h.solve(st, ref, seq, np.zeros(h.P, np.float32), 9, 4, timing=timing, nominal_freq=nominal,
                            optimize_swing=1))
        h.close()
    for o in outs[1:]:
        assert np.array_equal(outs[0]["costs"], o["costs"])
        assert outs[0]["best_index"] == o["best_index"] and outs[0]["best_cost"] == o["best_cost"]
        if method == "random_sampling":
            assert np.array_equal(outs[0]["best"], o["best"])
        else:
            assert np.max(np.abs(outs[0]["best"] - o["best"])) <= 1e-5 * max(1.0, float(np.max(np.abs(o["best"]))))


@pytest.mark.parametrize("method,param,ga,horizon,splines,n", [
    ("mppi", "zero_order", False, 20, 1, 4500),       # P = 240: inputs exceed the kernel parameter block -> H2D copy path
    ("random_sampling", "linear_spline", True, 8, 3, 2000),
    ("cem_mppi", "cubic_spline", False, 16, 4, 3000),  # P = 192
    ("mppi", "cubic_spline", True, 5, 1, 10000),       # short horizon, full-size sample count (one wave of quad blocks)
    # fixed gait with a horizon that is NOT a multiple of 4: the float4 step tables of the thread kernel must stay
    # 16-byte aligned (s_basis sat at s_st + 9 H floats), and the pair kernel takes any horizon
    ("mppi", "linear_spline", False, 10, 2, 3000),
    ("random_sampling", "cubic_spline", False, 5, 1, 2500),
    ("cem_mppi", "cubic_spline", False, 15, 3, 2000),
    ("mppi", "zero_order", False, 10, 1, 4100),
    ("cem_mppi", "zero_order", False, 7, 1, 1500),
])
def test_other_horizons_and_spline_counts(method, param, ga, horizon, splines, n, kernel):
    """mpc_params['horizon'] and ['num_splines'] away from the defaults (12, 2): every kernel, chained over two
    iterations, against the oracle on the very same noise."""
    rng = np.random.default_rng(horizon * 100 + splines)
    mass, inertia, _, _ = ROBOTS["go2"]
    kw = dict(mass=mass, inertia=np.array(inertia), parametrization=param, gait_adaptive=ga, horizon=horizon,
              num_splines=splines)
    h = _cabi.Handle(_cabi.make_config(sampling_method=method, num_parallel_computations=n, **kw))
    ocfg = so.make_cfg(**kw)
    P = h.P
    assert P == {"zero_order": 12 * horizon, "linear_spline": 12 * (splines + 1), "cubic_spline": 48 * splines}[param]
    st, ref, seq, timing, nominal = synth_problem("go2", "trot", rng, horizon=horizon)
    best = rng.normal(size=P).astype(np.float32)
    sigma = np.full(P, 3.0, np.float32)
    d_eps = torch.empty((n, P), dtype=torch.float32, device="cuda")
    d_freqs = torch.empty(n, dtype=torch.float32, device="cuda")
    for it in range(2):
        h.generate_noise(11, it, d_eps.data_ptr(), d_freqs.data_ptr(), sigma_in=sigma if method == "cem_mppi" else None,
                         nominal_freq=nominal, optimize_swing=1)
        eps = d_eps.cpu().numpy()
        freqs = d_freqs.cpu().numpy() if ga else None
        r = h.solve(st, ref, seq, best, seed=11, offset=it, timing=timing, nominal_freq=nominal, optimize_swing=1,
                    sigma_in=sigma if method == "cem_mppi" else None)
        o = so.solve(ocfg, method, st, ref, seq, best, eps, timing=timing, freqs=freqs, fast=False)
        compare(r, o, method, tag=f"H{horizon}S{splines}it{it}", eps=eps, best_in=best)
        best = r["best"]
        if method == "cem_mppi":
            sigma = r["sigma"]
    # the two iterations chained on the device give the same final mean
    best0 = rng.normal(size=P).astype(np.float32)
    a = h.solve(st, ref, seq, best0, 11, 5, timing=timing, nominal_freq=nominal, optimize_swing=1,
                sigma_in=np.full(P, 3.0, np.float32) if method == "cem_mppi" else None)
    b = h.solve(st, ref, seq, a["best"], 11, 6, timing=timing, nominal_freq=nominal, optimize_swing=1, sigma_in=a["sigma"])
    c2 = h.solve(st, ref, seq, best0, 11, 5, timing=timing, nominal_freq=nominal, optimize_swing=1,
                 sigma_in=np.full(P, 3.0, np.float32) if method == "cem_mppi" else None, iterations=2)
    assert np.array_equal(b["best"], c2["best"]) and np.array_equal(b["costs"], c2["costs"])
    h.close()
