"""GPU: the batched gait clock (qpmpc_contact_sequences / qpmpc_gait_run through the C ABI and the
PeriodicGaitGenerator drop-in) against the fixtures produced by the reference's own class and against the CPU oracle
on a large seeded batch.  Bar: contact flags and clock state bit-exact (float64 on both sides)."""
import numpy as np
import pytest

from oracle import gait_oracle as go  # checker only
from quadruped_pympc_tamols_b200 import _cabi
from quadruped_pympc_tamols_b200.helpers.periodic_gait_generator import GaitType, PeriodicGaitGenerator
from tests.util import GOLDEN as GOLDEN_DIR

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def handle():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    h = _cabi.Handle(_cabi.make_config(mass=15.0, inertia=np.eye(3), num_parallel_computations=64))
    yield h
    h.close()


def _lists(g, k):
    return [d for d in g["dts"][k] if d > 0], [int(n) for n in g["lengths"][k] if n > 0]


def test_contact_sequences_match_reference_fixtures_as_one_batch(handle):
    """All 42 reference cases (7 gaits, init hold states, uniform and non-uniform dt lists) as ONE batch per dt list."""
    g = np.load(GOLDEN_DIR + "/gait_sequences.npz")
    groups = {}
    for k in range(len(g["seq"])):
        groups.setdefault((tuple(g["dts"][k]), tuple(g["lengths"][k])), []).append(k)
    for _, ks in groups.items():
        ks = np.array(ks)
        H = int(g["horizon"][ks[0]])
        gen = PeriodicGaitGenerator(g["duty"][ks], g["freq"][ks], g["gait"][ks], H, num_robots=len(ks), handle=handle)
        assert np.array_equal(gen.phase_offset, g["offset"][ks])
        gen.set_phase_signal(g["phase"][ks], g["init"][ks])
        dts, lengths = _lists(g, ks[0])
        seq = gen.compute_contact_sequence(dts, lengths)
        assert seq.dtype.is_floating_point and tuple(seq.shape) == (len(ks), 4, H)
        assert np.array_equal(seq.cpu().numpy(), g["seq"][ks])
        assert np.array_equal(gen.phase_signal, g["phase"][ks]) and np.array_equal(gen.init, g["init"][ks])  # :112


def test_run_matches_reference_fixtures(handle):
    g = np.load(GOLDEN_DIR + "/gait_sequences.npz")
    for rdt in np.unique(g["run_dt"]):
        ks = np.nonzero(g["run_dt"] == rdt)[0]
        gen = PeriodicGaitGenerator(g["duty"][ks], g["freq"][ks], g["gait"][ks], 12, num_robots=len(ks), handle=handle)
        gen.set_phase_signal(g["phase"][ks], g["init"][ks])
        for j in range(8):
            c = gen.run(float(rdt), g["freq"][ks]).cpu().numpy()
            assert np.array_equal(c, g["run_contacts"][ks, j])
        assert np.array_equal(gen.phase_signal, g["run_phase"][ks]) and np.array_equal(gen.init, g["run_init"][ks])


def test_large_batch_matches_oracle_and_feeds_the_solver_layout(handle):
    """8192 clocks with random gaits / phases / frequencies: bit-exact against the oracle; full-stance robots all ones;
    a long run() history stays bit-exact too (the state is carried on the device)."""
    rng = np.random.default_rng(3)
    R, H = 8192, 12
    gait = rng.integers(0, 8, R)
    duty, freq = rng.uniform(0.4, 0.9, R), rng.uniform(0.5, 2.6, R)
    gen = PeriodicGaitGenerator(duty, freq, gait, H, num_robots=R, handle=handle)
    phase, init = rng.uniform(0, 1, (R, 4)), rng.random((R, 4)) < 0.2
    gen.set_phase_signal(phase, init)
    dts, lengths = [0.01, 0.02], [2, H]
    seq = gen.compute_contact_sequence(dts, lengths).cpu().numpy()
    full = gait == GaitType.FULL_STANCE.value
    want = go.contact_sequences(phase, init, gen.phase_offset, freq, duty, go.expand_dts(dts, lengths, H), full)
    assert np.array_equal(seq, want)
    assert np.all(seq[full] == 1.0) and seq.dtype == np.float32
    # full-stance clocks were reset by compute_contact_sequence (:97), the others kept
    ph, ini = gen.phase_signal, gen.init
    assert np.array_equal(ph[~full], phase[~full]) and np.array_equal(ph[full], gen.phase_offset[full])
    o_ph, o_in = ph.copy(), ini.copy()
    for _ in range(50):
        c = gen.run(0.002).cpu().numpy()
        oc = go.run(o_ph, o_in, gen.phase_offset, duty, 0.002, freq)
        assert np.array_equal(c, oc)
    assert np.array_equal(gen.phase_signal, o_ph) and np.array_equal(gen.init, o_in)


def test_bad_arguments_fail_loudly(handle):
    with pytest.raises(_cabi.QpmpcError):
        PeriodicGaitGenerator(0.6, 1.4, 0, 12)  # no handle: there is no CPU path
    b = _cabi.QpmpcGaitBatch(num_robots=1, horizon=12)
    with pytest.raises(_cabi.QpmpcError, match="missing device pointer"):
        handle.contact_sequences(b)
