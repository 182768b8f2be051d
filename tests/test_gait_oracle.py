"""CPU: the gait-clock oracle (oracle/gait_oracle.py) against the fixtures generated from the reference's own
PeriodicGaitGenerator (oracle/make_golden_gait.py), and the host-side dt expansion of the drop-in."""
import numpy as np

from oracle import gait_oracle as go
from quadruped_pympc_tamols_b200.helpers import periodic_gait_generator as pgg
from tests.util import GOLDEN


def test_oracle_matches_reference_fixtures():
    g = np.load(GOLDEN + "/gait_sequences.npz")
    assert len(g["seq"]) == 42 and set(g["gait"]) == set(range(7))
    for k in range(len(g["seq"])):
        dts = [d for d in g["dts"][k] if d > 0]
        lengths = [int(n) for n in g["lengths"][k] if n > 0]
        H = int(g["horizon"][k])
        assert np.array_equal(go.PHASE_OFFSETS[int(g["gait"][k])], g["offset"][k])
        seq = go.contact_sequences(g["phase"][k][None], g["init"][k][None], g["offset"][k][None], g["freq"][k],
                                   g["duty"][k], go.expand_dts(dts, lengths, H))[0]
        assert np.array_equal(seq, g["seq"][k])
        ph, ini = g["phase"][k][None].copy(), g["init"][k][None].copy()
        for j in range(8):
            assert np.array_equal(go.run(ph, ini, g["offset"][k][None], g["duty"][k], g["run_dt"][k], g["freq"][k])[0],
                                  g["run_contacts"][k][j])
        assert np.array_equal(ph[0], g["run_phase"][k]) and np.array_equal(ini[0], g["run_init"][k])


def test_dt_expansion_rule():
    """periodic_gait_generator.py:104-110: the list index moves at most once per column."""
    assert np.allclose(go.expand_dts([0.01, 0.02], [3, 12], 6), [0, 0.01, 0.01, 0.02, 0.02, 0.02])
    assert np.allclose(pgg.expand_dts([0.01, 0.02], [3, 12], 6)[:6], go.expand_dts([0.01, 0.02], [3, 12], 6))
    assert np.allclose(go.expand_dts([0.02], [12], 4), [0, 0.02, 0.02, 0.02])
    assert pgg.GaitType.FULL_STANCE.value == go.FULL_STANCE
