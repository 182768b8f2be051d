"""CPU: the seed-foothold oracle (oracle/seed_oracle.py) against the fixture generated from the reference's own
FootholdReferenceGenerator (oracle/make_golden_seed.py): 12 robots x 30 control steps, deque roll-over, clipping, CoM offset."""
import numpy as np

from oracle import seed_oracle as so
from tests.util import GOLDEN


def test_oracle_matches_reference_fixture():
    g = np.load(GOLDEN + "/seed_footholds.npz")
    T, R = g["ref_feet"].shape[:2]
    assert T > int(g["hist_len"])  # the deque rolls over
    hist = [[] for _ in range(R)]
    for t in range(T):
        out = so.seed_footholds(g["base_position"][t], g["base_euler"][t], g["base_lin_vel"][t], g["ref_lin_vel"][t],
                                g["hips"][t], g["com_height"][t], g["stance_time"][t], g["lift_off_z"][t], hist,
                                float(g["hip_height"]), float(g["hip_offset"]), float(g["gravity"]), int(g["hist_len"]),
                                g["com_offset_b"])
        assert np.max(np.abs(out - g["ref_feet"][t])) <= 1e-14  # float64; scipy's rotation differs in the last bit
    assert all(len(h) == int(g["hist_len"]) for h in hist)
