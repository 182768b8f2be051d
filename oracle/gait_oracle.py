"""CPU restatement of the reference's periodic gait clock, for a batch of robots.  TEST INFRASTRUCTURE: only tests/,
__graft_entry__.smoke() and bench.py's CPU legs may import this; the product path is the CUDA kernel behind
qpmpc_contact_sequences / qpmpc_gait_run (include/qpmpc.h).

Follows quadruped_pympc/helpers/periodic_gait_generator.py (paths relative to /root/reference):
  :22-45   phase offsets per GaitType
  :47-76   run(dt, new_step_freq): phase += dt * f, wrap to [0, 1), init hold, duty-factor test
  :93-114  compute_contact_sequence(contact_sequence_dts, contact_sequence_lenghts): column 0 = run(0.0, f), then one
           run(dt_j, f) per further column with the non-uniform dt list, clock state restored afterwards.
Pinned by tests/golden/gait_sequences.npz, produced by oracle/make_golden_gait.py from the reference's own class.
All arithmetic is float64 like the reference's Python floats / numpy arrays.
"""
import numpy as np

# GaitType values, helpers/quadruped_utils.py:12-22
TROT, PACE, BOUNDING, CIRCULARCRAWL, BFDIAGONALCRAWL, BACKDIAGONALCRAWL, FRONTDIAGONALCRAWL, FULL_STANCE = range(8)
PHASE_OFFSETS = {  # periodic_gait_generator.py:24-39 (anything else, FULL_STANCE included: the last row)
    TROT: [0.5, 1.0, 1.0, 0.5], PACE: [0.8, 0.3, 0.8, 0.3], BOUNDING: [0.5, 0.5, 0.0, 0.0],
    CIRCULARCRAWL: [0.0, 0.25, 0.75, 0.5], BFDIAGONALCRAWL: [0.0, 0.25, 0.5, 0.75],
    BACKDIAGONALCRAWL: [0.0, 0.5, 0.75, 0.25], FRONTDIAGONALCRAWL: [0.5, 1.0, 0.75, 1.25],
    FULL_STANCE: [0.0, 0.5, 0.5, 0.0],
}


def expand_dts(contact_sequence_dts, contact_sequence_lenghts, horizon):
    """The dt each column 1..H-1 advances by (:104-110; j moves at most one entry per column); entry 0 is 0.0."""
    out = np.zeros(horizon)
    j = 0
    for i in range(1, horizon):
        if i >= contact_sequence_lenghts[j]:
            j += 1
        out[i] = contact_sequence_dts[j]
    return out


def run(phase, init, phase_offset, duty, dt, step_freq):
    """:47-76 on arrays [R, 4]; phase and init are updated in place, the contact flags [R, 4] are returned."""
    phase += (np.asarray(dt, np.float64) * np.asarray(step_freq, np.float64))[..., None] if np.ndim(step_freq) else dt * step_freq
    np.mod(phase, 1.0, out=phase)
    contact = np.zeros(phase.shape)
    hold = init & (phase <= phase_offset)
    release = init & ~hold
    contact[hold | release] = 1.0
    phase[release] = 0.0
    init &= ~release
    free = ~(hold | release)
    d = np.broadcast_to(np.asarray(duty, np.float64)[..., None] if np.ndim(duty) else duty, phase.shape)
    contact[free & (phase < d)] = 1.0
    return contact


def contact_sequences(phase, init, phase_offset, step_freq, duty, step_dts, full_stance=None):
    """[R, 4, H] contact flags from clocks (phase [R, 4], init [R, 4] bool); the clocks are not modified (:112).
    step_dts [H] = expand_dts(...).  Robots flagged full_stance get ones (:95-98, truncated to H columns)."""
    phase = np.array(phase, np.float64)
    init = np.array(init, bool)
    phase_offset = np.asarray(phase_offset, np.float64)
    R, H = phase.shape[0], len(step_dts)
    out = np.zeros((R, 4, H))
    for n in range(H):
        out[:, :, n] = run(phase, init, phase_offset, duty, float(step_dts[n]) if n else 0.0, step_freq)
    if full_stance is not None:
        out[np.asarray(full_stance, bool)] = 1.0
    return out
