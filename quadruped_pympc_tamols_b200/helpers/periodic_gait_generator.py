"""Batched drop-in for quadruped_pympc/helpers/periodic_gait_generator.py (class PeriodicGaitGenerator, :8-118) --
SURVEY.md section 8(f) row N2: R gait clocks live on the device and one kernel launch produces the contact sequences of
all robots in the [R, 4, H] float32 layout `qpmpc_batch.d_contact_sequence` consumes, instead of a Python loop over R
generator objects.  Same constructor arguments, methods and semantics as upstream, each scalar argument optionally an
array with one entry per robot; float64 like upstream, contact flags bit-exact (tests/golden/gait_sequences.npz).

No CPU path: the clocks are torch CUDA tensors and every step is the kernel behind qpmpc_contact_sequences /
qpmpc_gait_run (include/qpmpc.h).
"""
import enum

import numpy as np

from .. import _cabi


class GaitType(enum.Enum):  # helpers/quadruped_utils.py:12-22
    TROT = 0
    PACE = 1
    BOUNDING = 2
    CIRCULARCRAWL = 3
    BFDIAGONALCRAWL = 4
    BACKDIAGONALCRAWL = 5
    FRONTDIAGONALCRAWL = 6
    FULL_STANCE = 7


# periodic_gait_generator.py:24-39 (any other value: the last row)
_PHASE_OFFSETS = {0: [0.5, 1.0, 1.0, 0.5], 1: [0.8, 0.3, 0.8, 0.3], 2: [0.5, 0.5, 0.0, 0.0], 3: [0.0, 0.25, 0.75, 0.5],
                  4: [0.0, 0.25, 0.5, 0.75], 5: [0.0, 0.5, 0.75, 0.25], 6: [0.5, 1.0, 0.75, 1.25]}
_DEFAULT_OFFSETS = [0.0, 0.5, 0.5, 0.0]


def expand_dts(contact_sequence_dts, contact_sequence_lenghts, horizon):
    """dt of column i = 1..H-1 with upstream's rule (:104-110: the list index moves at most once per column)."""
    out = np.zeros(_cabi.MAX_H)
    j = 0
    for i in range(1, horizon):
        if i >= contact_sequence_lenghts[j]:
            j += 1
        out[i] = contact_sequence_dts[j]
    return out


class PeriodicGaitGenerator:
    """`num_robots` periodic gait clocks on the GPU of `handle` (a quadruped_pympc_tamols_b200._cabi.Handle)."""

    def __init__(self, duty_factor, step_freq, gait_type, horizon, num_robots=1, handle=None):
        import torch  # device buffers only
        if handle is None:
            raise _cabi.QpmpcError("PeriodicGaitGenerator needs the qpmpc handle whose stream it runs on "
                                   "(there is no CPU path)")
        self._torch, self._h = torch, handle
        self._dev = torch.device("cuda", int(handle.cfg.device))
        self.num_robots = R = int(num_robots)
        self.horizon = int(horizon)
        self.n_contact = 4
        self.duty_factor = np.broadcast_to(np.asarray(duty_factor, np.float64), (R,)).copy()
        self.step_freq = np.broadcast_to(np.asarray(step_freq, np.float64), (R,)).copy()
        self.gait_type = np.broadcast_to(np.asarray([g.value if isinstance(g, GaitType) else g for g in np.atleast_1d(
            np.asarray(gait_type, object))]), (R,)).astype(np.int64).copy()
        self.previous_gait_type = self.gait_type.copy()
        self.start_and_stop_activated = False
        self._d_contact_now = torch.zeros((R, 4), dtype=torch.float32, device=self._dev)
        self._d_contact_seq = torch.zeros((R, 4, self.horizon), dtype=torch.float32, device=self._dev)
        self.reset()

    # ------------------------------------------------------------------ state, :22-45, 78-91
    def _up(self, a, dtype):
        return self._torch.as_tensor(np.array(a), dtype=dtype, device=self._dev)  # np.array: a writable copy

    def reset(self):
        t = self._torch
        self.phase_offset = np.array([_PHASE_OFFSETS.get(int(g), _DEFAULT_OFFSETS) for g in self.gait_type], np.float64)
        self._d_offset = self._up(self.phase_offset, t.float64)
        self._d_phase = self._up(self.phase_offset, t.float64)  # _phase_signal starts at the offsets (:43)
        self._d_init = t.zeros((self.num_robots, 4), dtype=t.uint8, device=self._dev)
        self._d_freq = self._up(self.step_freq, t.float64)
        self._d_duty = self._up(self.duty_factor, t.float64)
        self._d_full = self._up(self.gait_type == GaitType.FULL_STANCE.value, t.uint8)
        self.time_before_switch_freq = 0

    def set_phase_signal(self, phase_signal, init=None):
        t = self._torch
        phase_signal = np.asarray(phase_signal, np.float64).reshape(self.num_robots, 4)
        self._d_phase = self._up(phase_signal, t.float64)
        init = np.zeros((self.num_robots, 4), bool) if init is None else np.asarray(init, bool).reshape(self.num_robots, 4)
        self._d_init = self._up(init, t.uint8)

    @property
    def phase_signal(self):
        self._sync()
        return self._d_phase.cpu().numpy()

    @property
    def init(self):
        self._sync()
        return self._d_init.cpu().numpy().astype(bool)

    def _sync(self):
        self._torch.cuda.current_stream(self._dev).synchronize()
        self._torch.cuda.ExternalStream(self._h.stream, device=self._dev).synchronize()

    def _batch(self, out, step_dts=None):
        b = _cabi.QpmpcGaitBatch(num_robots=self.num_robots, horizon=self.horizon, d_phase=self._d_phase.data_ptr(),
                                 d_init=self._d_init.data_ptr(), d_phase_offset=self._d_offset.data_ptr(),
                                 d_step_freq=self._d_freq.data_ptr(), d_duty_factor=self._d_duty.data_ptr(),
                                 d_full_stance=self._d_full.data_ptr(), d_contact=out.data_ptr())
        if step_dts is not None:
            for i, v in enumerate(step_dts):
                b.step_dts[i] = float(v)
        return b

    # ------------------------------------------------------------------ the clock, :47-76
    def run(self, dt, new_step_freq=None):
        """Advance every clock by dt * new_step_freq and return the contact flags, a [R, 4] float32 CUDA tensor."""
        if new_step_freq is not None:
            self._d_freq = self._up(np.broadcast_to(np.asarray(new_step_freq, np.float64), (self.num_robots,)),
                                    self._torch.float64)
        self._torch.cuda.current_stream(self._dev).synchronize()  # uploads above ran on torch's stream
        self._h.gait_run(self._batch(self._d_contact_now), dt)
        return self._d_contact_now

    # ------------------------------------------------------------------ :93-114
    def compute_contact_sequence(self, contact_sequence_dts, contact_sequence_lenghts):
        """[R, 4, H] float32 CUDA tensor (the layout of qpmpc_batch.d_contact_sequence); the clocks are unchanged.
        FULL_STANCE robots get ones (upstream returns 2H columns of ones there; H are produced, and like upstream
        their clocks are reset)."""
        self._torch.cuda.current_stream(self._dev).synchronize()
        self._h.contact_sequences(self._batch(self._d_contact_seq, expand_dts(contact_sequence_dts,
                                                                              contact_sequence_lenghts, self.horizon)))
        full = self.gait_type == GaitType.FULL_STANCE.value
        if full.any():  # :97 self.reset() for those robots
            self._sync()
            ph = self._d_phase.cpu().numpy()
            ini = self._d_init.cpu().numpy()
            ph[full], ini[full] = self.phase_offset[full], 0
            self.set_phase_signal(ph, ini.astype(bool))
        return self._d_contact_seq

    def set_full_stance(self):
        self.gait_type[:] = GaitType.FULL_STANCE.value
        self.reset()

    def restore_previous_gait(self):
        self.gait_type = self.previous_gait_type.copy()
        self.reset()
