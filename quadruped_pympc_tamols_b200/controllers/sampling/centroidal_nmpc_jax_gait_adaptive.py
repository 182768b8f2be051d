"""Drop-in for quadruped_pympc/controllers/sampling/centroidal_nmpc_jax_gait_adaptive.py (class Sampling_MPC), the
variant selected when mpc_params['optimize_step_freq'] is True (interfaces/srbd_controller_interface.py:77-81).

Every rollout draws its own step frequency, builds its own trot contact sequence with the jax gait generator's rule
(helpers/periodic_gait_generator_jax.py:68-89,136-151; duty factor 0.65, ..._gait_adaptive.py:179), and pays
100 (f - 1.3)^2 on top of the stage costs (:500).  All of that happens inside the CUDA rollout kernel; this file only
changes the argument plumbing relative to the fixed-gait class.
"""
import numpy as np

from .centroidal_nmpc_jax import Sampling_MPC as _FixedGaitSamplingMPC


class Sampling_MPC(_FixedGaitSamplingMPC):
    _gait_adaptive = True

    def _timing(self, timing):
        return np.asarray(timing, np.float32).reshape(4)

    def compute_control_cem_mppi(self, state, reference, contact_sequence, best_control_parameters, key, sigma, timing,
                                 nominal_step_frequency, optimize_swing):
        """..._gait_adaptive.py:964-1131.  Nine positional arguments as upstream; the interface passes six
        (srbd_controller_interface.py:140-147), which raises TypeError there and here alike."""
        r = self._solve(state, reference, contact_sequence, best_control_parameters, key, sigma=sigma,
                        timing=self._timing(timing), nominal_step_frequency=nominal_step_frequency,
                        optimize_swing=optimize_swing)
        return (r["grf"], r["footholds"], r["predicted_state"], r["best"], r["best_cost"], r["best_freq"], r["costs"],
                r["sigma"])

    def compute_rollout(self, initial_state, reference, timing, control_parameters, step_frequency):
        """..._gait_adaptive.py:326-501: cost of one rollout with its own step frequency"""
        return self.vectorized_rollout(initial_state, reference, timing, control_parameters, step_frequency)[0]

    def vectorized_rollout(self, initial_state, reference, timing, control_parameters, step_frequency):
        """jax.vmap(compute_rollout, in_axes=(None, None, None, 0, 0)) (:186): params [n, P], freqs [n] -> costs [n]"""
        unused_contact = np.zeros((4, self.horizon), np.float32)  # the per-sample gait replaces it
        return self._device_costs(initial_state, reference, control_parameters, unused_contact,
                                  timing=self._timing(timing), freqs=step_frequency)

    jit_vectorized_rollout = vectorized_rollout
