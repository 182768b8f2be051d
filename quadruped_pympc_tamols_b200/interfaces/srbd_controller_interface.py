"""Drop-in for the SAMPLING branch of quadruped_pympc/interfaces/srbd_controller_interface.py
(class SRBDControllerInterface, :7-243) -- SURVEY.md section 8(f) row N1.

Upstream, `compute_control` runs `num_sampling_iterations` controller calls in a Python loop (:131-170), each one a
host <-> device round trip: with_newkey, jitted_compute_control, best_control_parameters (and the CEM sigma) handed
back to the next iteration through the host.  Here the loop is ONE call into the C ABI
(`qpmpc_solve_iterated`): the iterations are launched back to back on the controller's stream, the mean and sigma
stay on the device in between, and only the last iteration's results return.

Everything the branch does around the loop is kept as upstream: current contact from column 0 of the contact
sequence, `prepare_state_and_reference` with the previous MPC contact (lift-off zeroing, optional shift_solution),
footholds taken from the reference state, GRFs masked by the current contact, the 7-tuple that is returned.  The
gradient-based controller types (acados) are out of scope: constructing this class with such a
type raises, pointing at the reference's own interface.
"""
import numpy as np

from .._config import resolve_config
from ..helpers.legs_attr import LegsAttr


class SRBDControllerInterface:
    """Interface for the sampling controller on one B200 (same constructor state and `compute_control` signature as
    upstream; `config` / `device_index` are optional extras for processes that hold several controllers)."""

    def __init__(self, config=None, device_index=0):
        cfg = resolve_config(config)
        self._config = cfg
        self.type = cfg.mpc_params["type"]
        self.mpc_dt = cfg.mpc_params["dt"]
        self.horizon = cfg.mpc_params["horizon"]
        self.optimize_step_freq = cfg.mpc_params["optimize_step_freq"]
        self.step_freq_available = cfg.mpc_params["step_freq_available"]
        self.previous_contact_mpc = np.array([1, 1, 1, 1])  # :19

        if self.type != "sampling":
            raise NotImplementedError(
                f"mpc_params['type'] = {self.type!r}: only the sampling controller runs on this path; use "
                "quadruped_pympc.interfaces.srbd_controller_interface.SRBDControllerInterface for the acados types")
        if self.optimize_step_freq:  # :77-83
            from ..controllers.sampling.centroidal_nmpc_jax_gait_adaptive import Sampling_MPC
        else:
            from ..controllers.sampling.centroidal_nmpc_jax import Sampling_MPC
        self.controller = Sampling_MPC(config=cfg, device_index=device_index)

    def compute_control(self, state_current, ref_state, contact_sequence, inertia, pgg_phase_signal, pgg_step_freq,
                        optimize_swing, external_wrenches=np.zeros((6,))):
        """srbd_controller_interface.py:84-243 (sampling branch :117-183, common tail :224-243).

        Returns (nmpc_GRFs, nmpc_footholds, nmpc_joints_pos, nmpc_joints_vel, nmpc_joints_acc, best_sample_freq,
        nmpc_predicted_state) with LegsAttr GRFs / footholds and None for the joint entries, as upstream."""
        contact_sequence = np.asarray(contact_sequence)
        current_contact = np.array([contact_sequence[0][0], contact_sequence[1][0], contact_sequence[2][0],
                                    contact_sequence[3][0]])
        state, reference = self.controller.prepare_state_and_reference(state_current, ref_state, current_contact,
                                                                       self.previous_contact_mpc)
        self.previous_contact_mpc = current_contact

        c = self.controller
        if c.sampling_method == "cem_mppi":
            if self.optimize_step_freq:
                # upstream passes six arguments to a nine-argument jitted_compute_control here (:140-147 against
                # ..._gait_adaptive.py:964) and dies with TypeError; say so instead of guessing the missing three
                raise TypeError("compute_control_cem_mppi() missing 3 required positional arguments: 'timing', "
                                "'nominal_step_frequency', and 'optimize_swing' (cem_mppi with optimize_step_freq "
                                "is not callable through the upstream interface either)")
            out = c.compute_control_iterated(state, reference, contact_sequence,
                                             sigma0=self._config.mpc_params["sigma_cem_mppi"])
        else:
            out = c.compute_control_iterated(state, reference, contact_sequence, timing=pgg_phase_signal,
                                             nominal_step_frequency=pgg_step_freq, optimize_swing=optimize_swing)
        nmpc_GRFs, _, nmpc_predicted_state, _, _, best_sample_freq = out[:6]
        self.last_best_cost = float(out[4])
        self.last_costs = out[6]

        nmpc_footholds = LegsAttr(FL=ref_state["ref_foot_FL"][0], FR=ref_state["ref_foot_FR"][0],
                                  RL=ref_state["ref_foot_RL"][0], RR=ref_state["ref_foot_RR"][0])  # :172-177
        nmpc_GRFs = np.array(nmpc_GRFs)
        nmpc_GRFs = LegsAttr(FL=nmpc_GRFs[0:3] * current_contact[0], FR=nmpc_GRFs[3:6] * current_contact[1],
                             RL=nmpc_GRFs[6:9] * current_contact[2], RR=nmpc_GRFs[9:12] * current_contact[3])
        return nmpc_GRFs, nmpc_footholds, None, None, None, best_sample_freq, nmpc_predicted_state

    def compute_RTI(self):
        """Real-time-iteration phases exist only for the acados controllers (:245-260)."""
        raise NotImplementedError("compute_RTI applies to the gradient-based controllers, not to sampling")

    def close(self):
        self.controller.close()
