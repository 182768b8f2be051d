"""Drop-in for quadruped_pympc/controllers/sampling/centroidal_nmpc_jax.py (class Sampling_MPC).

Same attribute and method surface, positional argument orders and return tuples as the reference class
(centroidal_nmpc_jax.py:21-1097) as used by interfaces/srbd_controller_interface.py:113-180, with the XLA executable
`jitted_compute_control` replaced by ONE call into the sm_100a library through the C ABI of include/qpmpc.h
(qpmpc_solve: noise, rollouts, costs, update, first-step GRFs and predicted state).  No jax, no CPU fallback: when the
shared library is missing or no sm_100 GPU is present the constructor raises (the reference silently falls back to
jax-CPU, :43-48).

Differences a caller can observe, all forced by the absence of jax: arrays come back as numpy float32 instead of jax
arrays; `master_key` is a (seed, offset) Philox key instead of a threefry key (any object works upstream as long as
`with_newkey` advances it); the random stream is this library's own (Philox4x32-7 seeds + xoroshiro64+ bursts, csrc/philox.cuh), so individual samples differ from jax's while their
distribution is the same (parity is checked by feeding identical noise to both sides, tests/test_sampling_parity_gpu.py).
"""
import sys
import types

import numpy as np

from ... import _cabi
from ..._config import resolve_config

dtype_general = "float32"


class Sampling_MPC:
    """Sampling-based control law (random sampling / MPPI / CEM-MPPI over SRBD rollouts) on one B200."""

    _gait_adaptive = False

    def __init__(self, config=None, device_index=0):
        config = resolve_config(config)
        self._config = config
        mp = config.mpc_params
        self.num_parallel_computations = mp["num_parallel_computations"]
        self.sampling_method = mp["sampling_method"]
        self.control_parametrization = mp["control_parametrization"]
        self.num_sampling_iterations = mp["num_sampling_iterations"]
        self.dt = mp["dt"]
        self.horizon = mp["horizon"]
        self.state_dim = 24
        self.control_dim = 24
        self.reference_dim = self.state_dim
        self.max_sampling_forces_x = 10
        self.max_sampling_forces_y = 10
        self.max_sampling_forces_z = 30
        self.device = f"cuda:{device_index}"

        # parameter count per leg, centroidal_nmpc_jax.py:50-93 (anything that is not a spline name is zero order)
        if self.control_parametrization == "linear_spline":
            self.num_spline = mp["num_splines"]
            self.num_control_parameters_single_leg = (self.num_spline + 1) * 3
            spline = self.compute_linear_spline
        elif self.control_parametrization == "cubic_spline":
            self.num_spline = mp["num_splines"]
            self.num_control_parameters_single_leg = 4 * 3 * self.num_spline
            spline = self.compute_cubic_spline
        else:
            self.num_spline = mp.get("num_splines", 1)
            self.num_control_parameters_single_leg = self.horizon * 3
            spline = self.compute_zero_order_spline
        self.num_control_parameters = self.num_control_parameters_single_leg * 4
        self.spline_fun_FL = self.spline_fun_FR = self.spline_fun_RL = self.spline_fun_RR = spline

        if self.sampling_method == "random_sampling":
            self.compute_control = self.compute_control_random_sampling
            self.sigma_random_sampling = mp["sigma_random_sampling"]
        elif self.sampling_method == "mppi":
            self.compute_control = self.compute_control_mppi
            self.sigma_mppi = mp["sigma_mppi"]
        elif self.sampling_method == "cem_mppi":
            self.compute_control = self.compute_control_cem_mppi
            self.sigma_cem_mppi = np.ones(self.num_control_parameters, dtype=np.float32) * mp["sigma_cem_mppi"]
        else:  # :107-110
            print("Error: sampling method not recognized")
            sys.exit(1)
        self.jitted_compute_control = self.compute_control

        self.robot = types.SimpleNamespace(mass=config.mass, inertia=np.asarray(config.inertia), dt=self.dt)
        self.Q = np.diag(np.array(list(_cabi.Q_DIAG) + [0.0] * 12, np.float32))  # :118-130
        self.mu = mp["mu"]
        self.f_z_max = mp["grf_max"]
        self.f_z_min = mp["grf_min"]
        self.best_control_parameters = np.zeros((self.num_control_parameters,), dtype=np.float32)
        self.master_key = np.array([42, 0], dtype=np.uint64)  # jax.random.PRNGKey(42), :169
        self.step_freq_available = list(mp.get("step_freq_available", [1.4, 2.0, 2.4]))

        param = self.control_parametrization if self.control_parametrization in _cabi.PARAM else "zero_order"
        self._handle = _cabi.Handle(_cabi.make_config(
            mass=config.mass, inertia=config.inertia, horizon=self.horizon, dt=self.dt, parametrization=param,
            num_splines=self.num_spline, mu=self.mu, grf_min=self.f_z_min, grf_max=self.f_z_max,
            sampling_method=self.sampling_method, num_parallel_computations=self.num_parallel_computations,
            gait_adaptive=self._gait_adaptive,
            use_nonuniform_discretization=mp.get("use_nonuniform_discretization", False),
            horizon_fine_grained=mp.get("horizon_fine_grained", 2), dt_fine_grained=mp.get("dt_fine_grained", 0.01),
            sigma_random_sampling=mp.get("sigma_random_sampling", (0.2, 3, 10)), sigma_mppi=mp.get("sigma_mppi", 3),
            step_freq_available=self.step_freq_available, device=device_index))
        assert self._handle.P == self.num_control_parameters

    # ------------------------------------------------------------------ control parametrisations (host, float32)
    # Used by shift_solution and available to callers as upstream; the rollouts evaluate the same formulas on the GPU.
    def _chunk(self, step, horizon_leg):
        f32 = np.float32
        bounds = np.linspace(0, self.horizon, self.num_spline + 1).astype(np.float32)
        index = int(np.max(np.where(f32(step) >= bounds, np.arange(self.num_spline + 1), 0)))
        tau = f32(step) / (f32(horizon_leg) / f32(self.num_spline))
        return index, f32(tau - f32(index))

    @staticmethod
    def _at(p, i):
        """jnp scalar indexing: one negative wrap, then clamp (out-of-range reads do not raise under jax)."""
        n = len(p)
        i = i + n if i < 0 else i
        return np.float32(p[min(max(i, 0), n - 1)])

    def compute_linear_spline(self, parameters, step, horizon_leg):
        """centroidal_nmpc_jax.py:181-201"""
        index, q = self._chunk(step, horizon_leg)
        shift = self.num_spline + 1
        one = np.float32(1.0)
        return tuple((one - q) * self._at(parameters, index + shift * k) + q * self._at(parameters, index + shift * k + 1)
                     for k in range(3))

    def compute_cubic_spline(self, parameters, step, horizon_leg):
        """centroidal_nmpc_jax.py:204-257 (start_index = 10 * index as written upstream, :219)"""
        index, q = self._chunk(step, horizon_leg)
        f32 = np.float32
        q2, q3 = q * q, q * q * q
        a, b = f32(2) * q3 - f32(3) * q2 + f32(1), q3 - f32(2) * q2 + q
        c, d = f32(-2) * q3 + f32(3) * q2, q3 - q2
        s = 10 * index
        out = []
        for k in (0, 4, 8):
            p0, p1, p2, p3 = (self._at(parameters, s + k + j) for j in range(4))
            phi = f32(0.5) * ((p2 - p1) + (p1 - p0))
            phi_next = f32(0.5) * ((p3 - p2) + (p2 - p1))
            out.append(a * p1 + b * phi + c * p2 + d * phi_next)
        return tuple(out)

    def compute_zero_order_spline(self, parameters, step, horizon_leg):
        """centroidal_nmpc_jax.py:259-268 (index = int16(step))"""
        i = int(np.int16(step))
        return (self._at(parameters, i), self._at(parameters, i + self.horizon), self._at(parameters, i + 2 * self.horizon))

    # ------------------------------------------------------------------ key / sigma state, :498-511
    def with_newkey(self):
        self.master_key = np.array([self.master_key[0], self.master_key[1] + np.uint64(1)], dtype=np.uint64)
        return self

    def get_key(self):
        return self.master_key

    def with_newsigma(self, sigma):
        self.sigma_cem_mppi = sigma
        return self

    def get_sigma(self):
        return self.sigma_cem_mppi

    def reset(self):
        print("Resetting the controller")

    # ------------------------------------------------------------------ host-side preparation, :513-627
    def shift_solution(self, best_control_parameters, step):
        """centroidal_nmpc_jax.py:513-561: entries 0, 2, 4 of every leg block are overwritten with the leg's force at
        `step` (the upstream indices, kept as they are)."""
        best = np.array(best_control_parameters, dtype=np.float32)
        n = self.num_control_parameters_single_leg
        for leg, fun in enumerate((self.spline_fun_FL, self.spline_fun_FR, self.spline_fun_RL, self.spline_fun_RR)):
            block = best[leg * n:(leg + 1) * n]
            fx, fy, fz = fun(block.copy(), step, self.horizon)
            block[0], block[2], block[4] = fx, fy, fz
        return best

    def prepare_state_and_reference(self, state_current, reference_state, current_contact, previous_contact,
                                    mpc_frequency=100):
        """centroidal_nmpc_jax.py:563-627: dicts -> state[24], reference[24]; swing feet take their reference
        foothold; a leg's parameter block is zeroed on its stance -> swing edge."""
        if self._config.mpc_params.get("shift_solution", False):
            self.best_control_parameters = self.shift_solution(self.best_control_parameters, 1.0 / mpc_frequency)
        feet = ("FL", "FR", "RL", "RR")
        state = np.concatenate([np.asarray(state_current[k], np.float64).reshape(-1) for k in
                                ("position", "linear_velocity", "orientation", "angular_velocity")] +
                               [np.asarray(state_current["foot_" + f], np.float64).reshape(3) for f in feet]).reshape((24,))
        for leg, f in enumerate(feet):
            if current_contact[leg] == 0.0:
                state[12 + 3 * leg:15 + 3 * leg] = np.asarray(reference_state["ref_foot_" + f]).reshape((3,))
        reference = np.concatenate([np.asarray(reference_state[k], np.float64).reshape(-1) for k in
                                    ("ref_position", "ref_linear_velocity", "ref_orientation", "ref_angular_velocity")] +
                                   [np.asarray(reference_state["ref_foot_" + f], np.float64).reshape(3) for f in feet]
                                   ).reshape((24,))
        self.best_control_parameters = np.array(self.best_control_parameters)
        n = self.num_control_parameters_single_leg
        for leg in range(4):
            if previous_contact[leg] == 1 and current_contact[leg] == 0:
                self.best_control_parameters[leg * n:(leg + 1) * n] = 0.0
        return state, reference

    # ------------------------------------------------------------------ the hot path
    def _solve(self, state, reference, contact_sequence, best_control_parameters, key, sigma=None, timing=None,
               nominal_step_frequency=0.0, optimize_swing=0):
        key = np.asarray(key, dtype=np.uint64).reshape(-1)
        r = self._handle.solve(state, reference, contact_sequence, best_control_parameters, int(key[0]), int(key[-1]),
                               timing=timing, nominal_freq=float(nominal_step_frequency),
                               optimize_swing=int(optimize_swing), sigma_in=sigma, want_costs="lazy")
        return r

    def compute_control_random_sampling(self, state, reference, contact_sequence, best_control_parameters, key, timing,
                                        nominal_step_frequency, optimize_swing):
        """centroidal_nmpc_jax.py:629-787 -> (GRFs[12], footholds[12], predicted_state[24], best[P], best_cost,
        best_freq, costs[N])"""
        r = self._solve(state, reference, contact_sequence, best_control_parameters, key, timing=self._timing(timing),
                        nominal_step_frequency=nominal_step_frequency, optimize_swing=optimize_swing)
        return r["grf"], r["footholds"], r["predicted_state"], r["best"], r["best_cost"], r["best_freq"], r["costs"]

    def compute_control_mppi(self, state, reference, contact_sequence, best_control_parameters, key, timing,
                             nominal_step_frequency, optimize_swing):
        """centroidal_nmpc_jax.py:789-932, same tuple as random sampling"""
        return self.compute_control_random_sampling(state, reference, contact_sequence, best_control_parameters, key,
                                                    timing, nominal_step_frequency, optimize_swing)

    def compute_control_cem_mppi(self, state, reference, contact_sequence, best_control_parameters, key, sigma):
        """centroidal_nmpc_jax.py:934-1094 -> the 7-tuple plus new_sigma[P]"""
        r = self._solve(state, reference, contact_sequence, best_control_parameters, key, sigma=sigma)
        return (r["grf"], r["footholds"], r["predicted_state"], r["best"], r["best_cost"], r["best_freq"], r["costs"],
                r["sigma"])

    def compute_control_iterated(self, state, reference, contact_sequence, timing=None, nominal_step_frequency=0.0,
                                 optimize_swing=0, sigma0=None, iterations=None):
        """The whole `for iter_sampling in range(num_sampling_iterations)` loop of
        SRBDControllerInterface.compute_control (interfaces/srbd_controller_interface.py:131-170) as ONE call: every
        iteration advances the key (with_newkey), starts from the previous iteration's mean and, for cem_mppi, from
        its sigma (iteration 0 from `sigma0` = mpc_params['sigma_cem_mppi'], :134-135) -- chained on the device, no
        host round trip in between.  Updates best_control_parameters, master_key and sigma_cem_mppi exactly as the
        loop would and returns the last iteration's tuple (7 entries, 8 with the new sigma for cem_mppi)."""
        K = int(self.num_sampling_iterations if iterations is None else iterations)
        cem = self.sampling_method == "cem_mppi"
        if cem and sigma0 is None:
            sigma0 = self._config.mpc_params["sigma_cem_mppi"]
        sigma = np.broadcast_to(np.asarray(sigma0, np.float32), (self.num_control_parameters,)) if cem else None
        key = np.asarray(self.master_key, dtype=np.uint64).reshape(-1)
        r = self._handle.solve(state, reference, contact_sequence, self.best_control_parameters, int(key[0]),
                               int(key[-1]) + 1, timing=self._timing(timing), nominal_freq=float(nominal_step_frequency),
                               optimize_swing=int(optimize_swing), sigma_in=sigma, want_costs="lazy", iterations=K)
        self.master_key = np.array([key[0], key[-1] + np.uint64(K)], dtype=np.uint64)  # K times with_newkey
        self.best_control_parameters = r["best"]
        out = (r["grf"], r["footholds"], r["predicted_state"], r["best"], r["best_cost"], r["best_freq"], r["costs"])
        if cem:
            self.sigma_cem_mppi = r["sigma"]
            out = out + (r["sigma"],)
        return out

    def _timing(self, timing):
        return None  # the fixed-gait controller ignores timing / nominal frequency / optimize_swing (:629-639)

    def _device_costs(self, state, reference, params, contact_sequence, timing=None, freqs=None):
        import torch  # device buffers only
        dev = torch.device(self.device)
        params = np.ascontiguousarray(np.asarray(params, np.float32).reshape(-1, self.num_control_parameters))
        n = params.shape[0]
        d_params = torch.from_numpy(params).to(dev)
        d_costs = torch.empty(n, dtype=torch.float32, device=dev)
        d_freqs = None
        if freqs is not None:
            d_freqs = torch.from_numpy(np.ascontiguousarray(np.broadcast_to(np.asarray(freqs, np.float32), (n,)))).to(dev)
        torch.cuda.synchronize(dev)
        self._handle.rollout_costs(state, reference, contact_sequence, d_params.data_ptr(), n, d_costs.data_ptr(),
                                   timing=timing, d_freqs=d_freqs.data_ptr() if d_freqs is not None else None)
        return d_costs.cpu().numpy()

    def compute_rollout(self, initial_state, reference, control_parameters, contact_sequence):
        """centroidal_nmpc_jax.py:316-496: cost of one rollout"""
        return self._device_costs(initial_state, reference, control_parameters, contact_sequence)[0]

    def vectorized_rollout(self, initial_state, reference, control_parameters, contact_sequence):
        """jax.vmap(compute_rollout, in_axes=(None, None, 0, None)) (:176): params [n, P] -> costs [n]"""
        return self._device_costs(initial_state, reference, control_parameters, contact_sequence)

    jit_vectorized_rollout = vectorized_rollout

    def close(self):
        self._handle.close()
