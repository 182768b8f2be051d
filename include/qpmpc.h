/*
 * include/qpmpc.h -- C ABI of the B200 (sm_100a) sampling-MPC hot path.
 *
 * The reference (Magicyw/Quadruped-PyMPC-TAMOLS) has no FFI layer: its boundary for this path is the
 * Python class `Sampling_MPC` (quadruped_pympc/controllers/sampling/centroidal_nmpc_jax.py:21 and
 * centroidal_nmpc_jax_gait_adaptive.py:23) whose `jitted_compute_control` is an XLA executable.
 * Every entry point below names the reference interface it replaces.  The Python drop-in classes in
 * quadruped_pympc_tamols_b200/controllers/sampling/ bind these symbols with ctypes (INTEGRATION.md shows
 * the stub a reference maintainer would add).
 *
 * Conventions: plain pointers and sizes only; every function returns 0 on success and a negative
 * QPMPC_E_* code otherwise, with a message available from qpmpc_last_error().  A handle owns its device
 * scratch and one CUDA stream; it is thread-compatible (one caller at a time), like the reference object.
 * Pointers named d_* are DEVICE pointers, all others are HOST pointers.
 */
#ifndef QPMPC_H
#define QPMPC_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define QPMPC_MAX_H 32           /* maximum mpc_params['horizon'] */
#define QPMPC_MAX_FREQ 8         /* maximum len(mpc_params['step_freq_available']) */
#define QPMPC_MAX_P (12 * QPMPC_MAX_H)

enum { QPMPC_PARAM_ZERO_ORDER = 0, QPMPC_PARAM_LINEAR = 1, QPMPC_PARAM_CUBIC = 2 };
enum { QPMPC_METHOD_RANDOM = 0, QPMPC_METHOD_MPPI = 1, QPMPC_METHOD_CEM = 2 };
enum {
    QPMPC_OK = 0,
    QPMPC_E_ARG = -1,     /* bad argument / unsupported configuration */
    QPMPC_E_CUDA = -2,    /* a CUDA runtime call failed (message has the cudaError string) */
    QPMPC_E_NODEVICE = -3 /* no sm_100 device: there is NO CPU fallback (reference falls back silently,
                             centroidal_nmpc_jax.py:43-48; this library fails loudly instead) */
};

/* Everything Sampling_MPC.__init__ reads from config.py (centroidal_nmpc_jax.py:23-177,
 * centroidal_nmpc_jax_gait_adaptive.py:25-187) plus the model constants of
 * centroidal_model_jax.py:19-53. */
typedef struct qpmpc_config {
    int32_t horizon;                   /* mpc_params['horizon'] (<= QPMPC_MAX_H) */
    int32_t parametrization;           /* mpc_params['control_parametrization'] -> QPMPC_PARAM_* */
    int32_t num_splines;               /* mpc_params['num_splines'] */
    int32_t gait_adaptive;             /* mpc_params['optimize_step_freq']: 1 selects the *_gait_adaptive.py semantics */
    int32_t sampling_method;           /* mpc_params['sampling_method'] -> QPMPC_METHOD_* */
    int32_t num_parallel_computations; /* mpc_params['num_parallel_computations'] = rollouts per robot */
    int32_t max_robots;                /* capacity of the batched entry points (>= 1) */
    int32_t device;                    /* CUDA device ordinal */
    float mass;                        /* config.mass */
    float inertia[9];                  /* config.inertia, row major */
    float mu, f_z_min, f_z_max;        /* mpc_params['mu'], ['grf_min'], ['grf_max'] */
    float dts[QPMPC_MAX_H];            /* centroidal_model_jax.py:42-53 (non-uniform discretisation folded in) */
    float q_diag[12];                  /* diagonal of Q for the 12 body states, centroidal_nmpc_jax.py:118-130 */
    float duty_factor, mpc_dt;         /* PeriodicGaitGeneratorJax(duty_factor=0.65, mpc_dt=dt), ..._gait_adaptive.py:179 */
    float sigma_random_sampling[3];    /* mpc_params['sigma_random_sampling'] */
    float sigma_mppi;                  /* mpc_params['sigma_mppi'] */
    int32_t num_step_freq;             /* len(mpc_params['step_freq_available']) */
    float step_freq_available[QPMPC_MAX_FREQ];
    int32_t cem_elite, cem_sqrt;       /* 10,1 (centroidal_nmpc_jax.py:1075-1081) / 200,0 (..._gait_adaptive.py:1115-1121) */
    float cem_lo, cem_hi;              /* 0.2,5 / 0.5,3 */
} qpmpc_config;

/* The fixed-size part of what compute_control_* returns (centroidal_nmpc_jax.py:776-787). */
typedef struct qpmpc_result {
    float grf[12];             /* nmpc_GRFs */
    float footholds[12];       /* nmpc_footholds (always zero upstream) */
    float predicted_state[24]; /* nmpc_predicted_state */
    float best_cost;
    float best_freq;           /* 1.4 / 1.65 constants, or the sampled frequency of the best rollout */
    int32_t best_index;
    int32_t reserved;
} qpmpc_result;

typedef struct qpmpc_handle qpmpc_handle;

/* Sampling_MPC.__init__ (centroidal_nmpc_jax.py:23-177). */
int qpmpc_create(const qpmpc_config *cfg, qpmpc_handle **out);
/* Object lifetime end (Python GC upstream). */
int qpmpc_destroy(qpmpc_handle *h);
/* self.num_control_parameters (centroidal_nmpc_jax.py:52-93). */
int qpmpc_num_params(const qpmpc_handle *h);
/* Message of the last failing call on this thread. */
const char *qpmpc_last_error(void);
/* Library / build identification ("qpmpc-b200 sm_100a ..."). */
const char *qpmpc_version(void);
/* The CUDA stream (cudaStream_t as void*) the handle launches on; lets callers time with events. */
void *qpmpc_stream(qpmpc_handle *h);
/* Number of kernels launched by this handle since creation (bench.py's gpu_launches claim). */
int64_t qpmpc_launch_count(const qpmpc_handle *h);
/* Which rollout kernel the most recent solve / rollout_costs call launched: "pair" (two rollouts per thread, packed
 * FP32: zero-order fixed-gait launches of the throughput regime), "quad" (four lanes per rollout: latency regime) or
 * "thread" (one thread per rollout).  Measurement and test hook; no reference counterpart. */
const char *qpmpc_last_kernel(const qpmpc_handle *h);

/*
 * jitted_compute_control for ONE robot with host buffers in and out:
 *   compute_control_random_sampling  centroidal_nmpc_jax.py:629-787  (gait adaptive: ..._gait_adaptive.py:634-806)
 *   compute_control_mppi             centroidal_nmpc_jax.py:789-932  (..._gait_adaptive.py:808-962)
 *   compute_control_cem_mppi         centroidal_nmpc_jax.py:934-1094 (..._gait_adaptive.py:964-1131)
 * state[24], reference[24], contact_sequence[4*horizon] (row major [leg][n]), best_in[P],
 * (seed, offset) = the noise key (the reference's `key`; with_newkey advances offset),
 * timing[4] / nominal_step_frequency / optimize_swing as upstream (ignored when !gait_adaptive),
 * sigma_in[P] (cem_mppi only, may be NULL otherwise).
 * d_eps [N,P] / d_freqs [N]: optional DEVICE buffers holding reference-supplied noise
 * (additional_random_parameters with sigma applied, step_frequencies_vec); NULL = in-kernel Philox.
 * Outputs: *out, best_out[P], sigma_out[P] (cem only, may be NULL), costs_out[N] (may be NULL).
 * One launch: the inputs ride in the kernel parameter block, the results return through mapped pinned memory
 * (QPMPC_ZEROCOPY=0: H2D copy, kernel, D2H copy, stream synchronise).  Returns when the results are complete.
 */
int qpmpc_solve(qpmpc_handle *h, const float *state, const float *reference, const float *contact_sequence,
                const float *best_in, uint64_t seed, uint64_t offset, const float *timing,
                float nominal_step_frequency, int32_t optimize_swing, const float *sigma_in, const float *d_eps,
                const float *d_freqs, qpmpc_result *out, float *best_out, float *sigma_out, float *costs_out);

/* `iterations` chained solves without a host round trip: the sampling branch of
 * SRBDControllerInterface.compute_control (interfaces/srbd_controller_interface.py:123-170), whose loop
 *   for iter_sampling in range(num_sampling_iterations): controller.with_newkey(); ... = jitted_compute_control(...)
 * feeds each iteration's best_control_parameters (and, for cem_mppi, its new sigma) to the next one.
 * Iteration i draws its noise with the key (seed, first_offset + i), reads the mean (and sigma) the previous
 * iteration left ON THE DEVICE, and only the last iteration's outputs travel back: `iterations` kernel launches
 * back to back on the handle's stream, one wait.  Arguments and outputs as qpmpc_solve (outputs = those of the last
 * iteration; costs_out = its costs).  Reference-supplied noise buffers are not supported here (Philox only).
 */
int qpmpc_solve_iterated(qpmpc_handle *h, const float *state, const float *reference, const float *contact_sequence,
                         const float *best_in, uint64_t seed, uint64_t first_offset, int32_t iterations,
                         const float *timing, float nominal_step_frequency, int32_t optimize_swing,
                         const float *sigma_in, qpmpc_result *out, float *best_out, float *sigma_out,
                         float *costs_out);

/* Device-resident problem description for the batched entry points; every pointer is a DEVICE pointer
 * with `num_robots` leading rows. */
typedef struct qpmpc_batch {
    int32_t num_robots;
    int32_t robot_offset;            /* global id of row 0: robot r of this batch draws from Philox subsequence
                                        robot_offset + r, so results do not depend on how robots are sharded */
    const float *d_state;            /* [R,24] */
    const float *d_reference;        /* [R,24] */
    const float *d_contact_sequence; /* [R,4,H] */
    const float *d_best_in;          /* [R,P] */
    const float *d_sigma_in;         /* [R,P]  cem_mppi only, else NULL */
    const float *d_timing;           /* [R,4]  gait adaptive only, else NULL */
    const float *d_nominal_freq;     /* [R]    gait adaptive only, else NULL */
    const int32_t *d_optimize_swing; /* [R]    gait adaptive only, else NULL */
    const float *d_eps;              /* [R,N,P] reference-supplied noise or NULL (Philox) */
    const float *d_freqs;            /* [R,N]   reference-supplied step frequencies or NULL */
    qpmpc_result *d_result;          /* [R] out */
    float *d_best_out;               /* [R,P] out */
    float *d_sigma_out;              /* [R,P] out, cem_mppi only, else NULL */
    float *d_costs;                  /* [R,N] out, or NULL to use handle scratch */
} qpmpc_batch;

/* R independent jitted_compute_control calls in one launch sequence (the reference runs these as separate
 * OS processes: simulation/batched_simulations.py:43-58).  Asynchronous on qpmpc_stream(h): the caller
 * synchronises.  No collective: robots are independent. */
int qpmpc_solve_batched(qpmpc_handle *h, const qpmpc_batch *batch, uint64_t seed, uint64_t offset);

/* costs[N] of the most recent qpmpc_solve / qpmpc_solve_iterated on this handle, copied to the host on request.
 * Upstream the 7th return value of jitted_compute_control is a DEVICE array nobody in the control loop reads
 * (srbd_controller_interface.py:131-168 uses best_cost only); the Python class therefore leaves the costs on the
 * device and fetches them through this call only when they are looked at.  Synchronous. */
int qpmpc_copy_costs(qpmpc_handle *h, float *costs_out);

/* jit_vectorized_rollout (centroidal_nmpc_jax.py:176-177, ..._gait_adaptive.py:186-187):
 * d_params [n,P] (DEVICE) -> d_costs [n] (DEVICE), raw costs without the NaN/Inf saturation.
 * state / reference / contact_sequence / timing are HOST; d_freqs [n] DEVICE (gait adaptive) or NULL.
 * Synchronous. */
int qpmpc_rollout_costs(qpmpc_handle *h, const float *state, const float *reference, const float *contact_sequence,
                        const float *timing, const float *d_params, const float *d_freqs, int32_t n, float *d_costs);

/* Writes the noise the fused solve would draw for (seed, offset, robot) to DEVICE buffers:
 * d_eps_out [N,P] = additional_random_parameters (row 0 zero, sigma applied: centroidal_nmpc_jax.py:647-677,
 * 806-812, 951-958), d_freqs_out [N] = step_frequencies_vec (..._gait_adaptive.py:687-692, 834-838, 994-999;
 * may be NULL).  Bit-identical to the in-kernel stream, so the oracle can consume the same noise.  Synchronous. */
int qpmpc_generate_noise(qpmpc_handle *h, uint64_t seed, uint64_t offset, int32_t robot, const float *sigma_in,
                         float nominal_step_frequency, int32_t optimize_swing, float *d_eps_out, float *d_freqs_out);

/* Measurement hooks (no reference counterpart: the reference only has wall-clock prints,
 * interfaces/wb_interface.py:231,241).  With profiling on, each solve records CUDA events around its two kernels on
 * the handle's stream; qpmpc_profile_last waits for the last solve and returns both durations in milliseconds. */
int qpmpc_set_profiling(qpmpc_handle *h, int32_t on);
int qpmpc_profile_last(qpmpc_handle *h, float *rollout_ms, float *finalize_ms);
/* FP32 FMA-chain microbenchmark on `device`: the measured CUDA-core peak (TFLOP/s, FMA = 2) used as the denominator of
 * the rollout kernel's roofline (nothing here is a dense contraction, so tensor-core peaks do not apply). */
int qpmpc_measure_fp32_peak(int32_t device, float *tflops);

/* ---- the periodic gait clock for a batch of robots, the step in front of the rollouts.
 * Replaces PeriodicGaitGenerator.run / compute_contact_sequence (helpers/periodic_gait_generator.py:47-76, 93-114)
 * looped in Python over the robots; float64 like the reference, contact flags bit-exact.
 * All pointers are DEVICE pointers with `num_robots` leading rows; legs in FL, FR, RL, RR order. */
typedef struct qpmpc_gait_batch {
    int32_t num_robots;
    int32_t horizon;                 /* columns of the contact sequence (<= QPMPC_MAX_H) */
    double *d_phase;                 /* [R,4] phase signals in [0,1)            (_phase_signal) */
    uint8_t *d_init;                 /* [R,4] 1 while a leg still waits for its phase offset (_init) */
    const double *d_phase_offset;    /* [R,4] per-gait leg delays               (phase_offset, :24-39) */
    const double *d_step_freq;       /* [R]   step frequency                    (step_freq) */
    const double *d_duty_factor;     /* [R]   duty factor */
    const uint8_t *d_full_stance;    /* [R]   1 = GaitType.FULL_STANCE: all ones (:95-98), or NULL */
    double step_dts[QPMPC_MAX_H];    /* dt of column i (entry 0 unused): contact_sequence_dts / _lenghts expanded
                                        with the reference's rule (:104-110) */
    float *d_contact;                /* out: [R,4,horizon] for qpmpc_contact_sequences (the layout of
                                        qpmpc_batch.d_contact_sequence), [R,4] for qpmpc_gait_run */
} qpmpc_gait_batch;

/* compute_contact_sequence for every robot: column 0 = run(0.0, f), column i = run(step_dts[i], f); the clocks
 * (d_phase, d_init) are left untouched, as upstream restores them (:112).  Asynchronous on the handle's stream. */
int qpmpc_contact_sequences(qpmpc_handle *h, const qpmpc_gait_batch *g);
/* run(dt, step_freq) for every robot: advances d_phase / d_init in place and writes the current contact flags. */
int qpmpc_gait_run(qpmpc_handle *h, const qpmpc_gait_batch *g, double dt);

/* ---- the step behind the solve for a batch of robots, the stance torque mapping
 * tau[leg] = -feet_jac[leg][:, legs_qvel_idx[leg]].T @ nmpc_GRFs[leg]  (interfaces/wb_interface.py:368-371), computed from
 * the result records qpmpc_solve_batched left on the device (no host round trip between solve and torques).
 * d_result [R] (qpmpc_batch.d_result), d_feet_jac [R,4,3,3] float64 = each leg's 3 x 3 translational Jacobian block,
 * row-major (upstream takes it from MuJoCo), d_tau [R,4,3] float64 out.  The swing legs' entries are overwritten by the
 * swing controller upstream (:377-396), exactly as there.  Asynchronous on the handle's stream. */
int qpmpc_stance_torques(qpmpc_handle *h, int32_t num_robots, const qpmpc_result *d_result, const double *d_feet_jac,
                         double *d_tau);

/* ---- SURVEY.md 8(f) N4, second half: dataset rows for a batch of robots, packed on the device.
 * Replaces MatLogger.record_step (simulation/simulation.py:241-304) looped in Python over robots: one float64 row
 *   [time, observable 0 flattened, observable 1 flattened, ...]
 * per robot and step, the observables being the controller's (quadruped_pympc_wrapper.py:205-243: nmpc_GRFs from the
 * qpmpc_result records, reference height / angles / feet from the reference rows, phase signal from the gait clocks,
 * ...) and whatever state arrays the simulator side keeps on the device.  Each observable is a SEGMENT: a device array
 * with one row per robot (any of four element types, any row stride -- e.g. the grf field of qpmpc_result with a stride
 * of sizeof(qpmpc_result) / 4), flattened row-major and cast to float64 exactly like MatLogger._flatten_value (:119-148).
 * d_rows [num_robots, num_columns] receives the rows of this step; the caller lays the steps out as it likes (the
 * Python writer uses [steps, robots, columns] and saves MatLogger's {header, data} table per robot).
 * Asynchronous on the handle's stream. */
#define QPMPC_MAX_OBS_SEGMENTS 24
enum { QPMPC_OBS_F32 = 0, QPMPC_OBS_F64 = 1, QPMPC_OBS_I32 = 2, QPMPC_OBS_U8 = 3 };
typedef struct qpmpc_obs_segment {
    const void *d_src;    /* device array, one row per robot */
    int64_t robot_stride; /* elements between the rows of consecutive robots */
    int32_t dtype;        /* QPMPC_OBS_* */
    int32_t count;        /* values per robot = columns this observable contributes */
} qpmpc_obs_segment;
typedef struct qpmpc_obs_layout {
    int32_t num_segments;
    int32_t num_columns;  /* 1 (time) + sum of the segment counts (checked) */
    qpmpc_obs_segment seg[QPMPC_MAX_OBS_SEGMENTS];
} qpmpc_obs_layout;
int qpmpc_pack_observables(qpmpc_handle *h, const qpmpc_obs_layout *layout, int32_t num_robots, double time,
                           double *d_rows);

#ifdef __cplusplus
}
#endif
#endif /* QPMPC_H */
