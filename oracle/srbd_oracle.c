/*
 * oracle/srbd_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement (plain C, every intermediate float32, no FMA contraction: build with
 * -ffp-contract=off) of the reference's sampling-MPC inner loop.  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may load this file's shared object; the CUDA
 * product path never does.
 *
 * Parity status: the reference's sampling path is jax code and jax is not installable here, so the
 * restatement is pinned against the reference's OWN Python source executed under a numpy-backed
 * stand-in for the jax API (oracle/jax_numpy_standin.py, fixtures in tests/golden/sampling_*.npz made by
 * oracle/make_golden_sampling.py).  Parity against XLA's own code generation is "unpinned".
 *
 * Reference lines followed (paths relative to /root/reference/quadruped_pympc):
 *   controllers/sampling/centroidal_model_jax.py:67-91   calculate_inverse  -> inverse3()
 *   controllers/sampling/centroidal_model_jax.py:93-162  fd                 -> srbd_fd()
 *   controllers/sampling/centroidal_model_jax.py:164-174 integrate_jax      -> srbd_integrate()
 *   controllers/sampling/centroidal_nmpc_jax.py:181-201  compute_linear_spline
 *   controllers/sampling/centroidal_nmpc_jax.py:204-257  compute_cubic_spline
 *   controllers/sampling/centroidal_nmpc_jax.py:259-268  compute_zero_order_spline
 *   controllers/sampling/centroidal_nmpc_jax.py:270-314  enforce_force_constraints
 *   controllers/sampling/centroidal_nmpc_jax.py:316-496  compute_rollout (fixed gait)
 *   controllers/sampling/centroidal_nmpc_jax_gait_adaptive.py:326-501 compute_rollout (per-sample gait)
 *   helpers/periodic_gait_generator_jax.py:68-89,136-151 run / compute_contact_sequence
 *   controllers/sampling/centroidal_nmpc_jax.py:629-787 / 789-932 / 934-1094 compute_control_*
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* -DORACLE_ALT_TRIG: a second implementation of the transcendentals (double-precision libm rounded to float, i.e.
 * correctly rounded in practice) for tools/fp32_envelope.py, which measures how far two faithful float32
 * implementations of the reference drift apart.  Never defined in the checker or the timed builds. */
#ifdef ORACLE_ALT_TRIG
#define sinf(x) ((float)sin((double)(x)))
#define cosf(x) ((float)cos((double)(x)))
#endif
#ifdef _OPENMP
#include <omp.h>
#endif

#define ORACLE_MAX_H 64

enum { PARAM_ZERO_ORDER = 0, PARAM_LINEAR = 1, PARAM_CUBIC = 2 };
enum { METHOD_RANDOM = 0, METHOD_MPPI = 1, METHOD_CEM = 2 };

typedef struct {
    int horizon;          /* mpc_params['horizon'] */
    int parametrization;  /* PARAM_* */
    int num_splines;      /* mpc_params['num_splines'] */
    int gait_adaptive;    /* 1 = centroidal_nmpc_jax_gait_adaptive.py */
    float mass;
    float inertia[9];     /* row major, config.inertia cast to float32 */
    float mu, f_z_min, f_z_max;
    float dts[ORACLE_MAX_H];
    float q_diag[24];     /* centroidal_nmpc_jax.py:118-130 */
    float duty_factor;    /* 0.65: centroidal_nmpc_jax_gait_adaptive.py:179 */
    float mpc_dt;         /* pgg mpc_dt = cfg dt */
    int cem_elite;        /* 10 (fixed gait :1075) or 200 (gait adaptive :1115) */
    int cem_sqrt;         /* 1 fixed gait, 0 gait adaptive */
    float cem_lo, cem_hi; /* clip bounds 0.2,5 / 0.5,3 */
} oracle_cfg;

int oracle_num_params(const oracle_cfg *c) {
    if (c->parametrization == PARAM_LINEAR) return (c->num_splines + 1) * 3 * 4;
    if (c->parametrization == PARAM_CUBIC) return 4 * 3 * c->num_splines * 4;
    return c->horizon * 3 * 4;
}
static int leg_params(const oracle_cfg *c) { return oracle_num_params(c) / 4; }

/* centroidal_model_jax.py:67-91, same cofactor order, nine divisions by DET */
static void inverse3(const float A[9], float out[9]) {
    float a11 = A[0], a12 = A[1], a13 = A[2], a21 = A[3], a22 = A[4], a23 = A[5], a31 = A[6], a32 = A[7], a33 = A[8];
    float DET = a11 * (a33 * a22 - a32 * a23) - a21 * (a33 * a12 - a32 * a13) + a31 * (a23 * a12 - a22 * a13);
    out[0] = (a33 * a22 - a32 * a23) / DET;
    out[1] = -(a33 * a12 - a32 * a13) / DET;
    out[2] = (a23 * a12 - a22 * a13) / DET;
    out[3] = -(a33 * a21 - a31 * a23) / DET;
    out[4] = (a33 * a11 - a31 * a13) / DET;
    out[5] = -(a23 * a11 - a21 * a13) / DET;
    out[6] = (a32 * a21 - a31 * a22) / DET;
    out[7] = -(a32 * a11 - a31 * a12) / DET;
    out[8] = (a22 * a11 - a21 * a12) / DET;
}

static void matvec3(const float M[9], const float v[3], float out[3]) {
    for (int i = 0; i < 3; ++i) out[i] = (M[3 * i] * v[0] + M[3 * i + 1] * v[1]) + M[3 * i + 2] * v[2];
}

/* skew(v) @ f, centroidal_model_jax.py:98-99 (zero entries of the matrix product dropped) */
static void skew_mul(const float v[3], const float f[3], float out[3]) {
    out[0] = (-v[2]) * f[1] + v[1] * f[2];
    out[1] = v[2] * f[0] + (-v[0]) * f[2];
    out[2] = (-v[1]) * f[0] + v[0] * f[1];
}

/* centroidal_model_jax.py:93-162 */
static void srbd_fd(const oracle_cfg *c, const float inertia_inv[9], const float st[24], const float force[12],
                    const float stance[4], float out[12]) {
    const float *com = st, *vel = st + 3, *w = st + 9;
    float roll = st[6], pitch = st[7], yaw = st[8];
    float temp[3];
    for (int k = 0; k < 3; ++k)
        temp[k] = ((force[k] * stance[0] + force[3 + k] * stance[1]) + force[6 + k] * stance[2]) + force[9 + k] * stance[3];
    float inv_m = 1.0f / c->mass;
    float acc[3] = {inv_m * temp[0] + 0.0f, inv_m * temp[1] + 0.0f, inv_m * temp[2] + (-9.81f)};

    float sr = sinf(roll), cr = cosf(roll), sp = sinf(pitch), cp = cosf(pitch), sy = sinf(yaw), cy = cosf(yaw);
    float conj[9] = {1.0f, 0.0f, -sp, 0.0f, cr, cp * sr, 0.0f, -sr, cp * cr};

    float tau[3] = {0.0f, 0.0f, 0.0f};
    for (int leg = 0; leg < 4; ++leg) {
        float r[3] = {st[12 + 3 * leg] - com[0], st[13 + 3 * leg] - com[1], st[14 + 3 * leg] - com[2]};
        float t[3];
        skew_mul(r, force + 3 * leg, t);
        for (int k = 0; k < 3; ++k) {
            float tk = t[k] * stance[leg];
            tau[k] = (leg == 0) ? tk : tau[k] + tk;
        }
    }
    float conj_inv[9], euler_rates[3];
    inverse3(conj, conj_inv);
    matvec3(conj_inv, w, euler_rates);

    float R[9] = {cp * cy, cp * sy, -sp,
                  sr * sp * cy - cr * sy, sr * sp * sy + cr * cy, sr * cp,
                  cr * sp * cy + sr * sy, cr * sp * sy - sr * cy, cr * cp};
    float Iw[3], wIw[3], a1[3], Rt[3], a2[3];
    matvec3(c->inertia, w, Iw);
    skew_mul(w, Iw, wIw);
    matvec3(inertia_inv, wIw, a1);
    matvec3(R, tau, Rt);
    matvec3(inertia_inv, Rt, a2);
    for (int k = 0; k < 3; ++k) {
        out[k] = vel[k];
        out[3 + k] = acc[k];
        out[6 + k] = euler_rates[k];
        out[9 + k] = -a1[k] + a2[k];
    }
}

/* centroidal_model_jax.py:164-174 */
static void srbd_integrate(const oracle_cfg *c, const float inertia_inv[9], float st[24], const float force[12],
                           const float stance[4], int n) {
    float d[12];
    srbd_fd(c, inertia_inv, st, force, stance, d);
    float dt = c->dts[n];
    for (int k = 0; k < 12; ++k) st[k] = st[k] + d[k] * dt;
}

/* jnp index semantics for a scalar index into a 1-D array: negative wraps once, then clamps */
static int jidx(int i, int size) {
    if (i < 0) i += size;
    if (i < 0) i = 0;
    if (i > size - 1) i = size - 1;
    return i;
}

/* chunk index shared by the linear and cubic splines, centroidal_nmpc_jax.py:186-188 / 210-212 */
static int chunk_index(const oracle_cfg *c, float step) {
    int S = c->num_splines, idx = 0;
    for (int k = 0; k <= S; ++k) {
        float b = (float)((double)c->horizon * (double)k / (double)S); /* linspace(0, H, S+1) */
        int v = (step >= b) ? k : 0;
        if (v > idx) idx = v;
    }
    return idx;
}

/* step is an integer counter in the rollout and 0.0 in the GRF tail; horizon_leg is float */
static void spline_eval(const oracle_cfg *c, const float *p, int np, float step, float horizon_leg, float f[3]) {
    int H = c->horizon, S = c->num_splines;
    if (c->parametrization == PARAM_ZERO_ORDER) { /* :259-268 */
        int index = (int)(int16_t)step;
        f[0] = p[jidx(index, np)];
        f[1] = p[jidx(index + H, np)];
        f[2] = p[jidx(index + 2 * H, np)];
        return;
    }
    int index = chunk_index(c, step);
    float tau = step / (horizon_leg / (float)S);
    tau = tau - (float)index;
    float q = (tau - 0.0f) / (1.0f - 0.0f);
    if (c->parametrization == PARAM_LINEAR) { /* :181-201 */
        int shift = S + 1;
        f[0] = (1 - q) * p[jidx(index, np)] + q * p[jidx(index + 1, np)];
        f[1] = (1 - q) * p[jidx(index + shift, np)] + q * p[jidx(index + shift + 1, np)];
        f[2] = (1 - q) * p[jidx(index + shift * 2, np)] + q * p[jidx(index + shift * 2 + 1, np)];
        return;
    }
    /* cubic :204-257, start_index = 10*index as written */
    int s = 10 * index;
    float a = 2 * q * q * q - 3 * q * q + 1;
    float b = (q * q * q - 2 * q * q + q) * 1.0f;
    float cc = -2 * q * q * q + 3 * q * q;
    float d = (q * q * q - q * q) * 1.0f;
    for (int ax = 0; ax < 3; ++ax) {
        int k = 4 * ax;
        float p0 = p[jidx(s + k, np)], p1 = p[jidx(s + k + 1, np)], p2 = p[jidx(s + k + 2, np)], p3 = p[jidx(s + k + 3, np)];
        float phi = (1.0f / 2.0f) * (((p2 - p1) / 1.0f) + ((p1 - p0) / 1.0f));
        float phi_next = (1.0f / 2.0f) * (((p3 - p2) / 1.0f) + ((p2 - p1) / 1.0f));
        f[ax] = a * p1 + b * phi + cc * p2 + d * phi_next;
    }
}

/* centroidal_nmpc_jax.py:270-314; the selects keep the bound when the comparison is false (NaN -> bound) */
static void enforce_force_constraints(const oracle_cfg *c, float f[12]) {
    for (int leg = 0; leg < 4; ++leg) {
        float fz = f[3 * leg + 2];
        fz = (fz > c->f_z_min) ? fz : c->f_z_min;
        f[3 * leg + 2] = fz;
    }
    for (int leg = 0; leg < 4; ++leg) {
        float fz = f[3 * leg + 2];
        fz = (fz < c->f_z_max) ? fz : c->f_z_max;
        f[3 * leg + 2] = fz;
    }
    for (int leg = 0; leg < 4; ++leg) {
        float fz = f[3 * leg + 2];
        for (int ax = 0; ax < 2; ++ax) {
            float v = f[3 * leg + ax];
            v = (v > -c->mu * fz) ? v : -c->mu * fz;
            v = (v < c->mu * fz) ? v : c->mu * fz;
            f[3 * leg + ax] = v;
        }
    }
}

/* force post-processing shared by the rollout (:376-409) and the GRF tail (:712-741) */
static void shape_forces(const oracle_cfg *c, float f[12], const float stance[4]) {
    float n_stance = ((stance[0] + stance[1]) + stance[2]) + stance[3];
    float ref_force = (float)((double)c->mass * 9.81) / n_stance;
    for (int leg = 0; leg < 4; ++leg) f[3 * leg + 2] = ref_force + f[3 * leg + 2];
    for (int leg = 0; leg < 4; ++leg) {
        f[3 * leg + 0] = f[3 * leg + 0] * stance[leg] / 3.0f; /* 30/10 */
        f[3 * leg + 1] = f[3 * leg + 1] * stance[leg] / 3.0f;
        f[3 * leg + 2] = f[3 * leg + 2] * stance[leg];
    }
    enforce_force_constraints(c, f);
}

/* helpers/periodic_gait_generator_jax.py:68-89,136-151 (trot-only clock; all legs advance alike) */
static void gait_contact_sequence(const oracle_cfg *c, const float timing[4], float step_freq, float *seq /*[4][H]*/) {
    float t[4] = {timing[0], timing[1], timing[2], timing[3]};
    for (int n = 0; n < c->horizon; ++n)
        for (int leg = 0; leg < 4; ++leg) {
            t[leg] = (t[leg] >= 1.0f) ? 0.0f : t[leg];
            t[leg] = t[leg] + c->mpc_dt * step_freq;
            seq[leg * c->horizon + n] = (t[leg] < c->duty_factor) ? 1.0f : 0.0f;
        }
}

static float stage_cost(const oracle_cfg *c, const float st[24], const float ref[24]) {
    /* state_error.T @ Q @ state_error with the diagonal Q of :118-130 */
    float acc = 0.0f;
    for (int k = 0; k < 24; ++k) {
        float e = st[k] - ref[k];
        acc = acc + (e * c->q_diag[k]) * e;
    }
    return acc;
}

/* one rollout; fixed gait: centroidal_nmpc_jax.py:316-496, per-sample gait: ..._gait_adaptive.py:326-501 */
static float rollout_cost(const oracle_cfg *c, const float inertia_inv[9], const float state0[24], const float ref[24],
                          const float *params, const float *contact_seq_in /*[4][H] fixed gait*/,
                          const float timing[4], float step_freq) {
    int H = c->horizon, np = leg_params(c);
    float seq_local[4 * ORACLE_MAX_H];
    const float *seq = contact_seq_in;
    float horizon_leg[4] = {(float)H, (float)H, (float)H, (float)H};
    if (c->gait_adaptive) {
        gait_contact_sequence(c, timing, step_freq, seq_local);
        seq = seq_local;
        for (int leg = 0; leg < 4; ++leg) {
            float s = 0.0f;
            for (int n = 0; n < H; ++n) s = s + seq[leg * H + n];
            horizon_leg[leg] = s + 1.0f; /* :345-348 */
        }
    }
    float st[24];
    memcpy(st, state0, sizeof st);
    float cost = 0.0f;
    int n_[4] = {-1, -1, -1, -1};
    for (int n = 0; n < H; ++n) {
        float stance[4], f[12];
        for (int leg = 0; leg < 4; ++leg) {
            stance[leg] = seq[leg * H + n];
            if (c->gait_adaptive) n_[leg] = (int)((float)n_[leg] + 1 * stance[leg]); /* :353-356, int32 array */
            else n_[leg] = n;                                                       /* :348-351 */
            spline_eval(c, params + leg * np, np, (float)n_[leg], horizon_leg[leg], f + 3 * leg);
        }
        shape_forces(c, f, stance);
        srbd_integrate(c, inertia_inv, st, f, stance, n);
        cost = cost + stage_cost(c, st, ref);
    }
    if (c->gait_adaptive) cost += (step_freq - 1.3f) * 100 * (step_freq - 1.3f); /* :500 */
    return cost;
}

static void inertia_inverse(const oracle_cfg *c, float out[9]) { inverse3(c->inertia, out); }

/* jit(vmap(compute_rollout)): params [N,P] -> costs [N].  contact_seq used when !gait_adaptive,
 * timing/freqs used when gait_adaptive. */
void oracle_rollout_costs(const oracle_cfg *c, const float *state, const float *ref, const float *contact_seq,
                          const float *timing, const float *params, const float *freqs, int N, float *costs) {
    float inertia_inv[9];
    inertia_inverse(c, inertia_inv);
    int P = oracle_num_params(c);
#pragma omp parallel for schedule(static)
    for (int i = 0; i < N; ++i)
        costs[i] = rollout_cost(c, inertia_inv, state, ref, params + (size_t)i * P, contact_seq, timing,
                                freqs ? freqs[i] : 0.0f);
}

typedef struct {
    float grf[12];
    float footholds[12];
    float predicted_state[24];
    float best_cost;
    float best_freq;
    int best_index;
} oracle_result;

static int cmp_cost_idx(const void *a, const void *b) {
    const float *x = (const float *)a, *y = (const float *)b;
    if (x[0] < y[0]) return -1;
    if (x[0] > y[0]) return 1;
    return (x[1] < y[1]) ? -1 : (x[1] > y[1]);
}

/*
 * One compute_control_* call with an explicit noise matrix.
 *   eps  [N,P]  = additional_random_parameters (row 0 zero, sigma already applied)
 *   freqs[N]    = step_frequencies_vec (gait adaptive only)
 *   best [P]    in: previous best, out: updated best
 *   sigma_out[P] (cem only; may be NULL)
 * Sums of the MPPI update are accumulated in double (XLA's reduction order is unspecified).
 */
void oracle_solve(const oracle_cfg *c, int method, const float *state, const float *ref, const float *contact_seq,
                  const float *timing, const float *eps, const float *freqs, int N, float *best, float *costs,
                  float *sigma_out, oracle_result *out) {
    int P = oracle_num_params(c), np = leg_params(c), H = c->horizon;
    float *params = (float *)malloc((size_t)N * P * sizeof(float));
    for (size_t i = 0; i < (size_t)N; ++i)
        for (int p = 0; p < P; ++p) params[i * P + p] = best[p] + eps[i * P + p];
    oracle_rollout_costs(c, state, ref, contact_seq, timing, params, freqs, N, costs);
    for (int i = 0; i < N; ++i) { /* :686-687 */
        if (isnan(costs[i])) costs[i] = 1000000.0f;
        if (isinf(costs[i])) costs[i] = 1000000.0f;
    }
    int bi = 0;
    for (int i = 1; i < N; ++i)
        if (costs[i] < costs[bi]) bi = i; /* nanargmin: first minimum */
    out->best_index = bi;
    out->best_cost = costs[bi];
    if (method == METHOD_RANDOM) {
        for (int p = 0; p < P; ++p) best[p] = params[(size_t)bi * P + p]; /* :692 */
    } else { /* :828-836 */
        float beta = costs[bi];
        double denom = 0.0;
        float *e = (float *)malloc((size_t)N * sizeof(float));
        for (int i = 0; i < N; ++i) {
            e[i] = expf((-1.0f / 1.0f) * (costs[i] - beta));
            denom += (double)e[i];
        }
        float denom_f = (float)denom;
        double *acc = (double *)calloc((size_t)P, sizeof(double));
        for (size_t i = 0; i < (size_t)N; ++i) {
            float w = e[i] / denom_f;
            if (w == 0.0f) continue;
            for (int p = 0; p < P; ++p) acc[p] += (double)(w * eps[i * P + p]);
        }
        for (int p = 0; p < P; ++p) best[p] = best[p] + (float)acc[p];
        free(acc);
        free(e);
    }
    if (c->gait_adaptive) out->best_freq = freqs[bi];
    else out->best_freq = (method == METHOD_CEM) ? 1.65f : 1.4f;

    /* GRF of the first step, :707-747 (uses the PASSED-IN contact sequence in both files) */
    float stance[4], f[12];
    for (int leg = 0; leg < 4; ++leg) {
        stance[leg] = contact_seq[leg * H + 0];
        spline_eval(c, best + leg * np, np, 0.0f, 1.0f, f + 3 * leg);
    }
    shape_forces(c, f, stance);
    memcpy(out->grf, f, sizeof out->grf);
    memset(out->footholds, 0, sizeof out->footholds);
    float inertia_inv[9];
    inertia_inverse(c, inertia_inv);
    memcpy(out->predicted_state, state, 24 * sizeof(float));
    srbd_integrate(c, inertia_inv, out->predicted_state, f, stance, 0); /* :751-784 */

    if (method == METHOD_CEM && sigma_out) { /* :1075-1081 / gait adaptive :1115-1121 */
        int K = c->cem_elite < N ? c->cem_elite : N;
        float *ci = (float *)malloc((size_t)N * 2 * sizeof(float));
        for (int i = 0; i < N; ++i) { ci[2 * i] = costs[i]; ci[2 * i + 1] = (float)i; }
        qsort(ci, (size_t)N, 2 * sizeof(float), cmp_cost_idx); /* stable by (cost, index) */
        for (int p = 0; p < P; ++p) {
            double mean = 0.0;
            for (int k = 0; k < K; ++k) mean += (double)eps[(size_t)ci[2 * k + 1] * P + p];
            mean /= K;
            double var = 0.0;
            for (int k = 0; k < K; ++k) {
                double d = (double)eps[(size_t)ci[2 * k + 1] * P + p] - mean;
                var += d * d;
            }
            float v = (float)(var / (K - 1)) + 1e-8f;
            if (c->cem_sqrt) v = sqrtf(v);
            v = (v > c->cem_hi) ? c->cem_hi : v;
            v = (v < c->cem_lo) ? c->cem_lo : v;
            sigma_out[p] = v;
        }
        free(ci);
    }
    free(params);
}

/* ---- CPU noise for the timed baseline only (bench.py): Philox4x32-10 + Box-Muller, the same construction the CUDA
 * path uses, so the CPU arm pays for its random numbers as the reference's jax.random.normal call does
 * (centroidal_nmpc_jax.py:806-812).  Not used by any parity check (those feed explicit noise). */
static void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
    for (int r = 0; r < 10; ++r) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c[0], p1 = (uint64_t)0xCD9E8D57u * c[2];
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0, n1 = (uint32_t)p1, n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1, n3 = (uint32_t)p0;
        c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
}
void oracle_fill_noise(float *eps, int N, int P, float sigma, uint64_t seed) {
#pragma omp parallel for schedule(static)
    for (int i = 0; i < N; ++i) {
        float *row = eps + (size_t)i * P;
        if (i == 0) { memset(row, 0, (size_t)P * sizeof(float)); continue; }
        for (int g = 0; g < P / 4; ++g) {
            uint32_t c[4] = {(uint32_t)g, (uint32_t)i, 0u, (uint32_t)(seed >> 32)};
            philox4x32_10(c, (uint32_t)seed, 0x1234567u);
            for (int h = 0; h < 2; ++h) {
                float u1 = ((float)(c[2 * h] >> 8) + 0.5f) * 5.9604644775390625e-08f;
                float u2 = (float)(c[2 * h + 1] >> 8) * 5.9604644775390625e-08f;
                float r = sqrtf(-2.0f * logf(u1)), a = 6.2831853071795864f * u2;
                row[4 * g + 2 * h] = sigma * r * cosf(a);
                row[4 * g + 2 * h + 1] = sigma * r * sinf(a);
            }
        }
    }
}

int oracle_sizeof_cfg(void) { return (int)sizeof(oracle_cfg); }
int oracle_sizeof_result(void) { return (int)sizeof(oracle_result); }
/* torchrun exports OMP_NUM_THREADS=1 to every rank; the CPU baseline arm asks for all host threads explicitly */
void oracle_set_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

int oracle_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
