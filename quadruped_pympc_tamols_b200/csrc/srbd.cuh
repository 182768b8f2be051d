// Device-side single-rigid-body rollout: force parametrisations, force shaping, SRBD dynamics, stage cost.
// From-scratch CUDA restatement of the math of (paths relative to /root/reference/quadruped_pympc):
//   controllers/sampling/centroidal_model_jax.py:93-174            fd + explicit Euler
//   controllers/sampling/centroidal_nmpc_jax.py:181-268            zero-order / linear / cubic parametrisations
//   controllers/sampling/centroidal_nmpc_jax.py:270-314            enforce_force_constraints
//   controllers/sampling/centroidal_nmpc_jax.py:316-496            compute_rollout (given contact sequence)
//   controllers/sampling/centroidal_nmpc_jax_gait_adaptive.py:326-501  compute_rollout (per-sample gait)
//   helpers/periodic_gait_generator_jax.py:68-89,136-151           trot clock -> contact sequence
// The building blocks in the reference's operation order (generic spline evaluation, force shaping, one SRBD step,
// stage cost); k_finalize uses them for the first-step GRF and the one-step prediction, rollout.cuh builds the hot
// loops on the same formulas.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "../../include/qpmpc.h"

namespace qp {

constexpr int MAX_SPLINES = 16;

// Sample-invariant constants of one controller (kernel parameter, lives in the constant bank).
struct DevCfg {
    int H, param, S, ga, method, N, P, np;  // np = parameters per leg
    float mass, inv_mass, mg;               // mg = float(mass * 9.81)
    float I[9], Iinv[9];
    float mu, fzmin, fzmax;
    float dts[QPMPC_MAX_H];
    float q[12];
    float duty, mpc_dt;
    float sig_rs[3], sig_mppi;
    int nfreq;
    float freqs[QPMPC_MAX_FREQ];
    int cem_elite, cem_sqrt;
    float cem_lo, cem_hi;
    float bounds[MAX_SPLINES + 1];  // linspace(0, H, S+1), centroidal_nmpc_jax.py:186 / 210
};

// jnp scalar indexing into a 1-D array of `size`: a negative index wraps once, then clamps.
__device__ __forceinline__ int jidx(int i, int size) {
    i = i < 0 ? i + size : i;
    return min(max(i, 0), size - 1);
}

// index of the spline chunk `step` falls in (centroidal_nmpc_jax.py:186-188 / 210-212)
__device__ __forceinline__ int chunk_index(const DevCfg &c, float step) {
    int idx = 0;
    for (int k = 1; k <= c.S; ++k) idx = (step >= c.bounds[k]) ? k : idx;
    return idx;
}

// normalised position inside the chunk (centroidal_nmpc_jax.py:190-193 / 214-217)
__device__ __forceinline__ float chunk_tau(const DevCfg &c, float step, float horizon_leg, int index) {
    return step / (horizon_leg / (float)c.S) - (float)index;
}

struct CubicBasis {
    float a, b, c, d;
};
__device__ __forceinline__ CubicBasis cubic_basis(float q) {  // centroidal_nmpc_jax.py:222-225
    const float q2 = q * q, q3 = q2 * q;
    return {2.f * q3 - 3.f * q2 + 1.f, q3 - 2.f * q2 + q, -2.f * q3 + 3.f * q2, q3 - q2};
}

// One leg's (fx, fy, fz) from its parameter block.  `par(j)` returns parameter j of this leg.
//   zero order: centroidal_nmpc_jax.py:259-268 (index = int16(step)); linear: :181-201; cubic: :204-257
//   (start_index = 10*index exactly as written upstream).
template <int PARAM, class Par>
__device__ __forceinline__ void spline_eval(const DevCfg &c, Par par, int istep, int index, float q,
                                            const CubicBasis &B, float f[3]) {
    const int np = c.np;
    if (PARAM == QPMPC_PARAM_ZERO_ORDER) {
        f[0] = par(jidx(istep, np));
        f[1] = par(jidx(istep + c.H, np));
        f[2] = par(jidx(istep + 2 * c.H, np));
    } else if (PARAM == QPMPC_PARAM_LINEAR) {
        const int shift = c.S + 1;
#pragma unroll
        for (int ax = 0; ax < 3; ++ax)
            f[ax] = (1.f - q) * par(jidx(index + shift * ax, np)) + q * par(jidx(index + shift * ax + 1, np));
    } else {
        const int s = 10 * index;
#pragma unroll
        for (int ax = 0; ax < 3; ++ax) {
            const int k = s + 4 * ax;
            const float p0 = par(jidx(k, np)), p1 = par(jidx(k + 1, np)), p2 = par(jidx(k + 2, np)),
                        p3 = par(jidx(k + 3, np));
            const float phi = 0.5f * ((p2 - p1) + (p1 - p0));
            const float phin = 0.5f * ((p3 - p2) + (p2 - p1));
            f[ax] = B.a * p1 + B.b * phi + B.c * p2 + B.d * phin;
        }
    }
}

// Gravity compensation, contact mask with the 30/10 x,y scale, then the box + friction-pyramid selects
// (centroidal_nmpc_jax.py:376-409 and 270-314).  The selects keep the BOUND when the comparison is false, so a
// NaN force (zero legs in stance: mg/0 * 0) collapses to the bound exactly as upstream.
__device__ __forceinline__ void shape_forces(const DevCfg &c, float f[4][3], const float st[4], float ref_fz) {
#pragma unroll
    for (int l = 0; l < 4; ++l) {
        float fz = (ref_fz + f[l][2]) * st[l];
        float fx = f[l][0] * st[l] * (1.0f / 3.0f);
        float fy = f[l][1] * st[l] * (1.0f / 3.0f);
        fz = fz > c.fzmin ? fz : c.fzmin;
        fz = fz < c.fzmax ? fz : c.fzmax;
        const float lim = c.mu * fz;
        fx = fx > -lim ? fx : -lim;
        fx = fx < lim ? fx : lim;
        fy = fy > -lim ? fy : -lim;
        fy = fy < lim ? fy : lim;
        f[l][0] = fx;
        f[l][1] = fy;
        f[l][2] = fz;
    }
}

// x[0:12] += fd(x, f, stance) * dt  (centroidal_model_jax.py:93-174).  The Euler-rate map is the closed form of
// inv(conj_euler_rates) @ w (upstream inverts the 3x3 with a general adjugate; det = cos(pitch), so both go
// non-finite at pitch = +-90 deg and the cost saturates to 1e6 either way).
__device__ __forceinline__ void srbd_step(const DevCfg &c, float x[12], const float *feet /*[12]*/,
                                          const float f[4][3], const float st[4], float dt) {
    float tx = 0.f, ty = 0.f, tz = 0.f, lx = 0.f, ly = 0.f, lz = 0.f;
#pragma unroll
    for (int l = 0; l < 4; ++l) {
        const float fx = f[l][0] * st[l], fy = f[l][1] * st[l], fz = f[l][2] * st[l];
        lx += fx;
        ly += fy;
        lz += fz;
        const float rx = feet[3 * l] - x[0], ry = feet[3 * l + 1] - x[1], rz = feet[3 * l + 2] - x[2];
        tx += ry * fz - rz * fy;
        ty += rz * fx - rx * fz;
        tz += rx * fy - ry * fx;
    }
    float sr, cr, sp, cp, sy, cy;
    sincosf(x[6], &sr, &cr);
    sincosf(x[7], &sp, &cp);
    sincosf(x[8], &sy, &cy);
    const float wx = x[9], wy = x[10], wz = x[11];
    // euler rates
    const float yaw_rate = (sr * wy + cr * wz) / cp;
    const float pitch_rate = cr * wy - sr * wz;
    const float roll_rate = wx + sp * yaw_rate;
    // b_R_w @ tau
    const float bx = cp * cy * tx + cp * sy * ty - sp * tz;
    const float by = (sr * sp * cy - cr * sy) * tx + (sr * sp * sy + cr * cy) * ty + sr * cp * tz;
    const float bz = (cr * sp * cy + sr * sy) * tx + (cr * sp * sy - sr * cy) * ty + cr * cp * tz;
    // w x (I w)
    const float Iwx = c.I[0] * wx + c.I[1] * wy + c.I[2] * wz;
    const float Iwy = c.I[3] * wx + c.I[4] * wy + c.I[5] * wz;
    const float Iwz = c.I[6] * wx + c.I[7] * wy + c.I[8] * wz;
    const float gx = bx - (wy * Iwz - wz * Iwy);
    const float gy = by - (wz * Iwx - wx * Iwz);
    const float gz = bz - (wx * Iwy - wy * Iwx);
    const float ax = c.Iinv[0] * gx + c.Iinv[1] * gy + c.Iinv[2] * gz;
    const float ay = c.Iinv[3] * gx + c.Iinv[4] * gy + c.Iinv[5] * gz;
    const float az = c.Iinv[6] * gx + c.Iinv[7] * gy + c.Iinv[8] * gz;
    // explicit Euler, all derivatives from the OLD state
    x[0] += x[3] * dt;
    x[1] += x[4] * dt;
    x[2] += x[5] * dt;
    x[3] += (c.inv_mass * lx) * dt;
    x[4] += (c.inv_mass * ly) * dt;
    x[5] += (c.inv_mass * lz - 9.81f) * dt;
    x[6] += roll_rate * dt;
    x[7] += pitch_rate * dt;
    x[8] += yaw_rate * dt;
    x[9] += ax * dt;
    x[10] += ay * dt;
    x[11] += az * dt;
}

// (x+ - ref)^T Q (x+ - ref) with the diagonal Q of centroidal_nmpc_jax.py:118-130 (feet weights are zero).
__device__ __forceinline__ float stage_cost(const DevCfg &c, const float x[12], const float *ref) {
    float acc = 0.f;
#pragma unroll
    for (int k = 0; k < 12; ++k) {
        const float e = x[k] - ref[k];
        acc += (e * c.q[k]) * e;
    }
    return acc;
}

}  // namespace qp
