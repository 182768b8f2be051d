// Device-side single-rigid-body rollout: controller constants, spline chunk indexing and the generic spline evaluation.
// From-scratch CUDA restatement of the math of (paths relative to /root/reference/quadruped_pympc):
//   controllers/sampling/centroidal_model_jax.py:93-174            fd + explicit Euler
//   controllers/sampling/centroidal_nmpc_jax.py:181-268            zero-order / linear / cubic parametrisations
//   controllers/sampling/centroidal_nmpc_jax.py:270-314            enforce_force_constraints
//   controllers/sampling/centroidal_nmpc_jax.py:316-496            compute_rollout (given contact sequence)
//   controllers/sampling/centroidal_nmpc_jax_gait_adaptive.py:326-501  compute_rollout (per-sample gait)
//   helpers/periodic_gait_generator_jax.py:68-89,136-151           trot clock -> contact sequence
// DevCfg, the chunk index / position rules and the generic spline evaluation in the reference's operation order (the
// fused finalize uses it for the first-step GRF); rollout.cuh builds the hot loops (hoisted chunk coefficients, force
// shaping, body step, stage cost) on the same formulas, qpmpc.cu holds the quad step.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "../../include/qpmpc.h"

namespace qp {

constexpr int MAX_SPLINES = 16;

// sin and cos of an Euler angle without control flow: three-term Cody-Waite reduction by pi/2 and the two degree-7/8
// minimax polynomials on [-pi/4, pi/4] (<= 1.6 ulp, 9.3e-8 absolute over |a| < 1e5 against float64, the class of the
// library sincosf; checked by tests/test_sincos_fast.py on the host arithmetic and by the GPU parity tests).
// The library sincosf carries an out-of-line Payne-Hanek path behind a branch: three of them per step cut the step
// body into seven basic blocks and cost BSSY/BSYNC plus instruction-fetch stalls (profiles/r4a_ncu_source_c5.txt).
// |a| >= 105615 (the library's own switch point), NaN and Inf are evaluated at 0.  For roll and pitch such a rollout
// is already lost -- the angle itself stays in the state, its cost term (weights 500) is > 5e12 or NaN (-> 1e6 by the
// saturation).  Yaw has weight 0 in the reference's Q: a rollout whose yaw alone runs past 1e5 rad keeps a finite cost
// computed with sin = 0, cos = 1 where the library's Payne-Hanek path would still be exact.  Reaching 1e5 rad in a
// 0.24 s horizon takes a yaw rate of 4e5 rad/s, i.e. angular-velocity cost terms (weights 20..50) above 1e12: not a
// rollout any update listens to, but a documented divergence.
__device__ __forceinline__ void sincos_fast(float a, float *sp, float *cp) {
    const float b = fabsf(a) < 105615.0f ? a : 0.0f;
    const int q = __float2int_rn(b * 0.636619772f);
    const float k = (float)q;
    float r = fmaf(k, -1.5707962513e+0f, b);
    r = fmaf(k, -7.5497894159e-8f, r);
    r = fmaf(k, -5.3903029535e-15f, r);
    const float r2 = r * r;
    float c = fmaf(r2, 2.443315711809948e-5f, -1.388731625493765e-3f);
    c = fmaf(c, r2, 4.166664568298827e-2f);
    c = fmaf(c * r2, r2, fmaf(-0.5f, r2, 1.0f));
    float s = fmaf(r2, -1.9515295891e-4f, 8.3321608736e-3f);
    s = fmaf(s, r2, -1.6666654611e-1f);
    s = fmaf(s * r2, r, r);
    const bool odd = q & 1;
    const float ss = odd ? c : s, cc = odd ? s : c;
    *sp = (q & 2) ? -ss : ss;
    *cp = ((q + 1) & 2) ? -cc : cc;
}

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

// Force shaping (shape_leg), the body step and the stage cost live in rollout.cuh / qpmpc.cu (quad_leg_step).

}  // namespace qp
