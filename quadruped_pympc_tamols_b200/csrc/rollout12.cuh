// Fast rollout for the horizon the reference ships (mpc_params['horizon'] = 12, num_splines = 2, config.py:74,193):
// the 12 integration steps are fully unrolled, so every spline chunk index, every position inside a chunk and every
// parameter offset is a compile-time constant; the 12 body states, the 12 foot coordinates, the 12 reference values
// and the Hermite coefficients of the current chunk live in registers.  Same math as srbd.cuh (which stays the
// generic path for any other horizon / spline count), restated from
//   controllers/sampling/centroidal_model_jax.py:93-174, controllers/sampling/centroidal_nmpc_jax.py:181-496,
//   controllers/sampling/centroidal_nmpc_jax_gait_adaptive.py:326-501, helpers/periodic_gait_generator_jax.py:68-89.
// Deviations from the generic path, all far inside the 1e-4 cost tolerance (DESIGN.md "Numerics"):
//   * sin/cos through the SFU (MUFU.SIN/COS after a two-constant Cody-Waite reduction, abs. error ~5e-7) instead of
//     the 1-ulp library sincosf; the Euler-rate division and the GA chunk position use the SFU reciprocal;
//   * fd's second multiplication of the already masked forces by the contact flag is dropped for fx, fy (it is an
//     exact identity there) and kept for fz (a non-zero grf_min makes the clamp lift a swing leg's fz).
#pragma once
#include "srbd.cuh"

namespace qp {

constexpr int H12 = 12;

__device__ __forceinline__ void sincos_sfu(float x, float &s, float &c) {
    const float k = rintf(x * 0.15915494309189535f);
    float r = fmaf(k, -6.2831854820251465f, x);
    r = fmaf(k, 1.7484555314695172e-07f, r);
    s = __sinf(r);
    c = __cosf(r);
}

// contact-masked, friction-limited force of one leg: centroidal_nmpc_jax.py:382-402 then 270-314.
// st in {0,1}; st3 = st / 3.  max/min keep the BOUND for a NaN input exactly like the reference's selects.
__device__ __forceinline__ void shape_leg(const DevCfg &c, float f[3], float st, float st3, float ref_fz) {
    float fz = (ref_fz + f[2]) * st;
    float fx = f[0] * st3, fy = f[1] * st3;
    fz = fminf(fmaxf(fz, c.fzmin), c.fzmax);
    const float lim = c.mu * fz;
    f[0] = fminf(fmaxf(fx, -lim), lim);
    f[1] = fminf(fmaxf(fy, -lim), lim);
    f[2] = fz;
}

// x[0:12] += fd(x, f) * dt, cost += (x+ - ref)^T Q (x+ - ref).  `fzs` = fz masked once more by the contact flag.
__device__ __forceinline__ void body_step(const DevCfg &c, float x[12], const float feet[12], const float ref[12],
                                          const float f[4][3], const float st[4], float dt, float &cost) {
    float tx = 0.f, ty = 0.f, tz = 0.f, lx = 0.f, ly = 0.f, lz = 0.f;
#pragma unroll
    for (int l = 0; l < 4; ++l) {
        const float fx = f[l][0], fy = f[l][1], fz = f[l][2] * st[l];
        lx += fx;
        ly += fy;
        lz += fz;
        const float rx = feet[3 * l] - x[0], ry = feet[3 * l + 1] - x[1], rz = feet[3 * l + 2] - x[2];
        tx += ry * fz - rz * fy;
        ty += rz * fx - rx * fz;
        tz += rx * fy - ry * fx;
    }
    float sr, cr, sp, cp, sy, cy;
#ifdef QPMPC_SFU_TRIG
    sincos_sfu(x[6], sr, cr);
    sincos_sfu(x[7], sp, cp);
    sincos_sfu(x[8], sy, cy);
#else  // the softmin update needs costs good to ~1e-7 relative (DESIGN.md "Numerics"): 1-ulp sincosf
    sincosf(x[6], &sr, &cr);
    sincosf(x[7], &sp, &cp);
    sincosf(x[8], &sy, &cy);
#endif
    const float wx = x[9], wy = x[10], wz = x[11];
    const float yaw_rate = __fdividef(sr * wy + cr * wz, cp);
    const float pitch_rate = cr * wy - sr * wz;
    const float roll_rate = wx + sp * yaw_rate;
    const float bx = cp * cy * tx + cp * sy * ty - sp * tz;
    const float by = (sr * sp * cy - cr * sy) * tx + (sr * sp * sy + cr * cy) * ty + sr * cp * tz;
    const float bz = (cr * sp * cy + sr * sy) * tx + (cr * sp * sy - sr * cy) * ty + cr * cp * tz;
    const float Iwx = c.I[0] * wx + c.I[1] * wy + c.I[2] * wz;
    const float Iwy = c.I[3] * wx + c.I[4] * wy + c.I[5] * wz;
    const float Iwz = c.I[6] * wx + c.I[7] * wy + c.I[8] * wz;
    const float gx = bx - (wy * Iwz - wz * Iwy);
    const float gy = by - (wz * Iwx - wx * Iwz);
    const float gz = bz - (wx * Iwy - wy * Iwx);
    const float ax = c.Iinv[0] * gx + c.Iinv[1] * gy + c.Iinv[2] * gz;
    const float ay = c.Iinv[3] * gx + c.Iinv[4] * gy + c.Iinv[5] * gz;
    const float az = c.Iinv[6] * gx + c.Iinv[7] * gy + c.Iinv[8] * gz;
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
#pragma unroll
    for (int k = 0; k < 12; ++k) {
        const float e = x[k] - ref[k];
        cost += (e * c.q[k]) * e;
    }
}

// Coefficients of one spline chunk of one leg, read once per chunk from the thread's parameter column
// (col[j * stride] = parameter j of this rollout).  cubic: {p1, phi, p2, phi_next} per axis
// (centroidal_nmpc_jax.py:219-249, start_index = 10 * index); linear: {p[index + 3 ax], p[index + 3 ax + 1]} (:196-199).
template <int PARAM>
__device__ __forceinline__ void load_chunk(const float *col, int stride, int leg, int index, float k[3][4]) {
    if (PARAM == QPMPC_PARAM_CUBIC) {
        const float *p = col + (leg * 24 + 10 * index) * stride;
#pragma unroll
        for (int ax = 0; ax < 3; ++ax) {
            const float p0 = p[(4 * ax) * stride], p1 = p[(4 * ax + 1) * stride], p2 = p[(4 * ax + 2) * stride],
                        p3 = p[(4 * ax + 3) * stride];
            k[ax][0] = p1;
            k[ax][1] = 0.5f * ((p2 - p1) + (p1 - p0));
            k[ax][2] = p2;
            k[ax][3] = 0.5f * ((p3 - p2) + (p2 - p1));
        }
    } else {
        const float *p = col + (leg * 9 + index) * stride;
#pragma unroll
        for (int ax = 0; ax < 3; ++ax) {
            k[ax][0] = p[(3 * ax) * stride];
            k[ax][1] = p[(3 * ax + 1) * stride];
        }
    }
}

template <int PARAM>
__device__ __forceinline__ void eval_chunk(const float k[3][4], float q, float f[3]) {
    if (PARAM == QPMPC_PARAM_CUBIC) {
        const CubicBasis B = cubic_basis(q);
#pragma unroll
        for (int ax = 0; ax < 3; ++ax) f[ax] = B.a * k[ax][0] + B.b * k[ax][1] + B.c * k[ax][2] + B.d * k[ax][3];
    } else {
#pragma unroll
        for (int ax = 0; ax < 3; ++ax) f[ax] = (1.f - q) * k[ax][0] + q * k[ax][1];
    }
}

struct Problem12 {
    float x[12], feet[12], ref[12];
};
__device__ __forceinline__ void load_problem(Problem12 &pb, const float *s_state, const float *s_ref) {
#pragma unroll
    for (int k = 0; k < 12; ++k) {
        pb.x[k] = s_state[k];
        pb.feet[k] = s_state[12 + k];
        pb.ref[k] = s_ref[k];
    }
}

// Fixed contact sequence (centroidal_nmpc_jax.py:316-496), H = 12, S = 2.
// s_st [12][8]: per step {st[4], st[4] / 3}; s_reffz [12]: mg / legs in stance.
template <int PARAM>
__device__ __forceinline__ float rollout12_fixed(const DevCfg &c, const float *s_state, const float *s_ref,
                                                 const float *s_st, const float *s_reffz, const float *col, int stride) {
    Problem12 pb;
    load_problem(pb, s_state, s_ref);
    float cost = 0.f;
    float k[4][3][4];
#pragma unroll
    for (int n = 0; n < H12; ++n) {
        const int index = n >= 6 ? 1 : 0;                    // linspace(0, 12, 3) = {0, 6, 12}
        const float q = (float)n / 6.0f - (float)index;      // step / (horizon_leg / num_splines) - index
        if (PARAM != QPMPC_PARAM_ZERO_ORDER && (n == 0 || n == 6)) {
#pragma unroll
            for (int l = 0; l < 4; ++l) load_chunk<PARAM>(col, stride, l, index, k[l]);
        }
        const float4 st4 = *reinterpret_cast<const float4 *>(s_st + 8 * n);
        const float4 st34 = *reinterpret_cast<const float4 *>(s_st + 8 * n + 4);
        const float st[4] = {st4.x, st4.y, st4.z, st4.w}, st3[4] = {st34.x, st34.y, st34.z, st34.w};
        const float ref_fz = s_reffz[n];
        float f[4][3];
#pragma unroll
        for (int l = 0; l < 4; ++l) {
            if (PARAM == QPMPC_PARAM_ZERO_ORDER) {  // p[n], p[n + H], p[n + 2H] of the leg's block (:259-268)
#pragma unroll
                for (int ax = 0; ax < 3; ++ax) f[l][ax] = col[(l * 36 + ax * 12 + n) * stride];
            } else {
                eval_chunk<PARAM>(k[l], q, f[l]);
            }
            shape_leg(c, f[l], st[l], st3[l], ref_fz);
        }
        body_step(c, pb.x, pb.feet, pb.ref, f, st, c.dts[n], cost);
    }
    return cost;
}

// Per-sample gait (centroidal_nmpc_jax_gait_adaptive.py:326-501), H = 12, S = 2.
template <int PARAM>
__device__ __forceinline__ float rollout12_ga(const DevCfg &c, const float *s_state, const float *s_ref,
                                              const float *s_mgk, const float *s_timing, float freq, const float *col,
                                              int stride) {
    // trot clock of periodic_gait_generator_jax.py:68-89: restart at t >= 1, advance by mpc_dt * f, stance while t < duty
    uint32_t bits[4];
    float rq[4];
    const float inc = __fmul_rn(c.mpc_dt, freq);
#pragma unroll
    for (int l = 0; l < 4; ++l) {
        float t = s_timing[l];
        uint32_t b = 0;
#pragma unroll
        for (int n = 0; n < H12; ++n) {
            t = (t >= 1.0f) ? 0.0f : t;
            t = __fadd_rn(t, inc);
            b |= (t < c.duty ? 1u : 0u) << n;
        }
        bits[l] = b;
        // horizon_leg = sum(contact) + 1 (:345-348); position in chunk = step / (horizon_leg / 2) - index
        rq[l] = __fdividef(2.0f, (float)(__popc(b) + 1));
    }
    Problem12 pb;
    load_problem(pb, s_state, s_ref);
    float cost = 0.f;
    float k[4][3][4];
    int ncnt[4] = {-1, -1, -1, -1}, cur[4] = {-1, -1, -1, -1};
#pragma unroll
    for (int n = 0; n < H12; ++n) {
        float st[4], f[4][3];
        int nst = 0;
#pragma unroll
        for (int l = 0; l < 4; ++l) {
            const int on = (bits[l] >> n) & 1;
            st[l] = on ? 1.0f : 0.0f;
            nst += on;
            ncnt[l] += on;  // the leg's own step counter advances only in stance (:353-356)
        }
        const float ref_fz = s_mgk[nst];
#pragma unroll
        for (int l = 0; l < 4; ++l) {
            // a swing leg's spline value is multiplied by 0 below; any finite stand-in does, so clamp the counter at 0
            const int step = max(ncnt[l], 0);
            if (PARAM == QPMPC_PARAM_ZERO_ORDER) {
                const float *p = col + (l * 36 + step) * stride;
#pragma unroll
                for (int ax = 0; ax < 3; ++ax) f[l][ax] = p[(ax * 12) * stride];
            } else {
                const int index = step >= 6 ? 1 : 0;
                if (index != cur[l]) {  // at most twice per leg and rollout
                    load_chunk<PARAM>(col, stride, l, index, k[l]);
                    cur[l] = index;
                }
                eval_chunk<PARAM>(k[l], (float)step * rq[l] - (float)index, f[l]);
            }
            shape_leg(c, f[l], st[l], st[l] * (1.0f / 3.0f), ref_fz);
        }
        body_step(c, pb.x, pb.feet, pb.ref, f, st, c.dts[n], cost);
    }
    const float df = freq - 1.3f;
    return cost + df * 100.f * df;  // :500
}

}  // namespace qp
