// Thread-per-rollout SRBD integration for the throughput regime (batched robots), any horizon / spline count.
// The horizon loop is NOT unrolled on purpose: an unrolled 12-step body is ~100 KB of straight-line SASS, and with 12
// warps per SM at different program counters the 32 KB L1.5 instruction cache thrashes (measured: 25 % of the issue
// slots lost to stall_no_inst, profiles/r1g).  Instead the loop body stays small and everything step-dependent is
// either a per-robot table in shared memory (contact flags, gravity share, spline basis) or hoisted: the Hermite /
// linear coefficients of the current spline chunk of each leg sit in registers and are reloaded only when the chunk
// changes (twice per rollout at num_splines = 2).
// Math restated from
//   controllers/sampling/centroidal_model_jax.py:93-174, controllers/sampling/centroidal_nmpc_jax.py:181-496,
//   controllers/sampling/centroidal_nmpc_jax_gait_adaptive.py:326-501, helpers/periodic_gait_generator_jax.py:68-89.
// Deviations from the reference's operation order, all at float32 round-off level (DESIGN.md "Numerics"):
//   * the Euler-rate division and the per-sample-gait chunk position use the SFU reciprocal (2 ulp);
//   * fd's second multiplication of the already masked forces by the contact flag is explicit for fz (a non-zero
//     grf_min makes the clamp lift a swing leg's fz) and folded into the friction limit for fx, fy (shape_leg);
//   * sin/cos are sincos_fast (srbd.cuh): the library sincosf's accuracy class (1.6 ulp) without its slow-path
//     branch; the SFU versions (5e-7 absolute) moved costs by ~1e-6 relative, which the temperature-1 softmin of
//     the update amplifies beyond the 1e-3 bar.
#pragma once
#include "srbd.cuh"

namespace qp {

// contact-masked, friction-limited force of one leg: centroidal_nmpc_jax.py:382-402 then 270-314.
// st in {0,1}; st3 = st / 3.  max/min keep the BOUND for a NaN input exactly like the reference's selects.
__device__ __forceinline__ void shape_leg(const DevCfg &c, float f[3], float st, float st3, float ref_fz) {
    float fz = (ref_fz + f[2]) * st;
    float fx = f[0] * st3, fy = f[1] * st3;
    fz = fminf(fmaxf(fz, c.fzmin), c.fzmax);
    // friction limit masked by the contact flag: a swing leg clamps to [-0, 0], which is what fd's second multiplication
    // of fx, fy by the flag gives upstream -- also for a NaN parameter or a negative grf_min (lim = mu fz < 0 there)
    const float lim = (c.mu * st) * fz;
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
    sincos_fast(x[6], &sr, &cr);
    sincos_fast(x[7], &sp, &cp);
    sincos_fast(x[8], &sy, &cy);
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


// Coefficients of one spline chunk of one leg, read from the rollout's parameter column (col[j * STRIDE] = parameter j)
// when the chunk index changes.  cubic: {p1, phi, p2, phi_next} per axis (centroidal_nmpc_jax.py:219-249,
// start_index = 10 * index exactly as written upstream); linear: the two knots of the segment (:196-199).
// jidx reproduces jnp's scalar indexing (one negative wrap, then clamp) for chunk indices past the last knot.
template <int PARAM, int STRIDE>
__device__ __forceinline__ void load_chunk(const DevCfg &c, const float *col, int leg, int index, float k[3][4]) {
    const int np = c.np;
    const float *pl = col + leg * np * STRIDE;
    if (PARAM == QPMPC_PARAM_CUBIC) {
        const int s = 10 * index;
#pragma unroll
        for (int ax = 0; ax < 3; ++ax) {
            const int j = s + 4 * ax;
            const float p0 = pl[jidx(j, np) * STRIDE], p1 = pl[jidx(j + 1, np) * STRIDE],
                        p2 = pl[jidx(j + 2, np) * STRIDE], p3 = pl[jidx(j + 3, np) * STRIDE];
            k[ax][0] = p1;
            k[ax][1] = 0.5f * ((p2 - p1) + (p1 - p0));
            k[ax][2] = p2;
            k[ax][3] = 0.5f * ((p3 - p2) + (p2 - p1));
        }
    } else {
        const int shift = c.S + 1;
#pragma unroll
        for (int ax = 0; ax < 3; ++ax) {
            k[ax][0] = pl[jidx(index + shift * ax, np) * STRIDE];
            k[ax][1] = pl[jidx(index + shift * ax + 1, np) * STRIDE];
        }
    }
}

// basis = {a, b, c, d} of centroidal_nmpc_jax.py:222-225 (cubic) or {1 - q, q, -, -} (linear)
template <int PARAM>
__device__ __forceinline__ void eval_chunk(const float k[3][4], const float4 B, float f[3]) {
#pragma unroll
    for (int ax = 0; ax < 3; ++ax) {
        if (PARAM == QPMPC_PARAM_CUBIC)
            f[ax] = B.x * k[ax][0] + B.y * k[ax][1] + B.z * k[ax][2] + B.w * k[ax][3];
        else
            f[ax] = B.x * k[ax][0] + B.y * k[ax][1];
    }
}

template <int PARAM>
__device__ __forceinline__ float4 chunk_basis(float q) {
    if (PARAM == QPMPC_PARAM_CUBIC) {
        const CubicBasis B = cubic_basis(q);
        return make_float4(B.a, B.b, B.c, B.d);
    }
    return make_float4(1.f - q, q, 0.f, 0.f);
}

// spline chunk of a (non-negative) step: the number of knots linspace(0, H, S+1)[1..S] at or below it
// (centroidal_nmpc_jax.py:186-188 / 210-212).  b1, b2 = the first two knots held in registers: at the shipped
// num_splines = 2 no constant-bank lookups remain in the step loop.
__device__ __forceinline__ int chunk_of(const DevCfg &c, float stepf, float b1, float b2) {
    if (c.S <= 2) return (stepf >= b1 ? 1 : 0) + ((c.S == 2 && stepf >= b2) ? 1 : 0);
    int idx = 0;
    for (int j = 1; j <= c.S; ++j) idx += (stepf >= c.bounds[j]) ? 1 : 0;
    return idx;
}

struct Problem {
    float x[12], feet[12], ref[12];
};
__device__ __forceinline__ void load_problem(Problem &pb, const float *s_state, const float *s_ref) {
#pragma unroll
    for (int k = 0; k < 12; ++k) {
        pb.x[k] = s_state[k];
        pb.feet[k] = s_state[12 + k];
        pb.ref[k] = s_ref[k];
    }
}

// Per-robot step tables of the fixed-gait rollout (identical for every sample of the robot).
struct FixedTables {
    const float *st;     // [H][8]  {contact[4], contact[4] / 3}
    const float *reffz;  // [H]     mg / legs in stance
    const float *basis;  // [H][4]  chunk_basis at step n
    const int *index;    // [H]     spline chunk of step n
};

// Fixed contact sequence: centroidal_nmpc_jax.py:316-496.
template <int PARAM, int STRIDE>
__device__ __forceinline__ float rollout_fixed(const DevCfg &c, const float *s_state, const float *s_ref,
                                               const FixedTables &tb, const float *col) {
    Problem pb;
    load_problem(pb, s_state, s_ref);
    const int H = c.H, hs = H * STRIDE;
    float cost = 0.f;
    float k[4][3][4];
    int cur = -1;
    const float *pz = col;  // zero order: row of step n
#pragma unroll 1
    for (int n = 0; n < H; ++n) {
        float4 B = make_float4(0.f, 0.f, 0.f, 0.f);
        if (PARAM != QPMPC_PARAM_ZERO_ORDER) {
            const int idx = tb.index[n];
            if (idx != cur) {  // uniform over the block; twice per rollout at num_splines = 2
#pragma unroll
                for (int l = 0; l < 4; ++l) load_chunk<PARAM, STRIDE>(c, col, l, idx, k[l]);
                cur = idx;
            }
            B = *reinterpret_cast<const float4 *>(tb.basis + 4 * n);
        }
        const float4 st4 = *reinterpret_cast<const float4 *>(tb.st + 8 * n);
        const float4 st34 = *reinterpret_cast<const float4 *>(tb.st + 8 * n + 4);
        const float st[4] = {st4.x, st4.y, st4.z, st4.w}, st3[4] = {st34.x, st34.y, st34.z, st34.w};
        const float ref_fz = tb.reffz[n];
        float f[4][3];
#pragma unroll
        for (int l = 0; l < 4; ++l) {
            if (PARAM == QPMPC_PARAM_ZERO_ORDER) {  // p[n], p[n + H], p[n + 2H] of the leg's block (:259-268)
                const float *p = pz + l * 3 * hs;
                f[l][0] = p[0];
                f[l][1] = p[hs];
                f[l][2] = p[2 * hs];
            } else {
                eval_chunk<PARAM>(k[l], B, f[l]);
            }
            shape_leg(c, f[l], st[l], st3[l], ref_fz);
        }
        body_step(c, pb.x, pb.feet, pb.ref, f, st, c.dts[n], cost);
        pz += STRIDE;
    }
    return cost;
}

// Per-sample gait: centroidal_nmpc_jax_gait_adaptive.py:326-501 with the trot clock of
// periodic_gait_generator_jax.py:68-89 (restart at t >= 1, advance by mpc_dt * f, stance while t < duty).
// s_mgk[k] = mg / k for k = 0..4 (k = 0 -> +inf as upstream).
template <int PARAM, int STRIDE>
__device__ __forceinline__ float rollout_gait_adaptive(const DevCfg &c, const float *s_state, const float *s_ref,
                                                       const float *s_mgk, const float *s_timing, float freq,
                                                       const float *col) {
    const int H = c.H, hs = H * STRIDE;
    uint32_t bits[4];
    float rq[4];
    const float inc = __fmul_rn(c.mpc_dt, freq);
#pragma unroll
    for (int l = 0; l < 4; ++l) {
        float t = s_timing[l];
        uint32_t b = 0;
        for (int n = 0; n < H; ++n) {
            t = (t >= 1.0f) ? 0.0f : t;
            t = __fadd_rn(t, inc);
            b |= (t < c.duty ? 1u : 0u) << n;
        }
        bits[l] = b;
        // horizon_leg = sum(contact) + 1 (:345-348); position in chunk = step / (horizon_leg / S) - index
        rq[l] = __fdividef((float)c.S, (float)(__popc(b) + 1));
    }
    Problem pb;
    load_problem(pb, s_state, s_ref);
    const float b1 = c.bounds[1], b2 = c.bounds[2];
    float cost = 0.f;
    float k[4][3][4];
    int ncnt[4] = {-1, -1, -1, -1}, cur[4] = {-1, -1, -1, -1};  // ..._gait_adaptive.py:339
#pragma unroll 1
    for (int n = 0; n < H; ++n) {
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
                const float *p = col + l * 3 * hs + step * STRIDE;
                f[l][0] = p[0];
                f[l][1] = p[hs];
                f[l][2] = p[2 * hs];
            } else {
                const float stepf = (float)step;
                const int idx = chunk_of(c, stepf, b1, b2);
                if (idx != cur[l]) {
                    load_chunk<PARAM, STRIDE>(c, col, l, idx, k[l]);
                    cur[l] = idx;
                }
                eval_chunk<PARAM>(k[l], chunk_basis<PARAM>(stepf * rq[l] - (float)idx), f[l]);
            }
            shape_leg(c, f[l], st[l], st[l] * (1.0f / 3.0f), ref_fz);
        }
        body_step(c, pb.x, pb.feet, pb.ref, f, st, c.dts[n], cost);
    }
    const float df = freq - 1.3f;
    return cost + df * 100.f * df;  // :500
}

}  // namespace qp
