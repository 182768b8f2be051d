// Two rollouts per thread on sm_100's packed FP32 (FFMA2 / FMUL2 / FADD2 on 64-bit register pairs): the SRBD step
// of the throughput regime (batched robots, zero-order parametrisation, fixed gait).
// Why packed: the thread-per-rollout kernel was bound by ISSUE SLOTS, not by the FMA pipe (profiles/r4c: 8,170 warp
// instructions per rollout for 4,380 algorithmic FLOP, issue 72 % busy, FMA pipe 40 %).  A packed instruction does the
// same arithmetic for rollouts A (.x) and B (.y) in one slot; per-robot values (feet, reference, contact flags,
// inertia, weights) enter as broadcast scalar operands (R.F32 / UR.F32 / immediate), so they cost no pair registers.
// Measured instruction rates behind the design: tools/ubench.cu -> profiles/r5_ubench.txt.
// Every operation is an explicit round-to-nearest intrinsic: the arithmetic does not depend on the compiler's
// contraction choices.  Math restated from
//   controllers/sampling/centroidal_model_jax.py:93-174 (fd + Euler step),
//   controllers/sampling/centroidal_nmpc_jax.py:259-314 (zero-order forces, force limits), :316-496 (rollout, cost).
#pragma once
#include "noise.cuh"

namespace qp {

__device__ __forceinline__ v2 vadd(v2 a, v2 b) { return __fadd2_rn(a, b); }
__device__ __forceinline__ v2 vadd(v2 a, float s) { return __fadd2_rn(a, f2(s)); }
__device__ __forceinline__ v2 vneg(v2 a) { return f2(-a.x, -a.y); }
__device__ __forceinline__ v2 vsub(v2 a, v2 b) { return __fadd2_rn(a, vneg(b)); }
__device__ __forceinline__ v2 vmul(v2 a, v2 b) { return __fmul2_rn(a, b); }
__device__ __forceinline__ v2 vmul(v2 a, float s) { return __fmul2_rn(a, f2(s)); }
__device__ __forceinline__ v2 vfma(v2 a, v2 b, v2 c) { return __ffma2_rn(a, b, c); }
__device__ __forceinline__ v2 vfma(v2 a, float s, v2 c) { return __ffma2_rn(a, f2(s), c); }
__device__ __forceinline__ v2 vfma(v2 a, v2 b, float s) { return __ffma2_rn(a, b, f2(s)); }
__device__ __forceinline__ v2 vfma(v2 a, float s, float d) { return __ffma2_rn(a, f2(s), f2(d)); }

// sincos_fast (srbd.cuh) for both halves: same reduction constants and polynomials; the quadrant comes from the
// magic-number rounding (b * 2/pi + 1.5 * 2^23) - 1.5 * 2^23 = rint(b * 2/pi) for |b * 2/pi| < 2^22, whose low mantissa
// bits are q -- one packed FMA and one packed add instead of F2I + I2F per half on the quarter-rate conversion pipe.
// (The product is not rounded before the rint, so a quadrant can differ from the scalar version's at an exact tie of
// b * 2/pi: the reduced argument then sits 1e-7 beyond pi/4, where both polynomials are still accurate to 1e-7.)
__device__ __forceinline__ void sincos_fast2(v2 a, v2 &s, v2 &c) {
    const v2 b = f2(fabsf(a.x) < 105615.0f ? a.x : 0.0f, fabsf(a.y) < 105615.0f ? a.y : 0.0f);
    const v2 t = vfma(b, 0.636619772f, 12582912.0f);  // (written fused: ptxas would contract the pair anyway)
    const v2 k = vadd(t, -12582912.0f);
    v2 r = vfma(k, -1.5707962513e+0f, b);
    r = vfma(k, -7.5497894159e-8f, r);
    r = vfma(k, -5.3903029535e-15f, r);
    const v2 r2 = vmul(r, r);
    v2 cc = vfma(r2, 2.443315711809948e-5f, -1.388731625493765e-3f);
    cc = vfma(cc, r2, 4.166664568298827e-2f);
    cc = vfma(vmul(cc, r2), r2, vfma(r2, -0.5f, 1.0f));
    v2 ss = vfma(r2, -1.9515295891e-4f, 8.3321608736e-3f);
    ss = vfma(ss, r2, -1.6666654611e-1f);
    ss = vfma(vmul(ss, r2), r, r);
    const uint32_t qx = __float_as_uint(t.x), qy = __float_as_uint(t.y);
    const bool ox = qx & 1u, oy = qy & 1u;
    const uint32_t sx = __float_as_uint(ox ? cc.x : ss.x), cx = __float_as_uint(ox ? ss.x : cc.x);
    const uint32_t sy = __float_as_uint(oy ? cc.y : ss.y), cy = __float_as_uint(oy ? ss.y : cc.y);
    // sin: negate when bit 1 of q is set; cos: when bit 1 of q + 1 is set = bit 1 xor bit 0
    s = f2(__uint_as_float(sx ^ ((qx << 30) & 0x80000000u)), __uint_as_float(sy ^ ((qy << 30) & 0x80000000u)));
    c = f2(__uint_as_float(cx ^ (((qx << 30) ^ (qx << 31)) & 0x80000000u)),
           __uint_as_float(cy ^ (((qy << 30) ^ (qy << 31)) & 0x80000000u)));
}

__device__ __forceinline__ v2 rcp2(v2 a) {
    v2 y;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y.x) : "f"(a.x));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y.y) : "f"(a.y));
    return y;
}

// One SRBD step of rollouts A and B.  f[3 leg + axis] = the raw control parameters of this step (best + eps); st / st3 /
// ref_fz / feet / ref are per-robot scalars.  Force shaping: centroidal_nmpc_jax.py:377-402 then :270-314 (max / min keep
// the BOUND for a NaN input exactly like the reference's selects); fd multiplies the masked forces by the contact flag a
// second time (centroidal_model_jax.py:111-116, 132-135): fz explicitly, fx and fy through the masked friction limit.
template <bool QSPARSE>
__device__ __forceinline__ void pair_step(const DevCfg &c, v2 x[12], const float feet[12], const float ref[12], v2 f[12],
                                          const float st[4], const float st3[4], float ref_fz, float dt, v2 &cost) {
    const float must[4] = {c.mu * st[0], c.mu * st[1], c.mu * st[2], c.mu * st[3]};
    // net force and, about the ORIGIN, the moment of the four ground reaction forces; the torque about the centre of
    // mass is then sum (r_foot - com) x f = sum r_foot x f - com x sum f: the foot positions enter as broadcast scalars
    // (cheaper operand form than three register pairs) and the twelve lever-arm subtractions become one cross product
    v2 lx = f2(0.f), ly = f2(0.f), lz = f2(0.f), tx = f2(0.f), ty = f2(0.f), tz = f2(0.f);
#pragma unroll
    for (int l = 0; l < 4; ++l) {
        v2 fz = vmul(vadd(f[3 * l + 2], ref_fz), st[l]);
        v2 fx = vmul(f[3 * l], st3[l]), fy = vmul(f[3 * l + 1], st3[l]);
        fz = f2(fminf(fmaxf(fz.x, c.fzmin), c.fzmax), fminf(fmaxf(fz.y, c.fzmin), c.fzmax));
        // friction limit already masked by the contact flag: a swing leg clamps to [-0, 0], which is what the second
        // multiplication by the flag in fd gives for fx, fy (also for a NaN parameter); fz keeps its own
        const v2 lim = vmul(fz, must[l]);
        fx = f2(fminf(fmaxf(fx.x, -lim.x), lim.x), fminf(fmaxf(fx.y, -lim.y), lim.y));
        fy = f2(fminf(fmaxf(fy.x, -lim.x), lim.x), fminf(fmaxf(fy.y, -lim.y), lim.y));
        fz = vmul(fz, st[l]);
        lx = vadd(lx, fx);
        ly = vadd(ly, fy);
        lz = vadd(lz, fz);
        const float px = feet[3 * l], py = feet[3 * l + 1], pz = feet[3 * l + 2];
        tx = vfma(fz, py, tx);
        tx = vfma(fy, -pz, tx);
        ty = vfma(fx, pz, ty);
        ty = vfma(fz, -px, ty);
        tz = vfma(fy, px, tz);
        tz = vfma(fx, -py, tz);
    }
    tx = vfma(vneg(x[1]), lz, vfma(x[2], ly, tx));
    ty = vfma(vneg(x[2]), lx, vfma(x[0], lz, ty));
    tz = vfma(vneg(x[0]), ly, vfma(x[1], lx, tz));
    v2 sr, cr, sp, cp, sy, cy;
    sincos_fast2(x[6], sr, cr);
    sincos_fast2(x[7], sp, cp);
    sincos_fast2(x[8], sy, cy);
    const v2 wx = x[9], wy = x[10], wz = x[11];
    // Euler-rate map, closed form of inv(conj_euler_rates) @ w (centroidal_model_jax.py:124-137), SFU reciprocal
    const v2 yaw_rate = vmul(vfma(sr, wy, vmul(cr, wz)), rcp2(cp));
    const v2 pitch_rate = vfma(cr, wy, vneg(vmul(sr, wz)));
    const v2 roll_rate = vfma(sp, yaw_rate, wx);
    // b_R_w @ tau (centroidal_model_jax.py:141-155) as three planar rotations, yaw then pitch then roll: the same matrix
    // product (b_R_w = Rx(roll)^T Ry(pitch)^T Rz(yaw)^T) in 12 operations instead of 26 for the nine entries + mat-vec
    const v2 u1 = vfma(cy, tx, vmul(sy, ty));
    const v2 v1 = vfma(cy, ty, vneg(vmul(sy, tx)));
    const v2 bx = vfma(cp, u1, vneg(vmul(sp, tz)));
    const v2 w2 = vfma(sp, u1, vmul(cp, tz));
    const v2 by = vfma(cr, v1, vmul(sr, w2));
    const v2 bz = vfma(cr, w2, vneg(vmul(sr, v1)));
    const v2 Iwx = vfma(wx, c.I[0], vfma(wy, c.I[1], vmul(wz, c.I[2])));
    const v2 Iwy = vfma(wx, c.I[3], vfma(wy, c.I[4], vmul(wz, c.I[5])));
    const v2 Iwz = vfma(wx, c.I[6], vfma(wy, c.I[7], vmul(wz, c.I[8])));
    const v2 gx = vfma(wz, Iwy, vfma(vneg(wy), Iwz, bx));
    const v2 gy = vfma(wx, Iwz, vfma(vneg(wz), Iwx, by));
    const v2 gz = vfma(wy, Iwx, vfma(vneg(wx), Iwy, bz));
    const v2 ax = vfma(gx, c.Iinv[0], vfma(gy, c.Iinv[1], vmul(gz, c.Iinv[2])));
    const v2 ay = vfma(gx, c.Iinv[3], vfma(gy, c.Iinv[4], vmul(gz, c.Iinv[5])));
    const v2 az = vfma(gx, c.Iinv[6], vfma(gy, c.Iinv[7], vmul(gz, c.Iinv[8])));
    x[0] = vfma(x[3], dt, x[0]);
    x[1] = vfma(x[4], dt, x[1]);
    x[2] = vfma(x[5], dt, x[2]);
    x[3] = vfma(vmul(lx, c.inv_mass), dt, x[3]);
    x[4] = vfma(vmul(ly, c.inv_mass), dt, x[4]);
    x[5] = vfma(vfma(lz, c.inv_mass, -9.81f), dt, x[5]);
    x[6] = vfma(roll_rate, dt, x[6]);
    x[7] = vfma(pitch_rate, dt, x[7]);
    x[8] = vfma(yaw_rate, dt, x[8]);
    x[9] = vfma(ax, dt, x[9]);
    x[10] = vfma(ay, dt, x[10]);
    x[11] = vfma(az, dt, x[11]);
    // stage cost on the new state: sum q_k (x_k - ref_k)^2 over the body states (centroidal_nmpc_jax.py:450-451, 491).
    // QSPARSE: q[0] = q[1] = q[8] = 0 as in the reference's Q (:118-130), so those three terms are not evaluated.
    // Upstream a zero weight times a non-finite state still makes the cost NaN (-> 1e6); nothing is lost here, because
    // x, y and yaw can only become non-finite together with vx, vy and roll (x += vx dt; yaw' and roll' share the
    // 1 / cos(pitch) factor), whose weights the host requires to be positive before it selects a QSPARSE kernel.
#pragma unroll
    for (int k = 0; k < 12; ++k) {
        if (QSPARSE && (k == 0 || k == 1 || k == 8)) continue;
        const v2 e = vadd(x[k], -ref[k]);
        cost = vfma(vmul(e, e), c.q[k], cost);
    }
}

}  // namespace qp
