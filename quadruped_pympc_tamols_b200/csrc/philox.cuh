// Philox4x32-10 counter-based generator + Box-Muller, written so that the fused solve kernels and the
// noise-dump kernel produce bit-identical values (explicit _rn intrinsics and SFU instructions only: nothing here
// depends on the compiler's FMA contraction choices).
//
// Replaces the reference's jax.random.normal / uniform / choice draws
// (quadruped_pympc/controllers/sampling/centroidal_nmpc_jax.py:647-677, 806-812, 951-958 and
//  centroidal_nmpc_jax_gait_adaptive.py:687-692, 834-838, 994-999).  jax's threefry bit stream cannot be
// reproduced without jax, so parity is checked by handing the SAME noise to the oracle (qpmpc_generate_noise).
//
// Cost per group of four normals: 10 Philox rounds (2 IMAD.WIDE + 2 LOP3 each) + 4 int->float conversions with one
// FMA each + 2 x {LG2, FMUL, SQRT} for the radii (the standard deviation is folded into the square root's argument)
// + 2 x {FMUL, SIN, COS} + 4 FMUL: about 75 instructions, 10 of them on the SFU.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace qp {

__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
        c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
        k.x += 0x9E3779B9u;
        k.y += 0xBB67AE85u;
    }
    return c;
}

enum : uint32_t { STREAM_NORMAL = 0u, STREAM_UNIFORM = 1u, STREAM_FREQ = 2u };

struct NoiseKey {
    uint64_t seed, offset;
};

__device__ __forceinline__ uint4 draw(const NoiseKey &nk, uint32_t stream, uint32_t group, uint32_t sample,
                                      uint32_t robot) {
    const uint4 ctr = make_uint4(group | (stream << 28), sample, robot, (uint32_t)nk.offset);
    const uint2 key = make_uint2((uint32_t)nk.seed, (uint32_t)(nk.seed >> 32) ^ (uint32_t)(nk.offset >> 32));
    return philox4x32_10(ctr, key);
}

// u in [0,1): 24 random bits (the uniform block of random_sampling: U[-a, a) like jax.random.uniform)
__device__ __forceinline__ float u01_half_open(uint32_t x) { return __fmul_rn((float)(x >> 8), 5.9604644775390625e-08f); }
// u in (0,1]: (x + 0.5) / 2^32 in float; never 0, may round to 1 (then log2 u = 0: a zero radius)
__device__ __forceinline__ float u01_pos(uint32_t x) {
    return __fmaf_rn(__uint2float_rn(x), 2.3283064365386963e-10f, 1.1641532182693481e-10f);
}
// u in [0,1]: x / 2^32 in float (a phase of exactly one turn equals a phase of zero)
__device__ __forceinline__ float u01_turn(uint32_t x) { return __fmul_rn(__uint2float_rn(x), 2.3283064365386963e-10f); }

__device__ __forceinline__ float sfu_sqrt(float x) {
    float y;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float sfu_lg2(float x) {
    float y;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

constexpr float NEG_2LN2 = -1.3862943611198906f;  // sigma^2 * NEG_2LN2 * log2(u) = sigma^2 * (-2 ln u)

// four N(0, sigma^2) draws from one Philox block; k = sigma^2 * NEG_2LN2 (Box-Muller, both branches of both pairs)
__device__ __forceinline__ float4 normal4_scaled(uint4 x, float k) {
    const float r0 = sfu_sqrt(__fmul_rn(k, sfu_lg2(u01_pos(x.x))));
    const float r1 = sfu_sqrt(__fmul_rn(k, sfu_lg2(u01_pos(x.z))));
    float s0, c0, s1, c1;
    __sincosf(__fmul_rn(6.28318530717958647692f, u01_turn(x.y)), &s0, &c0);
    __sincosf(__fmul_rn(6.28318530717958647692f, u01_turn(x.w)), &s1, &c1);
    return make_float4(__fmul_rn(r0, c0), __fmul_rn(r0, s0), __fmul_rn(r1, c1), __fmul_rn(r1, s1));
}

}  // namespace qp
