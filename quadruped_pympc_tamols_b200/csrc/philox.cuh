// Random bits of the sampling controllers: Philox4x32 (counter-based) seeds, xoroshiro64+ bursts, Box-Muller on the SFU.
//
// Replaces the reference's jax.random.normal / uniform / choice draws
// (quadruped_pympc/controllers/sampling/centroidal_nmpc_jax.py:647-677, 806-812, 951-958 and
//  centroidal_nmpc_jax_gait_adaptive.py:687-692, 834-838, 994-999).  jax's threefry bit stream cannot be
// reproduced without jax, so parity is checked by handing the SAME noise to the oracle (qpmpc_generate_noise).
//
// Noise stream "v3" (this repo's own definition; qpmpc_version() carries the tag).  The P parameters of a rollout are
// drawn in BLOCKS of 12 values (P is a multiple of 12 for every parametrisation: 12 H, 12 (S + 1), 48 S).  Blocks 2 m
// and 2 m + 1 of sample i of robot r under key (seed, offset) share ONE counter-based seed,
//     (a, b, c, d) = Philox4x32-7(counter = (m | stream << 28, i, r, offset), key = seed)   -- random access, 128 bits
// block 2 m starts the 64-bit generator xoroshiro64+ from (a, b), block 2 m + 1 from (c, d), and
//     w_0 .. w_11 = its first 12 outputs                                                     -- a short sequential burst
// pair j of the block is the Box-Muller transform of (w_2j, w_2j+1) (uniform rows: value j from w_j).
// Why: measured on B200 (tools/ubench.cu, profiles/r5_ubench.txt) the 32 x 32 -> 64 multiply IMAD.WIDE issues once
// per ~4.5 cycles per scheduler and sits on the SAME pipe as the FP32 work; v1 (Philox4x32-10 per four values) cost 90
// FMA-pipe cycles per group, 3,240 per zero-order rollout, about as much as the whole SRBD integration.  Seven rounds
// (the Crush-resistant minimum of the Random123 paper) once per TWENTY-FOUR values cost 63 (v2 seeded every twelve);
// the generator steps are LOP3 / SHF / IADD on the otherwise lightly loaded ALU pipe, six per word (xoroshiro64: the
// 2 x 32-bit engine of Blackman & Vigna, constants 26 / 9 / 13, with the '+' output -- the weak low bits of a '+'
// scrambler are the ones the 24-bit float conversion rounds away; v2's xoshiro128+ took seven and four state registers
// per row).  A rollout that walks its blocks in order (k_rollout_pair) keeps the second half of the seed for the odd
// block; everything else (parameter columns of the small kernels, re-drawn rows) seeds per block and discards a half.
// Everything below uses explicit _rn intrinsics and SFU instructions only: nothing depends on the compiler's FMA
// contraction choices, so the scalar and the packed (two rollouts per thread, f32x2) generators are bit-identical.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace qp {

template <int ROUNDS>
__device__ __forceinline__ uint4 philox4x32(uint4 c, uint2 k) {
#pragma unroll
    for (int r = 0; r < ROUNDS; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
        c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
        k.x += 0x9E3779B9u;
        k.y += 0xBB67AE85u;
    }
    return c;
}

enum : uint32_t { STREAM_NORMAL = 0u, STREAM_UNIFORM = 1u, STREAM_FREQ = 2u };
constexpr int NOISE_BLOCK = 12;  // values per noise block

struct NoiseKey {
    uint64_t seed, offset;
};

// one Philox4x32-10 block (the step-frequency draw: one word per rollout)
__device__ __forceinline__ uint4 draw(const NoiseKey &nk, uint32_t stream, uint32_t group, uint32_t sample,
                                      uint32_t robot) {
    const uint4 ctr = make_uint4(group | (stream << 28), sample, robot, (uint32_t)nk.offset);
    const uint2 key = make_uint2((uint32_t)nk.seed, (uint32_t)(nk.seed >> 32) ^ (uint32_t)(nk.offset >> 32));
    return philox4x32<10>(ctr, key);
}

// xoroshiro64+ on one half of the state a Philox4x32-7 block left: 6 ALU instructions per 32-bit word
struct BlockRng {
    uint32_t s0, s1;
};
// the seeds of blocks 2 m (first) and 2 m + 1 (second)
__device__ __forceinline__ void block_rng_both(const NoiseKey &nk, uint32_t stream, uint32_t m, uint32_t sample, uint32_t robot,
                                               BlockRng &first, BlockRng &second) {
    const uint4 ctr = make_uint4(m | (stream << 28), sample, robot, (uint32_t)nk.offset);
    const uint2 key = make_uint2((uint32_t)nk.seed, (uint32_t)(nk.seed >> 32) ^ (uint32_t)(nk.offset >> 32));
    const uint4 s = philox4x32<7>(ctr, key);
    first = BlockRng{s.x, s.y};
    second = BlockRng{s.z, s.w};
}
__device__ __forceinline__ BlockRng block_rng(const NoiseKey &nk, uint32_t stream, uint32_t block, uint32_t sample,
                                              uint32_t robot) {
    BlockRng first, second;
    block_rng_both(nk, stream, block >> 1, sample, robot, first, second);
    return (block & 1u) ? second : first;
}
__device__ __forceinline__ uint32_t rng_next(BlockRng &g) {
    const uint32_t out = g.s0 + g.s1;
    const uint32_t x = g.s1 ^ g.s0;
    g.s0 = __funnelshift_l(g.s0, g.s0, 26) ^ x ^ (x << 9);
    g.s1 = __funnelshift_l(x, x, 13);
    return out;
}

// u in [0,1): 24 random bits (the uniform block of random_sampling: U[-a, a) like jax.random.uniform)
__device__ __forceinline__ float u01_half_open(uint32_t x) { return __fmul_rn((float)(x >> 8), 5.9604644775390625e-08f); }
constexpr float U01_SCALE = 2.3283064365386963e-10f;   // 2^-32
constexpr float U01_HALF = 1.1641532182693481e-10f;    // 2^-33: u = (x + 0.5) / 2^32 in (0, 1], never 0
constexpr float TURN_SCALE = 1.4629180792671596e-09f;  // 2 pi / 2^32: phase of a 32-bit word

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
__device__ __forceinline__ float sfu_sin(float x) {
    float y;
    asm("sin.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float sfu_cos(float x) {
    float y;
    asm("cos.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

constexpr float NEG_2LN2 = -1.3862943611198906f;  // sigma^2 * NEG_2LN2 * log2(u) = sigma^2 * (-2 ln u)

// two N(0, sigma^2) draws from two words (Box-Muller, both branches); k = sigma^2 * NEG_2LN2
__device__ __forceinline__ void normal_pair(uint32_t w0, uint32_t w1, float k, float &z0, float &z1) {
    const float u = __fmaf_rn(__uint2float_rn(w0), U01_SCALE, U01_HALF);
    const float r = sfu_sqrt(__fmul_rn(k, sfu_lg2(u)));
    const float ph = __fmul_rn(__uint2float_rn(w1), TURN_SCALE);
    z0 = __fmul_rn(r, sfu_cos(ph));
    z1 = __fmul_rn(r, sfu_sin(ph));
}

}  // namespace qp
