// additional_random_parameters of the sampling controllers, drawn in blocks of 12 values (philox.cuh, stream v3).
//   random_sampling centroidal_nmpc_jax.py:647-677 (the two Gaussian thirds share ONE key upstream, hence the same
//   normals scaled by sigma[0] and sigma[1]; the last third is U(-sigma[2], sigma[2]));
//   mppi :806-812; cem_mppi :951-958 (normals * sigma vector).  Row 0 is always zero.
// Block b holds the values of ORDER indices 12 b .. 12 b + 11.  For the zero-order parametrisation the order is by
// horizon step -- block n = the twelve (leg, axis) forces of step n, parameter (3 leg + axis) H + n -- so that a
// rollout can draw the block of step n while it integrates step n and never store its parameter vector; for the
// spline parametrisations the order is the parameter index itself.
// Two generators with bit-identical values: the scalar one (any mode, used wherever a row is re-drawn: parameter
// columns of the small kernels, random-sampling winner, CEM elites, live MPPI samples, the noise dump) and the packed
// one (two rollouts per thread, f32x2 arithmetic) of k_rollout_pair.
#pragma once
#include "philox.cuh"
#include "srbd.cuh"

namespace qp {

// parameter index of value k of block b
__device__ __forceinline__ int param_of(const DevCfg &c, int b, int k) {
    return c.param == QPMPC_PARAM_ZERO_ORDER ? k * c.H + b : NOISE_BLOCK * b + k;
}

// Everything that depends only on the row i, resolved once per rollout.
struct RowNoise {
    int mode;         // 0: N(0, sigma^2) with scalar sigma; 1: U[-a, a); 2: N(0,1) * sigma_vec[p]; 3: zero row
    uint32_t stream, sample;
    float k;          // mode 0: sigma^2 * NEG_2LN2; mode 1: a; mode 2: NEG_2LN2
};

__device__ __forceinline__ RowNoise row_noise(const DevCfg &c, int i) {
    RowNoise rn;
    rn.stream = STREAM_NORMAL;
    rn.sample = (uint32_t)(i - 1);
    if (i == 0) {
        rn.mode = 3;
        rn.k = 0.f;
    } else if (c.method == QPMPC_METHOD_RANDOM) {
        const int third = c.N / 3;
        if (i < 1 + 2 * third) {
            const bool first = i < 1 + third;
            const float s = first ? c.sig_rs[0] : c.sig_rs[1];
            rn.mode = 0;
            rn.sample = (uint32_t)(first ? i - 1 : i - 1 - third);
            rn.k = __fmul_rn(__fmul_rn(s, s), NEG_2LN2);
        } else {
            rn.mode = 1;
            rn.stream = STREAM_UNIFORM;
            rn.sample = (uint32_t)(i - 1 - 2 * third);
            rn.k = c.sig_rs[2];
        }
    } else if (c.method == QPMPC_METHOD_MPPI) {
        rn.mode = 0;
        rn.k = __fmul_rn(__fmul_rn(c.sig_mppi, c.sig_mppi), NEG_2LN2);
    } else {
        rn.mode = 2;
        rn.k = NEG_2LN2;
    }
    return rn;
}

// e[0..11] = block b of row rn (sigma applied).  sigma_at(k) = sigma_vec[param_of(b, k)] (mode 2 only).
template <class SigmaAt>
__device__ __forceinline__ void noise_block(const RowNoise &rn, const NoiseKey &nk, int b, uint32_t robot, float e[NOISE_BLOCK],
                                            SigmaAt sigma_at) {
    if (rn.mode == 3) {
#pragma unroll
        for (int k = 0; k < NOISE_BLOCK; ++k) e[k] = 0.f;
        return;
    }
    BlockRng g = block_rng(nk, rn.stream, (uint32_t)b, rn.sample, robot);
    if (rn.mode == 1) {
        const float a = rn.k, w = __fmul_rn(2.0f, a);
#pragma unroll
        for (int k = 0; k < NOISE_BLOCK; ++k) e[k] = __fmaf_rn(u01_half_open(rng_next(g)), w, -a);
        return;
    }
#pragma unroll
    for (int m = 0; m < NOISE_BLOCK / 2; ++m) {
        const uint32_t w0 = rng_next(g), w1 = rng_next(g);
        normal_pair(w0, w1, rn.k, e[2 * m], e[2 * m + 1]);
    }
    if (rn.mode == 2) {
#pragma unroll
        for (int k = 0; k < NOISE_BLOCK; ++k) e[k] = __fmul_rn(e[k], sigma_at(k));
    }
}

// ---- packed generator: rows A (.x) and B (.y) of one thread -------------------------------------------------
__device__ __forceinline__ float2 f2(float a, float b) { return make_float2(a, b); }
__device__ __forceinline__ float2 f2(float a) { return make_float2(a, a); }
typedef float2 v2;

// par[k] = best[k] + eps[k] for rows A and B.  NS = true: both rows are scalar-sigma normals (mppi) or the zero row;
// NS = false: any mode per row, the unused variants are computed and deselected (no control flow: the caller's SRBD
// arithmetic shares the basic block).  sigma_at(k) = sigma_vec[param_of(b, k)] (mode 2), unused otherwise.
// The last multiplication of a value and the addition of `best` are ONE fused multiply-add, written out: ptxas
// contracts mul.rn.f32x2 + add.rn.f32x2 pairs regardless of the rounding modifier (12.9), so leaving them apart would
// not keep them apart.  par therefore carries one rounding fewer than best + eps of the dumped noise (<= 1 ulp of par).
// carry (optional, two elements): a caller that draws blocks 0, 1, 2, ... of the two rows IN ORDER passes the same
// array every time; the even block seeds both blocks of its pair and leaves the odd one's seed there (one Philox per
// 24 values instead of one per 12).
template <bool NS, class SigmaAt>
__device__ __forceinline__ void noise_block_pair(const RowNoise &ra, const RowNoise &rb, const NoiseKey &nk, int b,
                                                 uint32_t robot, const float best[NOISE_BLOCK], v2 par[NOISE_BLOCK],
                                                 SigmaAt sigma_at, BlockRng *carry = nullptr) {
    BlockRng ga, gb;
    if (carry == nullptr) {
        ga = block_rng(nk, ra.stream, (uint32_t)b, ra.sample, robot);
        gb = block_rng(nk, rb.stream, (uint32_t)b, rb.sample, robot);
    } else if ((b & 1) == 0) {
        block_rng_both(nk, ra.stream, (uint32_t)b >> 1, ra.sample, robot, ga, carry[0]);
        block_rng_both(nk, rb.stream, (uint32_t)b >> 1, rb.sample, robot, gb, carry[1]);
    } else {
        ga = carry[0];
        gb = carry[1];
    }
    const float2 k2 = f2(ra.k, rb.k);
    const bool ua = ra.mode == 1, ub = rb.mode == 1;
    const float2 wid = f2(__fmul_rn(2.0f, ra.k), __fmul_rn(2.0f, rb.k)), lo = f2(-ra.k, -rb.k);
    // zero row (mode 3): NS relies on k = 0 -> radius -0 -> best + (+-0) = best exactly; the general path multiplies
    // by 0 where the other modes multiply by 1 (exact) or by sigma_vec[p]
    const float fa = ra.mode == 3 ? 0.0f : 1.0f, fb = rb.mode == 3 ? 0.0f : 1.0f;
#pragma unroll
    for (int m = 0; m < NOISE_BLOCK / 2; ++m) {
        const uint32_t a0 = rng_next(ga), a1 = rng_next(ga), b0 = rng_next(gb), b1 = rng_next(gb);
        const float2 u = __ffma2_rn(f2(__uint2float_rn(a0), __uint2float_rn(b0)), f2(U01_SCALE), f2(U01_HALF));
        const float2 ph = __fmul2_rn(f2(__uint2float_rn(a1), __uint2float_rn(b1)), f2(TURN_SCALE));
        const float2 rad = __fmul2_rn(k2, f2(sfu_lg2(u.x), sfu_lg2(u.y)));
        const float2 r = f2(sfu_sqrt(rad.x), sfu_sqrt(rad.y));
        const float2 cs = f2(sfu_cos(ph.x), sfu_cos(ph.y)), sn = f2(sfu_sin(ph.x), sfu_sin(ph.y));
        if (NS) {
            par[2 * m] = __ffma2_rn(r, cs, f2(best[2 * m]));
            par[2 * m + 1] = __ffma2_rn(r, sn, f2(best[2 * m + 1]));
        } else {
            float2 z0 = __fmul2_rn(r, cs), z1 = __fmul2_rn(r, sn);
            const float2 v0 = __ffma2_rn(f2(u01_half_open(a0), u01_half_open(b0)), wid, lo);
            const float2 v1 = __ffma2_rn(f2(u01_half_open(a1), u01_half_open(b1)), wid, lo);
            z0 = f2(ua ? v0.x : z0.x, ub ? v0.y : z0.y);
            z1 = f2(ua ? v1.x : z1.x, ub ? v1.y : z1.y);
            const float s0 = sigma_at(2 * m), s1 = sigma_at(2 * m + 1);
            par[2 * m] = __ffma2_rn(z0, f2(ra.mode == 2 ? s0 : fa, rb.mode == 2 ? s0 : fb), f2(best[2 * m]));
            par[2 * m + 1] = __ffma2_rn(z1, f2(ra.mode == 2 ? s1 : fa, rb.mode == 2 ? s1 : fb), f2(best[2 * m + 1]));
        }
    }
}

// the packed generator's eps alone (qpmpc_generate_noise with QPMPC_DUMP=pair: the bit-identity test of the generators)
template <bool NS, class SigmaAt>
__device__ __forceinline__ void noise_block_pair_eps(const RowNoise &ra, const RowNoise &rb, const NoiseKey &nk, int b,
                                                     uint32_t robot, v2 e[NOISE_BLOCK], SigmaAt sigma_at) {
    // best = 0: fma(r, cos, 0) = the rounded product (a -0 product becomes +0: the comparison is by value there)
    float zero[NOISE_BLOCK];
#pragma unroll
    for (int k = 0; k < NOISE_BLOCK; ++k) zero[k] = 0.f;
    noise_block_pair<NS>(ra, rb, nk, b, robot, zero, e, sigma_at);
}

}  // namespace qp
