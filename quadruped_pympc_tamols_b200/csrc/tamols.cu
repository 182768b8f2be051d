// tamols.cu -- TAMOLS candidate scoring + argmin for sm_100a behind the C ABI of include/tamols.h.
//
// One launch scores every cell of every leg job's candidate window:
//   k_tamols_score  grid (tiles, jobs), 256 threads = an 8 x 32 tile of candidates.  The tile plus a 4-cell halo of the
//                   height map is staged into shared memory with one TMA box load (cp.async.bulk.tensor.3d, UTMALDG)
//                   when the map's row pitch allows a tensor map, else by plain coalesced loads; the 14 stencil
//                   lookups of a candidate (its own height, 4 edge-gradient samples, the 3 x 3 roughness patch) hit
//                   shared memory, the 5 leg-collision samples (which lie between hip and candidate, i.e. outside the
//                   tile) go through L2.  Each thread scores one candidate in FP64 in the reference's operation order,
//                   the block reduces (score, index) lexicographically (= first minimum, like the strict `<` of the
//                   reference loop), and the last block of a job (atomic ticket) merges the per-tile partials and
//                   writes the adapted foothold and its constraint box.
// Reference semantics: quadruped_pympc/helpers/visual_foothold_adaptation.py:153-714 (see include/tamols.h).
// Compiled with -fmad=false so that no FMA contraction changes the float64 rounding relative to numpy.
// No CPU fallback: every entry point fails with TAMOLS_E_NODEVICE / TAMOLS_E_CUDA when the GPU path cannot run.
#include <cuda.h>
#include <cuda_runtime.h>

#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <type_traits>
#include <vector>

#include "../../include/tamols.h"

namespace tamols {

// A block of 256 threads scores a TW x TH tile of candidates: TH = 32 (4 cells per thread in the gate pass, 4x fewer
// blocks, partials and tickets) when the launch has enough tiles to fill the GPU, TH = 8 (one cell per thread) for a
// few legs, where latency matters and the extra blocks spread over more SMs.
constexpr int TW = 32, HALO = 4;
constexpr int BW = TW + 2 * HALO;  // 40 cells per staged row; staged rows = TH + 2 * HALO
constexpr int THREADS = 256;
#ifndef TAMOLS_EXACT_LOOKUPS
#define TAMOLS_EXACT_LOOKUPS 0  // 1: every stencil look-up through the per-point float64 nearest-cell computation (A/B)
#endif

struct Partial {
    double score;
    int32_t index;
    int32_t nfinite;
};

// A divisor that is the same for every candidate of a launch, with its IEEE reciprocal (div_exact below).
struct Recip {
    double b, y;  // y = 1.0 / b, correctly rounded (computed by the host's IEEE division)
};

// Constants derived from tamols_params once per launch on the host (derive_constants): the kernel's FP64 diet.
struct Derived {
    Recip two_delta;   // 2.0 * gradient_delta: _compute_edge_cost's central differences (:452-455)
    Recip six_delta2;  // 6.0 * (delta * delta): the plane fit of _compute_roughness_cost (:497-503)
    Recip nine;        // 9.0: mean height, mean residual, variance (:503, :511-516)
    // _check_kinematic_feasibility (:375-395) without the square root: for x >= 0, l_min <= sqrt(x) <= l_max with the
    // correctly rounded sqrt is the same predicate as reach_lo2 <= x <= reach_hi2, where reach_lo2 is the smallest
    // double whose rounded root reaches l_min and reach_hi2 the largest whose rounded root stays within l_max
    // (sqrt is monotonic; the host finds both by stepping ulps around l * l with its own IEEE sqrt)
    double reach_lo2, reach_hi2;
};

struct Args {
    tamols_params p;
    Derived k;
    const void *heights;
    const double *xy;
    const tamols_leg_job *jobs;
    tamols_result *results;
    double *scores;
    long long scores_stride;
    Partial *partials;
    unsigned int *tickets;
    unsigned int *host_flag;   // host-buffer call: mapped host block of {word, flag_value} pairs, 22 per job (else NULL)
    unsigned int flag_value;   // sequence number of this call
    int n_jobs;
    int rows, cols, pitch, max_tiles;
};

// per-block view of one leg job
struct Ctx {
    double ox, oy, rax, ray, cax, cay;  // grid
    double arx, ary, acx, acy;          // axes / |axis|^2: grid coordinate = d . a
    int map, r0, c0, nr, nc, flags;
    int br, bc, bh;                      // map coordinates of the staged box's first cell, staged rows
    long long map_base;                  // map * rows * pitch
    int pitch;
};

// height of map cell (ir, ic), clamped to the leg's window (the KD-tree upstream only holds the leg's own cells, so the
// nearest neighbour of a point outside is the nearest border cell), + the sensor offset of FastHeightMap.get_height
// BOX = false: no staged box (k_tamols_window, the result record), every cell through the read-only path (L1 / L2)
template <typename HT, bool BOX = true>
__device__ __forceinline__ double cell_height(const Args &a, const Ctx &c, const HT *s_box, int ir, int ic) {
    ir = min(max(ir, c.r0), c.r0 + c.nr - 1);
    ic = min(max(ic, c.c0), c.c0 + c.nc - 1);
    const int lr = ir - c.br, lc = ic - c.bc;
    HT h;
    if (BOX && (unsigned)lr < (unsigned)c.bh && (unsigned)lc < (unsigned)BW)
        h = s_box[lr * BW + lc];
    else
        h = __ldg(static_cast<const HT *>(a.heights) + c.map_base + (long long)ir * c.pitch + ic);
    return (double)h + a.p.sensor_offset;
}

template <typename HT, bool BOX = true>
__device__ __forceinline__ double lookup(const Args &a, const Ctx &c, const HT *s_box, double x, double y) {
    // FastHeightMap.get_height (:31-35): nearest cell of the leg's map + 0.02
    const double dx = x - c.ox, dy = y - c.oy;
    const double gr = dx * c.arx + dy * c.ary;
    const double gc = dx * c.acx + dy * c.acy;
    return cell_height<HT, BOX>(a, c, s_box, __double2int_rn(gr), __double2int_rn(gc));
}

// The 3 x 3 patch of _compute_roughness_cost / _compute_edge_cost samples the map at the candidate + (i, k) * delta.
// A candidate IS a cell centre, so in grid coordinates those nine points are the candidate's cell (R, C) plus nine
// offsets that are the same for every candidate of the job: their nearest cells are (R + dr, C + dc) with nine integer
// pairs computed once per block -- no float64 arithmetic per lookup (it was 4 DMUL + 2 DADD + 2 F2I for each of the 14
// stencil look-ups of a candidate, 40 % of the kernel's FP64 instructions).  Exactly the cells the per-point
// computation finds unless an offset sits within 1e-6 cells of a rounding tie (delta = half a cell, or such a yaw):
// the block then keeps the per-point path (`exact`).
struct Stencil {
    int dr[9], dc[9];
    int exact;
    int off[9];      // dr * pitch + dc: the nine cells relative to the candidate's own (k_tamols_window)
    int rmax, cmax;  // max |dr|, max |dc|: a candidate at least that far inside its window needs no clamping
};

__device__ __forceinline__ void candidate_xy(const Args &a, const Ctx &c, int R, int C, double &x, double &y) {
    if (a.xy) {  // coordinates as stored by the sensor (:243)
        const double2 q = __ldg(reinterpret_cast<const double2 *>(a.xy) + ((long long)c.map * a.rows + R) * a.cols + C);
        x = q.x;
        y = q.y;
    } else {
        x = (c.ox + (double)R * c.rax) + (double)C * c.cax;
        y = (c.oy + (double)R * c.ray) + (double)C * c.cay;
    }
}

// candidate construction (:188-196) for map cell (R, C): the cell's own coordinates and its own height (the nearest
// cell of a cell centre is the cell) + the foot clearance
template <typename HT, bool EXACT, bool BOX = true>
__device__ __forceinline__ void candidate_point(const Args &a, const Ctx &c, const HT *s_box, int R, int C, double cand[3]) {
    candidate_xy(a, c, R, C, cand[0], cand[1]);
    cand[2] = (EXACT ? lookup<HT, BOX>(a, c, s_box, cand[0], cand[1]) : cell_height<HT, BOX>(a, c, s_box, R, C)) + a.p.foot_clearance;
}

// a / k.b, bit-identical to the IEEE division, from the precomputed reciprocal: q = RN(a y) is within 2 ulp of a / b;
// one residual step r = a - b q (exact in an FMA), q + r y brings it within half an ulp plus 2^-100, and a second step
// is then Markstein's correction (q faithful, y = RN(1/b)  =>  RN(q + r y) = RN(a / b)).  Five FP64 instructions in
// place of the ~25 of div.rn.f64, and no call: nvcc's division takes an out-of-line slow path whenever the NUMERATOR
// is zero -- which on flat ground (stepping stones, stairs, floors) is every gradient, every plane coefficient and
// every variance, seven calls per candidate, 14 % of the kernel's executed instructions (profiles/r5_ncu_source_tamols).
// Outside 1e-290 < |a| < 1e290 (intermediate underflow / overflow, infinities, NaN) the true division is used; zero
// is returned as a * y, which has the sign a / b has (b > 0).  Explicit FMAs: this file is compiled with -fmad=false.
__device__ __forceinline__ double div_exact(double a, const Recip &k) {
    const double m = fabs(a);
    if (m > 1e-290 && m < 1e290) {
        double q = a * k.y;
        double r = __fma_rn(-k.b, q, a);
        q = __fma_rn(r, k.y, q);
        r = __fma_rn(-k.b, q, a);
        return __fma_rn(r, k.y, q);
    }
    if (m == 0.0) return a * k.y;
    return a / k.b;
}
// sqrt(x) with the zero case answered at once (sqrt.rn.f64 also leaves its fast path for a zero operand)
__device__ __forceinline__ double sqrt_z(double x) { return x == 0.0 ? x : sqrt(x); }

// _check_kinematic_feasibility (:375-395), the cheap first gate of _compute_tamols_score (:287-290): most cells of a
// window fail it (they lie outside the leg's reach)
__device__ __forceinline__ bool candidate_reachable(const Args &a, const Ctx &c, const tamols_leg_job &j, const double cand[3]) {
    const tamols_params &p = a.p;
    double d0 = cand[0] - j.hip[0], d1 = cand[1] - j.hip[1], d2 = cand[2] - j.hip[2];
    const double td2 = (d0 * d0 + d1 * d1) + d2 * d2;  // l_min <= sqrt(td2) <= l_max, see Derived
    if (!(a.k.reach_lo2 <= td2 && td2 <= a.k.reach_hi2)) return false;
    if (c.flags & TAMOLS_FLAG_HAS_VELOCITY) {
        d0 = cand[0] - (j.hip[0] + j.vel[0] * p.stance_duration);
        d1 = cand[1] - (j.hip[1] + j.vel[1] * p.stance_duration);
        d2 = cand[2] - (j.hip[2] + j.vel[2] * p.stance_duration);
        const double lo2 = (d0 * d0 + d1 * d1) + d2 * d2;
        if (!(a.k.reach_lo2 <= lo2 && lo2 <= a.k.reach_hi2)) return false;
    }
    return true;
}

// the rest of _compute_tamols_score (:291-373) for a reachable candidate
template <typename HT, bool EXACT, bool BOX = true>
__device__ double score_reachable(const Args &a, const Ctx &c, const tamols_leg_job &j, const HT *s_box,
                                  const double cand[3], int R, int C, const Stencil &sten, bool moving) {
    const tamols_params &p = a.p;
    const double inf = __longlong_as_double(0x7ff0000000000000LL);
    // _check_leg_collision (:397-420)
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        const double al = p.collision_alphas[k], be = 1.0 - al;
        const double px = be * j.hip[0] + al * cand[0];
        const double py = be * j.hip[1] + al * cand[1];
        const double pz = be * j.hip[2] + al * cand[2];
        const double h_ground = lookup<HT, BOX>(a, c, s_box, px, py) - p.sensor_offset;
        if (pz < h_ground + p.collision_margin) return inf;
    }
    const double d = p.gradient_delta;
    // _compute_roughness_cost (:468-521) and _compute_edge_cost (:422-466) share the 3 x 3 patch at spacing delta
    double h[9];
#pragma unroll
    for (int i = -1; i <= 1; ++i)
#pragma unroll
        for (int k = -1; k <= 1; ++k)
        {
            const int q = (i + 1) * 3 + (k + 1);
            if (EXACT) h[q] = lookup<HT, BOX>(a, c, s_box, cand[0] + (double)i * d, cand[1] + (double)k * d);
            else if (BOX) h[q] = cell_height<HT, true>(a, c, s_box, R + sten.dr[q], C + sten.dc[q]);
        }
    if (!EXACT && !BOX) {
        // no staged box: a candidate whose whole patch lies inside its window (all but a border strip) reads the nine
        // cells at fixed offsets from its own -- no clamps, one address each; the same cells, the same values
        const int ir = R - c.r0, ic = C - c.c0;
        if (ir >= sten.rmax && c.nr - 1 - ir >= sten.rmax && ic >= sten.cmax && c.nc - 1 - ic >= sten.cmax) {
            const HT *own = static_cast<const HT *>(a.heights) + c.map_base + (long long)R * c.pitch + C;
#pragma unroll
            for (int q = 0; q < 9; ++q) h[q] = (double)__ldg(own + sten.off[q]) + p.sensor_offset;
        } else {
#pragma unroll
            for (int q = 0; q < 9; ++q) h[q] = cell_height<HT, false>(a, c, s_box, R + sten.dr[q], C + sten.dc[q]);
        }
    }
    double edge;
    {
        const double gx = div_exact(fabs(h[2 * 3 + 1] - h[0 * 3 + 1]), a.k.two_delta);
        const double gy = div_exact(fabs(h[1 * 3 + 2] - h[1 * 3 + 0]), a.k.two_delta);
        const double g = sqrt_z(gx * gx + gy * gy);
        edge = g <= p.slope_threshold ? 0.0 : g - p.slope_threshold;
    }
    double rough;
    {
        double sxh = 0.0, syh = 0.0, sh = 0.0;
#pragma unroll
        for (int i = -1; i <= 1; ++i)
#pragma unroll
            for (int k = -1; k <= 1; ++k) {
                const double v = h[(i + 1) * 3 + (k + 1)];
                sxh += ((double)i * d) * v;
                syh += ((double)k * d) * v;
                sh += v;
            }
        const double pa = div_exact(sxh, a.k.six_delta2), pb = div_exact(syh, a.k.six_delta2), pc = div_exact(sh, a.k.nine);
        // residuals are evaluated twice (mean, then spread) instead of being kept: nine fewer float64 values live
        // across the loop, the same operations in the same order
        double mean = 0.0;
#pragma unroll
        for (int i = -1; i <= 1; ++i)
#pragma unroll
            for (int k = -1; k <= 1; ++k)
                mean += h[(i + 1) * 3 + (k + 1)] - ((((double)i * d) * pa + ((double)k * d) * pb) + pc);
        mean = div_exact(mean, a.k.nine);
        double var = 0.0;
#pragma unroll
        for (int i = -1; i <= 1; ++i)
#pragma unroll
            for (int k = -1; k <= 1; ++k) {
                const double res = h[(i + 1) * 3 + (k + 1)] - ((((double)i * d) * pa + ((double)k * d) * pb) + pc);
                var += (res - mean) * (res - mean);
            }
        rough = div_exact(var, a.k.nine);
    }
    // deviation from the seed (:344-345)
    const double e0 = cand[0] - j.seed[0], e1 = cand[1] - j.seed[1], e2 = cand[2] - j.seed[2];
    const double dev = (e0 * e0 + e1 * e1) + e2 * e2;
    // _compute_nominal_kinematic_cost (:523-553)
    double nom;
    {
        const double n0 = j.hip[0] - (cand[0] - 0.0), n1 = j.hip[1] - (cand[1] - 0.0);
        const double n2 = j.hip[2] - (cand[2] - (-p.h_des));
        nom = (n0 * n0 + n1 * n1) + n2 * n2;
    }
    // _compute_reference_tracking_cost (:555-609)
    double trk = 0.0;
    if (!(c.flags & TAMOLS_FLAG_HAS_VELOCITY)) {
        trk = e0 < 0.0 ? e0 * e0 : 0.0;
    } else {
        // moving = !(sqrt(vx^2 + vy^2) < 0.01), a property of the job: evaluated once per thread by the caller
        if (moving && ((j.vel[0] > 0.0 && e0 < 0.0) || (j.vel[0] < 0.0 && e0 > 0.0))) trk = e0 * e0;
    }
    // _compute_stability_cost (:611-714)
    double stab = 0.0;
    if (c.flags & TAMOLS_FLAG_STABILITY) {
        const double comx = j.base_xy[0] + j.vel[0] * p.estimated_swing_time;
        const double comy = j.base_xy[1] + j.vel[1] * p.estimated_swing_time;
        const double vx = j.diag_foot[0] - cand[0], vy = j.diag_foot[1] - cand[1];
        const double wx = comx - cand[0], wy = comy - cand[1];
        const double vv = vx * vx + vy * vy;
        double dist;
        if (vv < 1e-8) {
            dist = sqrt(wx * wx + wy * wy);
        } else {
            double t = (wx * vx + wy * vy) / vv;
            t = t < 0.0 ? 0.0 : (t > 1.0 ? 1.0 : t);
            const double ex = comx - (cand[0] + t * vx), ey = comy - (cand[1] + t * vy);
            dist = sqrt(ex * ex + ey * ey);
        }
        if (dist > p.stability_margin) stab = (dist - p.stability_margin) * (dist - p.stability_margin);
    }
    // total (:368-371), kinematic_cost = 0.0 (:335)
    return (((((0.0 + edge * p.weight_edge_avoidance) + rough * p.weight_roughness) + dev * p.weight_deviation) +
             nom * p.weight_nominal_kinematic) + trk * p.weight_reference_tracking) + stab * p.weight_stability;
}

// ---------------------------------------------------------------- TMA / mbarrier helpers (sm_100a PTX) -------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t phase) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(phase) : "memory");
}
__device__ __forceinline__ void tma_load_3d(void *dst, const CUtensorMap *tmap, uint64_t *bar, int x, int y, int z) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
        ::"r"(smem_u32(dst)), "l"(tmap), "r"(smem_u32(bar)), "r"(x), "r"(y), "r"(z) : "memory");
}

__device__ __forceinline__ void argmin_combine(double &s, int &i, double os, int oi) {
    if (os < s || (os == s && oi < i)) {
        s = os;
        i = oi;
    }
}

// The two passes over one tile (gate, then full score of the survivors), compiled once per look-up flavour.
template <typename HT, bool EXACT, int TH>
__device__ __forceinline__ void tile_passes(const Args &a, const Ctx &c, const tamols_leg_job &j, const HT *s_box,
                                            const Stencil &sten, unsigned short *s_list, int *s_nlist, int job, int tile_r,
                                            int tile_c, double &s, int &idx, int &nfin) {
    constexpr int CPT = TW * TH / THREADS;
    const int t = threadIdx.x;
    const double inf = __longlong_as_double(0x7ff0000000000000LL);
    // ---- pass 1: every cell of the tile through the reachability gate; survivors are compacted into s_list so that
    // pass 2 (collision line, edge, roughness, ...; ~10x the work) runs on full warps instead of a ragged annulus.
    // The compaction order is arbitrary; the result is not: the reduction below is a lexicographic (score, index) min.
#pragma unroll
    for (int k = 0; k < CPT; ++k) {
        const int e = k * THREADS + t;
        const int r = tile_r * TH + e / TW, q = tile_c * TW + e % TW;
        bool reach = false;
        if (r < c.nr && q < c.nc) {
            double cand[3];
            candidate_point<HT, EXACT>(a, c, s_box, c.r0 + r, c.c0 + q, cand);
            reach = candidate_reachable(a, c, j, cand);
            if (!reach && a.scores) a.scores[(long long)job * a.scores_stride + r * c.nc + q] = inf;
        }
        const unsigned m = __ballot_sync(0xffffffffu, reach);
        int base = 0;
        if ((t & 31) == 0 && m) base = atomicAdd(s_nlist, __popc(m));
        base = __shfl_sync(0xffffffffu, base, 0);
        if (reach) s_list[base + __popc(m & ((1u << (t & 31)) - 1u))] = (unsigned short)e;
    }
    __syncthreads();
    // ---- pass 2: full score of the reachable candidates
    s = inf;
    idx = 0x7fffffff;
    nfin = 0;
    const int nlist = *s_nlist;
    const bool moving = !(sqrt(j.vel[0] * j.vel[0] + j.vel[1] * j.vel[1]) < 0.01);  // :567-571
    for (int i = t; i < nlist; i += THREADS) {
        const int e = s_list[i];
        const int r = tile_r * TH + e / TW, q = tile_c * TW + e % TW;
        double cand[3];
        candidate_point<HT, EXACT>(a, c, s_box, c.r0 + r, c.c0 + q, cand);
        double sc = score_reachable<HT, EXACT>(a, c, j, s_box, cand, c.r0 + r, c.c0 + q, sten, moving);
        const int ci = r * c.nc + q;
        if (a.scores) a.scores[(long long)job * a.scores_stride + ci] = sc;
        if (sc < inf) {  // NaN is never selected by the reference's `score < best_score`
            ++nfin;
            argmin_combine(s, idx, sc, ci);
        }
    }

}

// The result record of one job from its (score, index, count) -- best candidate, constraint box (:206-228) -- by one thread.
template <typename HT>
__device__ __forceinline__ void write_result(const Args &a, const Ctx &c, int job, double s, int idx, int nfin) {
    const double inf = __longlong_as_double(0x7ff0000000000000LL);
    // host-buffer call: the record is built locally and leaves as {word, sequence} pairs (below)
    tamols_result local;
    tamols_result &out = a.host_flag != nullptr ? local : a.results[job];
    out.num_finite = nfin;
    out.best_score = s;
    double xyz[3] = {0.0, 0.0, 0.0};
    if (s < inf) {
        out.best_index = idx;
        candidate_xy(a, c, c.r0 + idx / c.nc, c.c0 + idx % c.nc, xyz[0], xyz[1]);
        xyz[2] = lookup<HT, false>(a, c, nullptr, xyz[0], xyz[1]) + a.p.foot_clearance;  // no staged box: through L2
    } else {
        out.best_index = -1;
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        out.best_xyz[k] = xyz[k];
        out.vertex1[k] = xyz[k];
        out.vertex2[k] = xyz[k];
    }
    out.vertex1[0] -= a.p.constraint_box_dx;
    out.vertex1[1] -= a.p.constraint_box_dy;
    out.vertex2[0] += a.p.constraint_box_dx;
    out.vertex2[1] += a.p.constraint_box_dy;
    if (a.host_flag != nullptr) {
        // Results go to mapped host memory as 8-byte {word, sequence} stores: every pair carries its own arrival
        // flag, so neither a system-wide fence nor a job counter is needed; the host waits for each pair.
        constexpr int RW = (int)(sizeof(tamols_result) / 4);
        const uint32_t *w = reinterpret_cast<const uint32_t *>(&local);
        uint2 *dst = reinterpret_cast<uint2 *>(a.host_flag) + (size_t)job * RW;
#pragma unroll
        for (int k = 0; k < RW; ++k)
            asm volatile("st.volatile.global.v2.u32 [%0], {%1, %2};" ::"l"(dst + k), "r"(w[k]), "r"(a.flag_value) : "memory");
    }
}

// ---------------------------------------------------------------- the kernel ---------------------------------
struct GridInv {
    double arx, ary, acx, acy;
};

template <typename HT, bool USE_TMA, int TH>
__global__ void __launch_bounds__(THREADS, TH == 32 ? 4 : 2) k_tamols_score(const __grid_constant__ CUtensorMap tmap, const Args a) {
    constexpr int BH = TH + 2 * HALO;
    __shared__ __align__(128) HT s_box[BH * BW];
    __shared__ __align__(8) uint64_t s_bar;
    __shared__ __align__(8) tamols_leg_job s_job;  // the job record, read once from global memory
    __shared__ GridInv s_inv;
    __shared__ Stencil s_sten;
    __shared__ unsigned short s_list[TW * TH];     // tile-local indices of the reachable candidates
    __shared__ int s_nlist;
    __shared__ double s_rs[THREADS / 32];
    __shared__ int s_ri[THREADS / 32], s_rn[THREADS / 32];
    __shared__ unsigned int s_ticket;

    const int job = blockIdx.y, t = threadIdx.x;
    // window shape first (two words): blocks beyond this job's own tile count leave at once
    const int nr_ = a.jobs[job].nr, nc_ = a.jobs[job].nc;
    const int tiles_c = (nc_ + TW - 1) / TW, tiles_r = (nr_ + TH - 1) / TH;
    const int ntiles = tiles_c * tiles_r;
    if ((int)blockIdx.x >= ntiles) return;
    {   // the 136-byte job record -> shared memory, 34 lanes; lane 0 then derives the inverse grid axes once per block
        static_assert(sizeof(tamols_leg_job) % 4 == 0, "job record is copied word-wise");
        const uint32_t *src = reinterpret_cast<const uint32_t *>(a.jobs + job);
        if (t < (int)(sizeof(tamols_leg_job) / 4)) reinterpret_cast<uint32_t *>(&s_job)[t] = __ldg(src + t);
        if (t == 0) s_nlist = 0;
    }
    __syncthreads();
    const tamols_leg_job &j = s_job;
    Ctx c;
    c.map = j.map; c.r0 = j.r0; c.c0 = j.c0; c.nr = j.nr; c.nc = j.nc; c.flags = j.flags;
    const int tile_r = blockIdx.x / tiles_c, tile_c = blockIdx.x % tiles_c;
    c.br = c.r0 + tile_r * TH - HALO;
    c.bh = BH;
    // TMA needs the box to start on a 16-byte boundary of the row (measured on B200: an odd f64 column raises
    // 'illegal instruction'), so the left halo is HALO .. HALO + 16/sizeof(HT) - 1 cells
    c.bc = (c.c0 + tile_c * TW - HALO) & ~(int)(16 / sizeof(HT) - 1);
    c.pitch = a.pitch;
    c.map_base = (long long)c.map * a.rows * a.pitch;

    // ---- stage tile + halo
    if (USE_TMA) {
        if (t == 0) {
            mbar_init(&s_bar, 1);
            mbar_expect_tx(&s_bar, (uint32_t)(BH * BW * sizeof(HT)));
            tma_load_3d(s_box, &tmap, &s_bar, c.bc, c.br, c.map);  // out-of-bounds cells are zero-filled, never read
        }
    } else {
        const HT *src = static_cast<const HT *>(a.heights) + c.map_base;
        for (int e = t; e < BH * BW; e += THREADS) {
            const int rr = c.br + e / BW, cc = c.bc + e % BW;
            s_box[e] = (rr >= 0 && rr < a.rows && cc >= 0 && cc < a.cols) ? __ldg(src + (long long)rr * a.pitch + cc) : HT(0);
        }
    }
    // ---- per-job constants while the copy is in flight
    c.ox = j.origin[0]; c.oy = j.origin[1];
    c.rax = j.row_axis[0]; c.ray = j.row_axis[1]; c.cax = j.col_axis[0]; c.cay = j.col_axis[1];
    if (t == 32) {
        const double r2 = c.rax * c.rax + c.ray * c.ray, c2 = c.cax * c.cax + c.cay * c.cay;
        s_inv.arx = c.rax / r2; s_inv.ary = c.ray / r2; s_inv.acx = c.cax / c2; s_inv.acy = c.cay / c2;
    }
    if (t >= 64 && t < 73) {  // nearest-cell offsets of the nine patch points, once per block (Stencil)
        const int q = t - 64, i = q / 3 - 1, k = q % 3 - 1;
        const double r2 = c.rax * c.rax + c.ray * c.ray, c2 = c.cax * c.cax + c.cay * c.cay;
        const double dx = (double)i * a.p.gradient_delta, dy = (double)k * a.p.gradient_delta;
        const double gr = dx * (c.rax / r2) + dy * (c.ray / r2), gc = dx * (c.cax / c2) + dy * (c.cay / c2);
        const int ir = __double2int_rn(gr), ic = __double2int_rn(gc);
        s_sten.dr[q] = ir;
        s_sten.dc[q] = ic;
        s_sten.off[q] = ir * a.pitch + ic;
        const bool tie = fabs(fabs(gr - (double)ir) - 0.5) < 1e-6 || fabs(fabs(gc - (double)ic) - 0.5) < 1e-6;
        const unsigned any = __ballot_sync(0x1ffu, tie);
        const int rmax = __reduce_max_sync(0x1ffu, abs(ir)), cmax = __reduce_max_sync(0x1ffu, abs(ic));
        if (q == 0) {
            s_sten.exact = (any != 0u || TAMOLS_EXACT_LOOKUPS) ? 1 : 0;
            s_sten.rmax = rmax;
            s_sten.cmax = cmax;
        }
    }
    __syncthreads();  // s_bar initialised (TMA) / s_box written (plain) / s_inv, s_sten written
    c.arx = s_inv.arx; c.ary = s_inv.ary; c.acx = s_inv.acx; c.acy = s_inv.acy;
    const bool exact = s_sten.exact != 0;
    if (USE_TMA) mbar_wait(&s_bar, 0);

    const double inf = __longlong_as_double(0x7ff0000000000000LL);
    double s;
    int idx, nfin;
    if (exact) tile_passes<HT, true, TH>(a, c, j, s_box, s_sten, s_list, &s_nlist, job, tile_r, tile_c, s, idx, nfin);
    else tile_passes<HT, false, TH>(a, c, j, s_box, s_sten, s_list, &s_nlist, job, tile_r, tile_c, s, idx, nfin);

    // ---- block reduction: lexicographic (score, index) minimum = first minimum in row-major order
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const double os = __shfl_xor_sync(0xffffffffu, s, o);
        const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
        argmin_combine(s, idx, os, oi);
        nfin += __shfl_xor_sync(0xffffffffu, nfin, o);
    }
    if ((t & 31) == 0) {
        s_rs[t >> 5] = s;
        s_ri[t >> 5] = idx;
        s_rn[t >> 5] = nfin;
    }
    __syncthreads();
    if (t == 0) {
        for (int w = 1; w < THREADS / 32; ++w) {
            argmin_combine(s, idx, s_rs[w], s_ri[w]);
            nfin += s_rn[w];
        }
        if (ntiles > 1) {
            Partial &pt = a.partials[(long long)job * a.max_tiles + blockIdx.x];
            pt.score = s;
            pt.index = idx;
            pt.nfinite = nfin;
            __threadfence();
            s_ticket = atomicAdd(&a.tickets[job], 1u);
        }
    }
    // A job that fits ONE tile (one robot's 13 x 7 sensor maps) is finished here: thread 0 already holds its
    // (score, index, count) -- no partial, no fence, no ticket, no merge (three dependent memory round trips less).
    if (ntiles > 1) {
        __syncthreads();
        if (s_ticket != (unsigned)(ntiles - 1)) return;

        // ---- last block of this job: merge the tile partials (:209-222)
        __threadfence();
        s = inf;
        idx = 0x7fffffff;
        nfin = 0;
        for (int b = t; b < ntiles; b += THREADS) {
            // written by other blocks: read through L2 (ld.global.cg), not this SM's L1
            const int4 raw = __ldcg(reinterpret_cast<const int4 *>(a.partials + (long long)job * a.max_tiles + b));
            argmin_combine(s, idx, __hiloint2double(raw.y, raw.x), raw.z);
            nfin += raw.w;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const double os = __shfl_xor_sync(0xffffffffu, s, o);
            const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
            argmin_combine(s, idx, os, oi);
            nfin += __shfl_xor_sync(0xffffffffu, nfin, o);
        }
        __syncthreads();
        if ((t & 31) == 0) {
            s_rs[t >> 5] = s;
            s_ri[t >> 5] = idx;
            s_rn[t >> 5] = nfin;
        }
        __syncthreads();
        if (t == 0) {
            for (int w = 1; w < THREADS / 32; ++w) {
                argmin_combine(s, idx, s_rs[w], s_ri[w]);
                nfin += s_rn[w];
            }
        }
    }
    if (t == 0) {
        write_result<HT>(a, c, job, s, idx, nfin);
        if (ntiles > 1) a.tickets[job] = 0;  // ready for the next launch on this stream
    }
}

// ---------------------------------------------------------------- the throughput kernel ----------------------
// One block of WTHREADS threads per JOB, the whole candidate window: the batched regime (thousands of jobs in one
// launch), where a job does not need to be spread over SMs.  Same per-candidate code as k_tamols_score (gate,
// score_reachable), same lexicographic (score, index) minimum, hence the same result bits; what changes is the shape
// of the work:
//   * no staged box -- every height comes through the read-only path (pass 1 reads the window row-major, coalesced;
//     the stencil and collision look-ups of pass 2 hit L1 / L2: the map's window region is a few tens of KB);
//   * ONE compacted list per band of at most WBAND cells (the whole window when it has <= WBAND cells): with ~830
//     reachable candidates of a 64 x 64 window and 128 threads a thread scores 6.5 candidates, so the ragged last round
//     costs 1 / 7 of pass 2 instead of the 20 % (205 candidates on 256 threads) .. 50 % (257+) of a 32 x 32 tile;
//   * no per-tile partials, arrival tickets or merge, a third of the block barriers per candidate.
// 64 registers x 128 threads: up to 8 blocks per SM (the tile kernel: 4 x 256 threads with one to two idle warps each).
constexpr int WBAND = 8192;  // cells per band: the list holds 16-bit cell offsets inside the band (16 KB of shared memory)

// candidate_reachable with the job's constants in registers: hip, hip + vel * stance_duration (the same product and
// sum, rounded once per job instead of once per cell)
struct Gate {
    double h0, h1, h2, v0, v1, v2;
    bool has_vel;
};
// window-relative rows / columns outside of which no cell can pass the gate (k_tamols_window's setup)
struct WinBox {
    int r0, r1, c0, c1;
};
__device__ __forceinline__ bool gate_reachable(const Args &a, const Gate &g, const double cand[3]) {
    double d0 = cand[0] - g.h0, d1 = cand[1] - g.h1, d2 = cand[2] - g.h2;
    const double td2 = (d0 * d0 + d1 * d1) + d2 * d2;
    if (!(a.k.reach_lo2 <= td2 && td2 <= a.k.reach_hi2)) return false;
    if (g.has_vel) {
        d0 = cand[0] - g.v0;
        d1 = cand[1] - g.v1;
        d2 = cand[2] - g.v2;
        const double lo2 = (d0 * d0 + d1 * d1) + d2 * d2;
        if (!(a.k.reach_lo2 <= lo2 && lo2 <= a.k.reach_hi2)) return false;
    }
    return true;
}

template <typename HT, bool EXACT, int WTHREADS>
__device__ __forceinline__ void window_passes(const Args &a, const Ctx &c, const tamols_leg_job &j, const Stencil &sten,
                                              const WinBox &box, unsigned short *s_list, int *s_nlist, int job, double &s,
                                              int &idx, int &nfin) {
    constexpr int NW = WTHREADS / 32;
    const int t = threadIdx.x, lane = t & 31, wid = t >> 5;
    const double inf = __longlong_as_double(0x7ff0000000000000LL);
    const bool moving = !(sqrt(j.vel[0] * j.vel[0] + j.vel[1] * j.vel[1]) < 0.01);  // :567-571
    s = inf;
    idx = 0x7fffffff;
    nfin = 0;
    // bands of whole rows, at most WBAND cells each (the host sends windows wider than WBAND to the tile kernel)
    const int band_rows = max(1, WBAND / c.nc);
    for (int rb = 0; rb < c.nr; rb += band_rows) {
        const int nrb = min(band_rows, c.nr - rb);
        Gate g;  // built per band: twelve registers that pass 2 does not have to carry
        g.h0 = j.hip[0]; g.h1 = j.hip[1]; g.h2 = j.hip[2];
        g.v0 = j.hip[0] + j.vel[0] * a.p.stance_duration;
        g.v1 = j.hip[1] + j.vel[1] * a.p.stance_duration;
        g.v2 = j.hip[2] + j.vel[2] * a.p.stance_duration;
        g.has_vel = (c.flags & TAMOLS_FLAG_HAS_VELOCITY) != 0;
        const HT *hmap = static_cast<const HT *>(a.heights) + c.map_base;
        // ---- pass 1: the reachability gate on every cell of the band, a warp per row (coalesced, no index division),
        // survivors compacted as 16-bit offsets inside the band (order arbitrary, result not)
        // (two copies of the loop: the plain one -- grid coordinates, no per-candidate score output, what a batched
        // launch runs -- keeps the two pointer tests and their address arithmetic out of the 30-odd instructions a cell costs)
        // plain loop only: rows and 32-cell segments outside the reach circles' bounding box are not visited
        const int box_r0 = box.r0, box_r1 = box.r1, box_c0 = box.c0, box_c1 = box.c1;  // computed once per job (k_tamols_window)
        const int seg0 = box_c0 & ~31;  // first 32-cell segment that touches the box
        auto pass1 = [&](auto plain_tag) {
            constexpr bool PLAIN = decltype(plain_tag)::value;
            const int r_first = PLAIN ? max(rb, box_r0) : rb, r_last = PLAIN ? min(rb + nrb - 1, box_r1) : rb + nrb - 1;
            const int q_first = PLAIN ? seg0 : 0, q_end = PLAIN ? min(c.nc, box_c1 + 1) : c.nc;
            // (row r of the band goes to warp (r - rb) % NW whatever the box: the compaction order stays arbitrary)
            for (int r = r_first + ((wid - (r_first - rb)) % NW + NW) % NW; r <= r_last; r += NW) {
                const int R = c.r0 + r;
                const double xr = c.ox + (double)R * c.rax, yr = c.oy + (double)R * c.ray;  // candidate_xy's first sum
                const HT *hrow = hmap + (long long)R * c.pitch + c.c0;
                for (int q0 = q_first; q0 < q_end; q0 += 32) {
                    const int q = q0 + lane;
                    bool reach = false;
                    if (q < c.nc) {
                        double cand[3];
                        if (!PLAIN && a.xy) candidate_xy(a, c, R, c.c0 + q, cand[0], cand[1]);
                        else {
                            cand[0] = xr + (double)(c.c0 + q) * c.cax;
                            cand[1] = yr + (double)(c.c0 + q) * c.cay;
                        }
                        // the cell's own height (the nearest cell of a cell centre is the cell) + offset + clearance
                        cand[2] = (EXACT ? lookup<HT, false>(a, c, nullptr, cand[0], cand[1])
                                         : (double)__ldg(hrow + q) + a.p.sensor_offset) + a.p.foot_clearance;
                        reach = gate_reachable(a, g, cand);
                        if (!PLAIN && !reach && a.scores) a.scores[(long long)job * a.scores_stride + r * c.nc + q] = inf;
                    }
                    const unsigned m = __ballot_sync(0xffffffffu, reach);
                    if (m) {  // warp-uniform
                        // one shared-memory atomic per warp and trip, issued as PTX: atomicAdd() under `lane == 0` makes
                        // nvcc wrap its own warp aggregation (vote, find-leader, shuffle: ~25 instructions) around it
                        unsigned base = 0;
                        if (lane == 0)
                            asm volatile("atom.shared.add.u32 %0, [%1], %2;" : "=r"(base) : "r"(smem_u32(s_nlist)), "r"(__popc(m)) : "memory");
                        base = __shfl_sync(0xffffffffu, base, 0);
                        if (reach) s_list[base + __popc(m & ((1u << lane) - 1u))] = (unsigned short)((r - rb) * c.nc + q);
                    }
                }
            }
        };
        if (a.xy == nullptr && a.scores == nullptr) pass1(std::true_type{});
        else pass1(std::false_type{});
        __syncthreads();
        // ---- pass 2: full score of the reachable candidates
        const int nlist = *s_nlist;
        for (int i = t; i < nlist; i += WTHREADS) {
            const int e = s_list[i], rl = e / c.nc, q = e - rl * c.nc, r = rb + rl, ci = r * c.nc + q;
            double cand[3];
            candidate_point<HT, EXACT, false>(a, c, nullptr, c.r0 + r, c.c0 + q, cand);
            const double sc = score_reachable<HT, EXACT, false>(a, c, j, nullptr, cand, c.r0 + r, c.c0 + q, sten, moving);
            if (a.scores) a.scores[(long long)job * a.scores_stride + ci] = sc;
            if (sc < inf) {  // NaN is never selected by the reference's `score < best_score`
                ++nfin;
                argmin_combine(s, idx, sc, ci);
            }
        }
        __syncthreads();  // every thread is done with this band's list
        if (t == 0) *s_nlist = 0;
        __syncthreads();
    }
}

template <typename HT, int WTHREADS>
__global__ void __launch_bounds__(WTHREADS, 1024 / WTHREADS) k_tamols_window(const Args a) {
    __shared__ __align__(8) tamols_leg_job s_job;
    __shared__ GridInv s_inv;
    __shared__ Stencil s_sten;
    __shared__ WinBox s_wbox;
    __shared__ unsigned short s_list[WBAND];
    __shared__ int s_nlist;
    __shared__ double s_rs[WTHREADS / 32];
    __shared__ int s_ri[WTHREADS / 32], s_rn[WTHREADS / 32];
    const int job = blockIdx.x, t = threadIdx.x;
    {
        const uint32_t *src = reinterpret_cast<const uint32_t *>(a.jobs + job);
        if (t < (int)(sizeof(tamols_leg_job) / 4)) reinterpret_cast<uint32_t *>(&s_job)[t] = __ldg(src + t);
        if (t == 0) s_nlist = 0;
    }
    __syncthreads();
    const tamols_leg_job &j = s_job;
    Ctx c;
    c.map = j.map; c.r0 = j.r0; c.c0 = j.c0; c.nr = j.nr; c.nc = j.nc; c.flags = j.flags;
    c.br = 0; c.bc = 0; c.bh = 0;  // no staged box
    c.pitch = a.pitch;
    c.map_base = (long long)c.map * a.rows * a.pitch;
    c.ox = j.origin[0]; c.oy = j.origin[1];
    c.rax = j.row_axis[0]; c.ray = j.row_axis[1]; c.cax = j.col_axis[0]; c.cay = j.col_axis[1];
    if (t == 0) {
        const double r2 = c.rax * c.rax + c.ray * c.ray, c2 = c.cax * c.cax + c.cay * c.cay;
        const double arx = c.rax / r2, ary = c.ray / r2, acx = c.cax / c2, acy = c.cay / c2;
        s_inv.arx = arx; s_inv.ary = ary; s_inv.acx = acx; s_inv.acy = acy;
        // The cells that can pass the reachability gate lie in the bounding box, in grid coordinates, of the reach
        // circle(s) -- horizontal distance to the hip (and to hip + vel * stance) at most sqrt(reach_hi2) -- widened by
        // one cell for the round-off of this estimate: pass 1 of a batched launch does not visit rows and 32-cell
        // segments outside it (they fail the exact gate, whose verdict is all pass 1 records).  Orthogonal axes
        // (include/tamols.h), so the circle's extent along an axis is radius / |axis|.
        WinBox b = {0, c.nr - 1, 0, c.nc - 1};
        const double rad = sqrt(a.k.reach_hi2);
        const double er = rad * sqrt(arx * arx + ary * ary) + 1.0, ec = rad * sqrt(acx * acx + acy * acy) + 1.0;
        auto clampd = [](double v) { return fmin(fmax(v, -1.0e9), 1.0e9); };
        auto shrink = [&](double px, double py) {
            const double gr = (px - c.ox) * arx + (py - c.oy) * ary, gc = (px - c.ox) * acx + (py - c.oy) * acy;
            b.r0 = max(b.r0, (int)clampd(floor(gr - er)) - c.r0);
            b.r1 = min(b.r1, (int)clampd(ceil(gr + er)) - c.r0);
            b.c0 = max(b.c0, (int)clampd(floor(gc - ec)) - c.c0);
            b.c1 = min(b.c1, (int)clampd(ceil(gc + ec)) - c.c0);
        };
        shrink(j.hip[0], j.hip[1]);
        if (c.flags & TAMOLS_FLAG_HAS_VELOCITY)
            shrink(j.hip[0] + j.vel[0] * a.p.stance_duration, j.hip[1] + j.vel[1] * a.p.stance_duration);
        s_wbox = b;
    }
    if (t >= 32 && t < 41) {  // nearest-cell offsets of the nine patch points (Stencil), as in k_tamols_score
        const int q = t - 32, i = q / 3 - 1, k = q % 3 - 1;
        const double r2 = c.rax * c.rax + c.ray * c.ray, c2 = c.cax * c.cax + c.cay * c.cay;
        const double dx = (double)i * a.p.gradient_delta, dy = (double)k * a.p.gradient_delta;
        const double gr = dx * (c.rax / r2) + dy * (c.ray / r2), gc = dx * (c.cax / c2) + dy * (c.cay / c2);
        const int ir = __double2int_rn(gr), ic = __double2int_rn(gc);
        s_sten.dr[q] = ir;
        s_sten.dc[q] = ic;
        s_sten.off[q] = ir * a.pitch + ic;
        const bool tie = fabs(fabs(gr - (double)ir) - 0.5) < 1e-6 || fabs(fabs(gc - (double)ic) - 0.5) < 1e-6;
        const unsigned any = __ballot_sync(0x1ffu, tie);
        const int rmax = __reduce_max_sync(0x1ffu, abs(ir)), cmax = __reduce_max_sync(0x1ffu, abs(ic));
        if (q == 0) {
            s_sten.exact = (any != 0u || TAMOLS_EXACT_LOOKUPS) ? 1 : 0;
            s_sten.rmax = rmax;
            s_sten.cmax = cmax;
        }
    }
    __syncthreads();
    c.arx = s_inv.arx; c.ary = s_inv.ary; c.acx = s_inv.acx; c.acy = s_inv.acy;
    double s;
    int idx, nfin;
    if (s_sten.exact != 0) window_passes<HT, true, WTHREADS>(a, c, j, s_sten, s_wbox, s_list, &s_nlist, job, s, idx, nfin);
    else window_passes<HT, false, WTHREADS>(a, c, j, s_sten, s_wbox, s_list, &s_nlist, job, s, idx, nfin);
    // ---- block reduction: lexicographic (score, index) minimum = first minimum in row-major order
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const double os = __shfl_xor_sync(0xffffffffu, s, o);
        const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
        argmin_combine(s, idx, os, oi);
        nfin += __shfl_xor_sync(0xffffffffu, nfin, o);
    }
    if ((t & 31) == 0) {
        s_rs[t >> 5] = s;
        s_ri[t >> 5] = idx;
        s_rn[t >> 5] = nfin;
    }
    __syncthreads();
    if (t == 0) {
        for (int w = 1; w < WTHREADS / 32; ++w) {
            argmin_combine(s, idx, s_rs[w], s_ri[w]);
            nfin += s_rn[w];
        }
        write_result<HT>(a, c, job, s, idx, nfin);
    }
}

// ---------------------------------------------------------------- seed footholds + job records
// One thread per robot.  foothold_reference_generator.py:84-157 in its operation order (float64, no contraction).
__global__ void k_seed_jobs(const __grid_constant__ tamols_seed_batch b) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= b.num_robots) return;
    const double *bp = b.d_base_position + 3 * r, *eu = b.d_base_euler + 3 * r;
    const double *v = b.d_base_lin_vel + 3 * r, *vr = b.d_ref_lin_vel + 3 * r;
    const double cy = cos(eu[2]), sy = sin(eu[2]);
    // R_W2H = [[cy, sy], [-sy, cy]]
    const double vhx = cy * v[0] + sy * v[1], vhy = -sy * v[0] + cy * v[1];
    const double rhx = cy * vr[0] + sy * vr[1], rhy = -sy * vr[0] + cy * vr[1];
    // deque(maxlen = hist_len).append, then the mean in deque order (oldest first)
    const int L = b.hist_len;
    double *hist = b.d_vel_hist + (size_t)r * L * 2;
    const int cnt = b.d_hist_count[r];
    hist[2 * (cnt % L)] = vhx;
    hist[2 * (cnt % L) + 1] = vhy;
    b.d_hist_count[r] = cnt + 1;
    const int n = min(cnt + 1, L), first = (cnt + 1 > L) ? (cnt + 1) % L : 0;
    double mx = 0.0, my = 0.0;
    for (int k = 0; k < n; ++k) {
        const int s = (first + k) % L;
        mx += hist[2 * s];
        my += hist[2 * s + 1];
    }
    mx /= (double)n;
    my /= (double)n;
    const double half = b.d_stance_time[r] / 2.0, lim = b.hip_height * 1.5;
    const double dx = fmin(fmax(half * rhx, -lim), lim), dy = fmin(fmax(half * rhy, -lim), lim);
    const double k = sqrt(b.d_com_height[r] / b.gravity);
    double ex = k * (mx - rhx), ey = k * (my - rhy);
    ex = ex > 0.05 ? 0.05 : ex;
    ex = ex < -0.05 ? -0.05 : ex;
    ey = ey > 0.05 ? 0.05 : ey;
    ey = ey < -0.05 ? -0.05 : ey;
    const double ox = dx + ex, oy = dy + ey;  // vel_offset + error_compensation
    // manual CoM offset: Rotation.from_euler("xyz", euler).as_matrix() @ com_pos_offset_b, x and y rows
    double cwx = 0.0, cwy = 0.0;
    if (b.d_com_offset_b != nullptr) {
        const double *co = b.d_com_offset_b + 3 * r;
        const double cr = cos(eu[0]), sr = sin(eu[0]), cp = cos(eu[1]), sp = sin(eu[1]);
        cwx = cp * cy * co[0] + (sr * sp * cy - cr * sy) * co[1] + (cr * sp * cy + sr * sy) * co[2];
        cwy = cp * sy * co[0] + (sr * sp * sy + cr * cy) * co[1] + (cr * sp * sy - sr * cy) * co[2];
    }
    for (int l = 0; l < 4; ++l) {
        const double *hp = b.d_hips + (size_t)(4 * r + l) * 3;
        const double px = hp[0] - bp[0], py = hp[1] - bp[1];
        double fx = cy * px + sy * py, fy = -sy * px + cy * py;
        fy += (l & 1) ? -b.hip_offset : b.hip_offset;  // FL, RL: +, FR, RR: -
        fx += ox;
        fy += oy;
        double wx = (cy * fx - sy * fy) + bp[0], wy = (sy * fx + cy * fy) + bp[1];  // R_W2H.T @ f + base
        wx += cwx;
        wy += cwy;
        const double wz = b.d_lift_off_z[4 * r + l];
        double *out = b.d_ref_feet + (size_t)(4 * r + l) * 3;
        out[0] = wx;
        out[1] = wy;
        out[2] = wz;
        if (b.d_jobs == nullptr) continue;
        tamols_leg_job j;
        j.map = 0;
        // nearest cell of the seed in grid coordinates, window centred on it and clamped to the map
        const double qx = wx - b.map_origin[0], qy = wy - b.map_origin[1];
        const double r2 = b.map_row_axis[0] * b.map_row_axis[0] + b.map_row_axis[1] * b.map_row_axis[1];
        const double c2 = b.map_col_axis[0] * b.map_col_axis[0] + b.map_col_axis[1] * b.map_col_axis[1];
        const int ri = (int)rint((qx * b.map_row_axis[0] + qy * b.map_row_axis[1]) / r2);
        const int ci = (int)rint((qx * b.map_col_axis[0] + qy * b.map_col_axis[1]) / c2);
        j.nr = min(b.win_rows, b.map_rows);
        j.nc = min(b.win_cols, b.map_cols);
        j.r0 = min(max(ri - j.nr / 2, 0), b.map_rows - j.nr);
        j.c0 = min(max(ci - j.nc / 2, 0), b.map_cols - j.nc);
        j.origin[0] = b.map_origin[0]; j.origin[1] = b.map_origin[1];
        j.row_axis[0] = b.map_row_axis[0]; j.row_axis[1] = b.map_row_axis[1];
        j.col_axis[0] = b.map_col_axis[0]; j.col_axis[1] = b.map_col_axis[1];
        j.seed[0] = wx; j.seed[1] = wy; j.seed[2] = wz;
        j.hip[0] = hp[0]; j.hip[1] = hp[1]; j.hip[2] = hp[2];
        j.vel[0] = v[0]; j.vel[1] = v[1]; j.vel[2] = v[2];
        int flags = TAMOLS_FLAG_HAS_VELOCITY;
        j.base_xy[0] = j.base_xy[1] = j.diag_foot[0] = j.diag_foot[1] = 0.0;
        const bool stance = b.d_contact != nullptr && b.d_contact[4 * r + l] == 1;
        if (b.d_feet != nullptr && !stance) {  // visual_foothold_adaptation.py:631-664: swing leg, diagonal partner
            const int d = 3 - l;                // FL <-> RR, FR <-> RL
            const double *df = b.d_feet + (size_t)(4 * r + d) * 3;
            flags |= TAMOLS_FLAG_STABILITY;
            j.base_xy[0] = bp[0]; j.base_xy[1] = bp[1];
            j.diag_foot[0] = df[0]; j.diag_foot[1] = df[1];
        }
        j.flags = flags;
        b.d_jobs[4 * r + l] = j;
    }
}

}  // namespace tamols

// ================================================================ host side ================================
using namespace tamols;

static thread_local std::string g_terr;
static int tfail(int code, const std::string &msg) {
    g_terr = msg;
    return code;
}
#define TCU(call)                                                                                   \
    do {                                                                                            \
        cudaError_t e_ = (call);                                                                    \
        if (e_ != cudaSuccess)                                                                      \
            return tfail(TAMOLS_E_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_));        \
    } while (0)

// Derived (above) from the handle's parameters.  Host IEEE arithmetic = the device's div.rn / sqrt.rn.
static double sqrt_threshold(double l, bool lower) {
    // lower: smallest x >= 0 with sqrt(x) >= l; else: largest x with sqrt(x) <= l  (sqrt correctly rounded, monotonic)
    if (!(l > 0.0) || !std::isfinite(l)) {  // degenerate limits keep the comparison's meaning
        if (l != l) return l;               // NaN: every comparison false, as upstream
        if (l <= 0.0) return lower ? 0.0 : (l == 0.0 ? 0.0 : -1.0);  // sqrt(x) >= l always; sqrt(x) <= l only for x = 0 (l = 0) / never
        return l;                           // +inf
    }
    double x = l * l;
    if (lower) {
        while (std::sqrt(x) >= l) x = std::nextafter(x, 0.0);
        while (std::sqrt(x) < l) x = std::nextafter(x, INFINITY);
    } else {
        while (std::sqrt(x) <= l) x = std::nextafter(x, INFINITY);
        while (std::sqrt(x) > l) x = std::nextafter(x, 0.0);
    }
    return x;
}
static Derived derive_constants(const tamols_params &p) {
    Derived k;
    const double d = p.gradient_delta;
    k.two_delta.b = 2.0 * d;
    k.six_delta2.b = 6.0 * (d * d);
    k.nine.b = 9.0;
    k.two_delta.y = 1.0 / k.two_delta.b;
    k.six_delta2.y = 1.0 / k.six_delta2.b;
    k.nine.y = 1.0 / k.nine.b;
    k.reach_lo2 = sqrt_threshold(p.l_min, true);
    k.reach_hi2 = sqrt_threshold(p.l_max, false);
    return k;
}

typedef CUresult (*encode_tiled_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                    const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

struct tamols_handle {
    tamols_params p;
    Derived derived;  // derive_constants(p), refreshed with p
    int device, max_jobs, num_sms;
    cudaStream_t stream;
    encode_tiled_fn encode;
    int64_t launches;
    unsigned int *d_tickets;   // [max_jobs] tile arrival tickets + [1] finished-job counter
    Partial *d_partials;
    size_t partial_cap;  // in Partial elements
    // cached tensor map
    CUtensorMap tmap;
    const void *tm_ptr;
    int tm_dtype, tm_maps, tm_rows, tm_cols, tm_pitch, tm_th;
    // staging for tamols_score_host
    char *d_stage, *h_stage;
    size_t stage_cap;
    char *m_out, *md_out;      // mapped pinned block of tamols_score_host: {word, sequence} pairs of the results
    std::vector<double> sensor_heights, sensor_xy;  // host scratch of tamols_score_sensors
    std::vector<tamols_leg_job> sensor_jobs;
    unsigned int seq;
    int env_tile_rows;  // TAMOLS_TILE_ROWS (0 = automatic) and TAMOLS_ZEROCOPY, read from the environment once per
    int env_zerocopy;   // handle: a getenv per call is measurable on a 30 us path
    const char *last_kernel;  // "window" | "tile32" | "tile8": what the last scoring launch ran (tamols_last_kernel)
    int env_kernel;     // TAMOLS_KERNEL: 0 automatic, 1 "tile" (k_tamols_score), 2 "window" (k_tamols_window)
};

extern "C" const char *tamols_last_error(void) { return g_terr.c_str(); }
extern "C" void *tamols_stream(tamols_handle *h) { return h ? (void *)h->stream : nullptr; }
extern "C" int64_t tamols_launch_count(const tamols_handle *h) { return h ? h->launches : 0; }
extern "C" const char *tamols_last_kernel(const tamols_handle *h) { return h && h->last_kernel ? h->last_kernel : ""; }

static int check_params(const tamols_params *p) {
    if (!p) return tfail(TAMOLS_E_ARG, "null params");
    if (!(p->gradient_delta > 0.0)) return tfail(TAMOLS_E_ARG, "gradient_delta must be > 0");
    return TAMOLS_OK;
}

extern "C" int tamols_create(const tamols_params *params, int32_t device, int32_t max_jobs, tamols_handle **out) {
    if (!out) return tfail(TAMOLS_E_ARG, "null argument");
    *out = nullptr;
    if (int rc = check_params(params)) return rc;
    if (max_jobs < 1) return tfail(TAMOLS_E_ARG, "max_jobs must be >= 1");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return tfail(TAMOLS_E_NODEVICE, "no CUDA device: this library has no CPU fallback");
    }
    if (device < 0 || device >= ndev) return tfail(TAMOLS_E_ARG, "device ordinal out of range");
    cudaDeviceProp prop;
    TCU(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10)
        return tfail(TAMOLS_E_NODEVICE, std::string("device is ") + prop.name + "; kernels are built for sm_100a only");
    TCU(cudaSetDevice(device));
    tamols_handle *h = new (std::nothrow) tamols_handle();
    if (!h) return tfail(TAMOLS_E_ARG, "out of host memory");
    memset(h, 0, sizeof *h);
    h->p = *params;
    h->derived = derive_constants(h->p);
    h->device = device;
    h->max_jobs = max_jobs;
    h->num_sms = prop.multiProcessorCount;
    void *fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
        h->encode = (encode_tiled_fn)fn;
    cudaGetLastError();
    bool ok = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking) == cudaSuccess &&
              cudaMalloc(&h->d_tickets, ((size_t)max_jobs + 1) * sizeof(unsigned int)) == cudaSuccess &&
              cudaMemset(h->d_tickets, 0, ((size_t)max_jobs + 1) * sizeof(unsigned int)) == cudaSuccess &&
              cudaHostAlloc(&h->m_out, 2 * (size_t)max_jobs * sizeof(tamols_result) + 64, cudaHostAllocMapped) == cudaSuccess &&
              cudaHostGetDevicePointer(&h->md_out, h->m_out, 0) == cudaSuccess;
    if (!ok) {
        std::string m = std::string("allocation failed: ") + cudaGetErrorString(cudaGetLastError());
        tamols_destroy(h);
        return tfail(TAMOLS_E_CUDA, m);
    }
    memset(h->m_out, 0, 2 * (size_t)max_jobs * sizeof(tamols_result) + 64);  // sequence 0 = nothing published yet
    h->seq = 0;
    h->env_tile_rows = 0;
    if (const char *e = getenv("TAMOLS_TILE_ROWS")) h->env_tile_rows = atoi(e) == 32 ? 32 : 8;
    if (const char *e = getenv("TAMOLS_KERNEL")) h->env_kernel = !strcmp(e, "tile") ? 1 : !strcmp(e, "window") ? 2 : 0;
    {
        const char *zc = getenv("TAMOLS_ZEROCOPY");
        h->env_zerocopy = !(zc && zc[0] == '0');
    }
    *out = h;
    return TAMOLS_OK;
}

extern "C" int tamols_destroy(tamols_handle *h) {
    if (!h) return TAMOLS_OK;
    cudaSetDevice(h->device);
    if (h->stream) {
        cudaStreamSynchronize(h->stream);
        cudaStreamDestroy(h->stream);
    }
    if (h->d_tickets) cudaFree(h->d_tickets);
    if (h->d_partials) cudaFree(h->d_partials);
    if (h->d_stage) cudaFree(h->d_stage);
    if (h->h_stage) cudaFreeHost(h->h_stage);
    if (h->m_out) cudaFreeHost(h->m_out);
    delete h;
    return TAMOLS_OK;
}

extern "C" int tamols_set_params(tamols_handle *h, const tamols_params *params) {
    if (!h) return tfail(TAMOLS_E_ARG, "null argument");
    if (int rc = check_params(params)) return rc;
    h->p = *params;
    h->derived = derive_constants(h->p);
    return TAMOLS_OK;
}

// builds (or reuses) the 3-D tensor map [maps][rows][cols] with a BW x (TH + 2 HALO) x 1 box; returns false when TMA cannot
// describe this buffer (pitch or base not 16-byte aligned, driver entry point missing) -> plain-load kernel
static bool get_tensor_map(tamols_handle *h, const void *d_heights, int dtype, int maps, int rows, int cols, int pitch,
                           int th) {
    const size_t es = dtype == TAMOLS_F64 ? 8 : 4;
    if (!h->encode || ((uintptr_t)d_heights & 15) || ((size_t)pitch * es) % 16) return false;
    if (h->tm_ptr == d_heights && h->tm_dtype == dtype && h->tm_maps == maps && h->tm_rows == rows &&
        h->tm_cols == cols && h->tm_pitch == pitch && h->tm_th == th)
        return true;
    const cuuint64_t gdim[3] = {(cuuint64_t)cols, (cuuint64_t)rows, (cuuint64_t)maps};
    const cuuint64_t gstride[2] = {(cuuint64_t)pitch * es, (cuuint64_t)pitch * es * (cuuint64_t)rows};
    const cuuint32_t box[3] = {BW, (cuuint32_t)(th + 2 * HALO), 1};
    const cuuint32_t estr[3] = {1, 1, 1};
    // CU_TENSOR_MAP_DATA_TYPE_FLOAT64 / FLOAT32: the copy is bit-exact for either
    const CUresult r = h->encode(&h->tmap, dtype == TAMOLS_F64 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT64 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32,
                                 3, const_cast<void *>(d_heights), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                 CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                                 CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        h->tm_ptr = nullptr;
        return false;
    }
    h->tm_ptr = d_heights; h->tm_dtype = dtype; h->tm_maps = maps; h->tm_rows = rows; h->tm_cols = cols; h->tm_pitch = pitch; h->tm_th = th;
    return true;
}

// k_tamols_window (one block per job) once the jobs alone fill the GPU several times over; k_tamols_score (a job's
// window split into tiles over many SMs) for the latency regime.  Same results either way.
static bool use_window_kernel(const tamols_handle *h, int n_jobs, int max_nc) {
    if (max_nc > WBAND) return false;  // a band is at least one whole row of 16-bit offsets
    if (h->env_kernel) return h->env_kernel == 2;
    return n_jobs >= 4 * h->num_sms;
}
static int launch_window(tamols_handle *h, const Args &a, int dtype) {
    // 128 threads, 8 blocks per SM.  Measured alternatives (8192 jobs, 64 x 64 windows, B200): 64 threads 530 us,
    // 128 threads at 6 / 8 / 10 blocks per SM (80 / 64 / 48 registers) 487 / 436 / 435 us.
    if (dtype == TAMOLS_F64) k_tamols_window<double, 128><<<a.n_jobs, 128, 0, h->stream>>>(a);
    else k_tamols_window<float, 128><<<a.n_jobs, 128, 0, h->stream>>>(a);
    TCU(cudaGetLastError());
    h->launches += 1;
    h->last_kernel = "window";
    return TAMOLS_OK;
}

static int launch_score(tamols_handle *h, const void *d_heights, int dtype, int maps, int rows, int cols, int pitch,
                        const double *d_xy, const tamols_leg_job *d_jobs, int n_jobs, int max_nr, int max_nc,
                        tamols_result *d_results, double *d_scores, int64_t scores_stride,
                        unsigned int *host_flag = nullptr, unsigned int flag_value = 0) {
    // tile height: 32-row tiles once they alone give every SM a block, else 8-row tiles (latency regime)
    const long long tiles32 = (long long)n_jobs * ((max_nr + 31) / 32) * ((max_nc + TW - 1) / TW);
    int th = tiles32 >= (long long)h->num_sms ? 32 : 8;
    // a window of at most 256 candidates that fits one 32 x 32 tile (one robot's 13 x 7 sensor maps): ONE tile per job,
    // which finishes without the partial / ticket / merge round trips
    if (max_nr <= 32 && max_nc <= TW && max_nr * max_nc <= THREADS && max_nr > 8) th = 32;
    if (h->env_tile_rows) th = h->env_tile_rows;
    if (use_window_kernel(h, n_jobs, max_nc)) {  // no tiles: no partials, tickets or tensor map
        Args a;
        a.p = h->p;
        a.k = h->derived;
        a.heights = d_heights; a.xy = d_xy; a.jobs = d_jobs; a.results = d_results; a.scores = d_scores;
        a.scores_stride = scores_stride; a.partials = nullptr; a.tickets = nullptr;
        a.rows = rows; a.cols = cols; a.pitch = pitch; a.max_tiles = 0;
        a.host_flag = host_flag; a.flag_value = flag_value; a.n_jobs = n_jobs;
        return launch_window(h, a, dtype);
    }
    const int max_tiles = ((max_nr + th - 1) / th) * ((max_nc + TW - 1) / TW);
    const size_t need = (size_t)n_jobs * max_tiles;
    if (need > h->partial_cap) {
        if (h->d_partials) {
            TCU(cudaStreamSynchronize(h->stream));
            TCU(cudaFree(h->d_partials));
            h->d_partials = nullptr;
            h->partial_cap = 0;
        }
        TCU(cudaMalloc(&h->d_partials, need * sizeof(Partial)));
        h->partial_cap = need;
    }
    Args a;
    a.p = h->p;
    a.k = h->derived;
    a.heights = d_heights; a.xy = d_xy; a.jobs = d_jobs; a.results = d_results; a.scores = d_scores;
    a.scores_stride = scores_stride; a.partials = h->d_partials; a.tickets = h->d_tickets;
    a.rows = rows; a.cols = cols; a.pitch = pitch; a.max_tiles = max_tiles;
    a.host_flag = host_flag; a.flag_value = flag_value; a.n_jobs = n_jobs;
    const bool tma = get_tensor_map(h, d_heights, dtype, maps, rows, cols, pitch, th);
    const dim3 grid(max_tiles, n_jobs);
#define TAMOLS_LAUNCH(HT, TMA, TH_) k_tamols_score<HT, TMA, TH_><<<grid, THREADS, 0, h->stream>>>(h->tmap, a)
    if (dtype == TAMOLS_F64) {
        if (th == 32) { if (tma) TAMOLS_LAUNCH(double, true, 32); else TAMOLS_LAUNCH(double, false, 32); }
        else { if (tma) TAMOLS_LAUNCH(double, true, 8); else TAMOLS_LAUNCH(double, false, 8); }
    } else {
        if (th == 32) { if (tma) TAMOLS_LAUNCH(float, true, 32); else TAMOLS_LAUNCH(float, false, 32); }
        else { if (tma) TAMOLS_LAUNCH(float, true, 8); else TAMOLS_LAUNCH(float, false, 8); }
    }
#undef TAMOLS_LAUNCH
    TCU(cudaGetLastError());
    h->launches += 1;
    h->last_kernel = th == 32 ? "tile32" : "tile8";
    return TAMOLS_OK;
}

extern "C" int tamols_score_device(tamols_handle *h, const void *d_heights, int32_t height_dtype, int32_t num_maps,
                                   int32_t rows, int32_t cols, int32_t pitch, const double *d_xy,
                                   const tamols_leg_job *d_jobs, int32_t n_jobs, int32_t max_nr, int32_t max_nc,
                                   tamols_result *d_results, double *d_scores, int64_t scores_stride) {
    if (!h || !d_heights || !d_jobs || !d_results) return tfail(TAMOLS_E_ARG, "null argument");
    if (height_dtype != TAMOLS_F64 && height_dtype != TAMOLS_F32) return tfail(TAMOLS_E_ARG, "unknown height_dtype");
    if (num_maps < 1 || rows < 1 || cols < 1 || pitch < cols) return tfail(TAMOLS_E_ARG, "bad map shape");
    if (n_jobs < 1 || n_jobs > h->max_jobs) return tfail(TAMOLS_E_ARG, "n_jobs exceeds max_jobs");
    if (max_nr < 1 || max_nc < 1 || max_nr > rows || max_nc > cols) return tfail(TAMOLS_E_ARG, "bad window bound");
    if (d_scores && scores_stride < (int64_t)max_nr * max_nc) return tfail(TAMOLS_E_ARG, "scores_stride too small");
    TCU(cudaSetDevice(h->device));
    return launch_score(h, d_heights, height_dtype, num_maps, rows, cols, pitch, d_xy, d_jobs, n_jobs, max_nr, max_nc,
                        d_results, d_scores, scores_stride);
}

// ---------------------------------------------------------------- tamols_score_device_static (include/tamols.h)
// The scratch of k_tamols_score (arrival tickets + per-tile partials) in a file-scope device array instead of the
// handle's cudaMalloc'ed buffers: the entry point allocates nothing.  Every GPU that loads the library holds its own
// copy of the array until the process exits.  In HOST code `g_static_scratch` is only the symbol's host-side stand-in,
// never a device address: the address a kernel may use comes from cudaGetSymbolAddress, resolved once per device with
// that device current (static_scratch below) and handed to the kernel through Args like any other pointer.
constexpr size_t STATIC_TICKET_BYTES = ((size_t)(TAMOLS_STATIC_MAX_JOBS + 1) * sizeof(unsigned int) + 255) / 256 * 256;
static_assert(sizeof(Partial) == 16, "TAMOLS_STATIC_SCRATCH_BYTES in include/tamols.h assumes 16-byte partials");
static_assert(STATIC_TICKET_BYTES + (size_t)TAMOLS_STATIC_MAX_PARTIALS * sizeof(Partial) == TAMOLS_STATIC_SCRATCH_BYTES,
              "scratch layout and TAMOLS_STATIC_SCRATCH_BYTES disagree");
static __device__ __align__(256) unsigned char g_static_scratch[TAMOLS_STATIC_SCRATCH_BYTES];

constexpr int STATIC_MAX_DEVICES = 64;
static std::mutex g_static_mu;
static void *g_static_ptr[STATIC_MAX_DEVICES];  // device address of g_static_scratch on device d, nullptr = not resolved yet

// device address of the array on `device`, which must be the CURRENT device (the symbol resolves per context)
static int static_scratch(int device, unsigned char **out) {
    if (device < 0 || device >= STATIC_MAX_DEVICES) return tfail(TAMOLS_E_ARG, "device ordinal beyond the static scratch table");
    std::lock_guard<std::mutex> lock(g_static_mu);
    if (!g_static_ptr[device]) {
        void *p = nullptr;
        TCU(cudaGetSymbolAddress(&p, g_static_scratch));
        if (!p || ((uintptr_t)p & 255)) return tfail(TAMOLS_E_CUDA, "cudaGetSymbolAddress returned an unusable address");
        g_static_ptr[device] = p;
    }
    *out = static_cast<unsigned char *>(g_static_ptr[device]);
    return TAMOLS_OK;
}

extern "C" int tamols_score_device_static(tamols_handle *h, const void *d_heights, int32_t height_dtype, int32_t num_maps,
                                          int32_t rows, int32_t cols, int32_t pitch, const double *d_xy,
                                          const tamols_leg_job *d_jobs, int32_t n_jobs, int32_t max_nr, int32_t max_nc,
                                          tamols_result *d_results, double *d_scores, int64_t scores_stride) {
    if (!h || !d_heights || !d_jobs || !d_results) return tfail(TAMOLS_E_ARG, "null argument");
    if (height_dtype != TAMOLS_F64 && height_dtype != TAMOLS_F32) return tfail(TAMOLS_E_ARG, "unknown height_dtype");
    if (num_maps < 1 || rows < 1 || cols < 1 || pitch < cols) return tfail(TAMOLS_E_ARG, "bad map shape");
    if (n_jobs < 1) return tfail(TAMOLS_E_ARG, "n_jobs must be >= 1");
    if (n_jobs > TAMOLS_STATIC_MAX_JOBS)
        return tfail(TAMOLS_E_ARG, "tamols_score_device_static: n_jobs = " + std::to_string(n_jobs) +
                                       " exceeds TAMOLS_STATIC_MAX_JOBS = " + std::to_string(TAMOLS_STATIC_MAX_JOBS) +
                                       " (use tamols_score_device, which allocates)");
    if (max_nr < 1 || max_nc < 1 || max_nr > rows || max_nc > cols) return tfail(TAMOLS_E_ARG, "bad window bound");
    if (d_scores && scores_stride < (int64_t)max_nr * max_nc) return tfail(TAMOLS_E_ARG, "scores_stride too small");
    // tile height: the rule of launch_score, restated here so that the existing path stays untouched
    const long long tiles32 = (long long)n_jobs * ((max_nr + 31) / 32) * ((max_nc + TW - 1) / TW);
    int th = tiles32 >= (long long)h->num_sms ? 32 : 8;
    if (max_nr <= 32 && max_nc <= TW && max_nr * max_nc <= THREADS && max_nr > 8) th = 32;
    if (h->env_tile_rows) th = h->env_tile_rows;
    const int max_tiles = ((max_nr + th - 1) / th) * ((max_nc + TW - 1) / TW);
    const long long need = (long long)n_jobs * max_tiles;
    if (need > TAMOLS_STATIC_MAX_PARTIALS)
        return tfail(TAMOLS_E_ARG, "tamols_score_device_static: " + std::to_string(n_jobs) + " jobs x " +
                                       std::to_string(max_tiles) + " tiles = " + std::to_string(need) +
                                       " partials exceed TAMOLS_STATIC_MAX_PARTIALS = " +
                                       std::to_string(TAMOLS_STATIC_MAX_PARTIALS) + " (use tamols_score_device, which allocates)");
    TCU(cudaSetDevice(h->device));
    unsigned char *scratch = nullptr;
    if (int rc = static_scratch(h->device, &scratch)) return rc;
    Args a;
    a.p = h->p;
    a.k = h->derived;
    a.heights = d_heights; a.xy = d_xy; a.jobs = d_jobs; a.results = d_results; a.scores = d_scores;
    a.scores_stride = scores_stride;
    a.tickets = reinterpret_cast<unsigned int *>(scratch);
    a.partials = reinterpret_cast<Partial *>(scratch + STATIC_TICKET_BYTES);
    a.rows = rows; a.cols = cols; a.pitch = pitch; a.max_tiles = max_tiles;
    a.host_flag = nullptr; a.flag_value = 0; a.n_jobs = n_jobs;
    // the array keeps whatever the last call (any handle, any stream, possibly one that aborted) left: the tickets must
    // start at zero, the partials are written by every tile before the merge reads them
    TCU(cudaMemsetAsync(a.tickets, 0, ((size_t)n_jobs + 1) * sizeof(unsigned int), h->stream));
    const bool tma = get_tensor_map(h, d_heights, height_dtype, num_maps, rows, cols, pitch, th);
    const dim3 grid(max_tiles, n_jobs);
#define TAMOLS_LAUNCH(HT, TMA, TH_) k_tamols_score<HT, TMA, TH_><<<grid, THREADS, 0, h->stream>>>(h->tmap, a)
    if (height_dtype == TAMOLS_F64) {
        if (th == 32) { if (tma) TAMOLS_LAUNCH(double, true, 32); else TAMOLS_LAUNCH(double, false, 32); }
        else { if (tma) TAMOLS_LAUNCH(double, true, 8); else TAMOLS_LAUNCH(double, false, 8); }
    } else {
        if (th == 32) { if (tma) TAMOLS_LAUNCH(float, true, 32); else TAMOLS_LAUNCH(float, false, 32); }
        else { if (tma) TAMOLS_LAUNCH(float, true, 8); else TAMOLS_LAUNCH(float, false, 8); }
    }
#undef TAMOLS_LAUNCH
    TCU(cudaGetLastError());
    h->launches += 1;
    h->last_kernel = th == 32 ? "tile32" : "tile8";
    return TAMOLS_OK;
}

static size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

extern "C" int tamols_score_host(tamols_handle *h, const double *heights, int32_t num_maps, int32_t rows, int32_t cols,
                                 const double *xy, const tamols_leg_job *jobs, int32_t n_jobs, tamols_result *results,
                                 double *scores, int64_t scores_stride) {
    if (!h || !heights || !jobs || !results) return tfail(TAMOLS_E_ARG, "null argument");
    if (num_maps < 1 || rows < 1 || cols < 1) return tfail(TAMOLS_E_ARG, "bad map shape");
    if (n_jobs < 1 || n_jobs > h->max_jobs) return tfail(TAMOLS_E_ARG, "n_jobs exceeds max_jobs");
    int max_nr = 0, max_nc = 0;
    for (int k = 0; k < n_jobs; ++k) {
        const tamols_leg_job &j = jobs[k];
        if (j.map < 0 || j.map >= num_maps || j.nr < 1 || j.nc < 1 || j.r0 < 0 || j.c0 < 0 || j.r0 + j.nr > rows ||
            j.c0 + j.nc > cols)
            return tfail(TAMOLS_E_ARG, "leg job window outside its map");
        max_nr = j.nr > max_nr ? j.nr : max_nr;
        max_nc = j.nc > max_nc ? j.nc : max_nc;
    }
    if (scores && scores_stride < (int64_t)max_nr * max_nc) return tfail(TAMOLS_E_ARG, "scores_stride too small");
    TCU(cudaSetDevice(h->device));
    // staging layout (device and pinned host mirror): heights (row pitch padded to 16 B so the TMA path applies) | xy |
    // jobs | results | scores
    const int pitch = (cols + 1) & ~1;
    const size_t cells = (size_t)num_maps * rows;
    const size_t o_h = 0, sz_h = cells * pitch * 8;
    const size_t o_xy = align_up(o_h + sz_h, 256), sz_xy = xy ? cells * cols * 16 : 0;
    const size_t o_j = align_up(o_xy + sz_xy, 256), sz_j = (size_t)n_jobs * sizeof(tamols_leg_job);
    const size_t o_r = align_up(o_j + sz_j, 256), sz_r = (size_t)n_jobs * sizeof(tamols_result);
    const size_t o_s = align_up(o_r + sz_r, 256), sz_s = scores ? (size_t)n_jobs * scores_stride * 8 : 0;
    const size_t total = o_s + sz_s;
    if (total > h->stage_cap) {
        TCU(cudaStreamSynchronize(h->stream));
        if (h->d_stage) cudaFree(h->d_stage);
        if (h->h_stage) cudaFreeHost(h->h_stage);
        h->d_stage = nullptr; h->h_stage = nullptr; h->stage_cap = 0; h->tm_ptr = nullptr;
        const size_t cap = align_up(total + total / 4, 4096);
        TCU(cudaMalloc(&h->d_stage, cap));
        TCU(cudaMallocHost(&h->h_stage, cap));
        h->stage_cap = cap;
    }
    char *hs = h->h_stage, *ds = h->d_stage;
    if (pitch == cols) {
        memcpy(hs + o_h, heights, sz_h);
    } else {
        for (size_t r = 0; r < cells; ++r) {
            double *dst = reinterpret_cast<double *>(hs + o_h) + r * pitch;
            memcpy(dst, heights + r * cols, (size_t)cols * 8);
            dst[cols] = 0.0;
        }
    }
    if (xy) memcpy(hs + o_xy, xy, sz_xy);
    memcpy(hs + o_j, jobs, sz_j);
    TCU(cudaMemcpyAsync(ds, hs, o_j + sz_j, cudaMemcpyHostToDevice, h->stream));  // one H2D copy
    double *d_scores = scores ? reinterpret_cast<double *>(ds + o_s) : nullptr;
    // Without per-candidate scores (the product call) the results go straight into a mapped pinned block and the last
    // finished job raises a flag there that the host spins on: no D2H copy command, no stream-synchronise wake-up.
    // With scores (tests) or TAMOLS_ZEROCOPY=0: device results + one D2H copy + cudaStreamSynchronize.
    const bool mapped_out = !scores && h->env_zerocopy;
    if (mapped_out) {
        volatile unsigned long long *pairs = reinterpret_cast<volatile unsigned long long *>(h->m_out);
        unsigned int seq = ++h->seq;
        if (seq == 0) seq = ++h->seq;  // the block starts zeroed: 0 never names a call
        if (int rc = launch_score(h, ds + o_h, TAMOLS_F64, num_maps, rows, cols, pitch,
                                  xy ? reinterpret_cast<const double *>(ds + o_xy) : nullptr,
                                  reinterpret_cast<const tamols_leg_job *>(ds + o_j), n_jobs, max_nr, max_nc,
                                  reinterpret_cast<tamols_result *>(ds + o_r), nullptr, 0,
                                  reinterpret_cast<unsigned int *>(h->md_out), seq))
            return rc;
        const size_t words = (size_t)n_jobs * (sizeof(tamols_result) / 4);
        uint32_t *dst = reinterpret_cast<uint32_t *>(results);
        unsigned long long spins = 0;
        for (size_t k = 0; k < words; ++k) {
            unsigned long long v;
            while ((unsigned int)((v = pairs[k]) >> 32) != seq) {
                __builtin_ia32_pause();
                if ((++spins & 0xFFFFF) == 0) {  // ~ms: make sure the kernel is still alive rather than spin forever
                    const cudaError_t q = cudaStreamQuery(h->stream);
                    if (q != cudaErrorNotReady && (unsigned int)(pairs[k] >> 32) != seq)
                        return tfail(TAMOLS_E_CUDA, q != cudaSuccess ? std::string("scoring kernel: ") + cudaGetErrorString(q)
                                                                     : std::string("scoring finished without publishing its results"));
                }
            }
            dst[k] = (uint32_t)v;
        }
        return TAMOLS_OK;

    }
    if (int rc = launch_score(h, ds + o_h, TAMOLS_F64, num_maps, rows, cols, pitch,
                              xy ? reinterpret_cast<const double *>(ds + o_xy) : nullptr,
                              reinterpret_cast<const tamols_leg_job *>(ds + o_j), n_jobs, max_nr, max_nc,
                              reinterpret_cast<tamols_result *>(ds + o_r), d_scores, scores_stride))
        return rc;
    TCU(cudaMemcpyAsync(hs + o_r, ds + o_r, (scores ? o_s + sz_s : o_r + sz_r) - o_r, cudaMemcpyDeviceToHost, h->stream));
    TCU(cudaStreamSynchronize(h->stream));
    memcpy(results, hs + o_r, sz_r);
    if (scores) {
        // candidates beyond a job's own window are not written by the kernel: report them as rejected
        for (int k = 0; k < n_jobs; ++k) {
            const int64_t n = (int64_t)jobs[k].nr * jobs[k].nc;
            const double *src = reinterpret_cast<const double *>(hs + o_s) + (size_t)k * scores_stride;
            double *dst = scores + (size_t)k * scores_stride;
            memcpy(dst, src, (size_t)n * 8);
            for (int64_t i = n; i < scores_stride; ++i) dst[i] = INFINITY;
        }
    }
    return TAMOLS_OK;
}

extern "C" int tamols_seed_jobs(tamols_handle *h, const tamols_seed_batch *b) {
    if (!h || !b) return tfail(TAMOLS_E_ARG, "null argument");
    if (b->num_robots < 1 || b->hist_len < 1) return tfail(TAMOLS_E_ARG, "num_robots and hist_len must be >= 1");
    if (!b->d_base_position || !b->d_base_euler || !b->d_base_lin_vel || !b->d_ref_lin_vel || !b->d_hips ||
        !b->d_com_height || !b->d_stance_time || !b->d_lift_off_z || !b->d_vel_hist || !b->d_hist_count || !b->d_ref_feet)
        return tfail(TAMOLS_E_ARG, "missing device pointer in tamols_seed_batch");
    if (b->d_jobs && (b->map_rows < 1 || b->map_cols < 1 || b->win_rows < 1 || b->win_cols < 1))
        return tfail(TAMOLS_E_ARG, "bad map / window shape");
    TCU(cudaSetDevice(h->device));
    const int threads = 128, blocks = (b->num_robots + threads - 1) / threads;
    k_seed_jobs<<<blocks, threads, 0, h->stream>>>(*b);
    TCU(cudaGetLastError());
    h->launches += 1;
    return TAMOLS_OK;
}

extern "C" int tamols_fit_grid(const double *xy, int32_t rows, int32_t cols, double origin[2], double row_axis[2],
                               double col_axis[2]) {
    if (!xy || !origin || !row_axis || !col_axis || rows < 1 || cols < 1) return tfail(TAMOLS_E_ARG, "bad argument");
    auto at = [&](int r, int c, int k) { return xy[((size_t)r * cols + c) * 2 + k]; };
    origin[0] = at(0, 0, 0);
    origin[1] = at(0, 0, 1);
    const bool hr = rows > 1, hc = cols > 1;
    if (hr) for (int k = 0; k < 2; ++k) row_axis[k] = (at(rows - 1, 0, k) - origin[k]) / (double)(rows - 1);
    if (hc) for (int k = 0; k < 2; ++k) col_axis[k] = (at(0, cols - 1, k) - origin[k]) / (double)(cols - 1);
    if (!hr && !hc) { row_axis[0] = 1.0; row_axis[1] = 0.0; col_axis[0] = 0.0; col_axis[1] = 1.0; }
    else if (!hr) { row_axis[0] = -col_axis[1]; row_axis[1] = col_axis[0]; }
    else if (!hc) { col_axis[0] = row_axis[1]; col_axis[1] = -row_axis[0]; }
    const double lr = hypot(row_axis[0], row_axis[1]), lc = hypot(col_axis[0], col_axis[1]);
    const double scale = lr < lc ? lr : lc;
    double worst = 0.0;
    for (int r = 0; r < rows; ++r)
        for (int c = 0; c < cols; ++c)
            for (int k = 0; k < 2; ++k) {
                const double m = origin[k] + (double)r * row_axis[k] + (double)c * col_axis[k];
                const double e = fabs(m - at(r, c, k));
                worst = e > worst ? e : worst;
            }
    if (!(scale > 0.0) || worst > 1e-6 * (scale > 1e-9 ? scale : 1e-9) + 1e-9)
        return tfail(TAMOLS_E_ARG, "heightmap cells are not a regular grid: the analytic nearest-cell lookup does not apply");
    if (fabs(row_axis[0] * col_axis[0] + row_axis[1] * col_axis[1]) > 1e-6 * scale * scale)
        return tfail(TAMOLS_E_ARG, "heightmap grid axes are not orthogonal");
    return TAMOLS_OK;
}

extern "C" int tamols_score_sensors(tamols_handle *h, const tamols_sensor_leg *legs, int32_t n_legs,
                                    tamols_result *results) {
    if (!h || !legs || !results) return tfail(TAMOLS_E_ARG, "null argument");
    if (n_legs < 1 || n_legs > h->max_jobs) return tfail(TAMOLS_E_ARG, "n_legs exceeds max_jobs");
    int R = 0, Cc = 0;
    for (int k = 0; k < n_legs; ++k) {
        if (!legs[k].data || legs[k].rows < 1 || legs[k].cols < 1) return tfail(TAMOLS_E_ARG, "bad sensor map");
        R = legs[k].rows > R ? legs[k].rows : R;
        Cc = legs[k].cols > Cc ? legs[k].cols : Cc;
    }
    const size_t cells = (size_t)R * Cc;
    h->sensor_heights.assign((size_t)n_legs * cells, 0.0);
    h->sensor_xy.assign((size_t)n_legs * cells * 2, 0.0);
    h->sensor_jobs.resize(n_legs);
    std::vector<double> dense;
    for (int k = 0; k < n_legs; ++k) {
        const tamols_sensor_leg &l = legs[k];
        dense.resize((size_t)l.rows * l.cols * 2);
        double *hz = h->sensor_heights.data() + k * cells, *hxy = h->sensor_xy.data() + k * cells * 2;
        for (int r = 0; r < l.rows; ++r)
            for (int c = 0; c < l.cols; ++c) {
                const double *cell = l.data + ((size_t)r * l.cols + c) * 3;
                hz[(size_t)r * Cc + c] = cell[2];
                hxy[((size_t)r * Cc + c) * 2] = cell[0];
                hxy[((size_t)r * Cc + c) * 2 + 1] = cell[1];
                dense[((size_t)r * l.cols + c) * 2] = cell[0];
                dense[((size_t)r * l.cols + c) * 2 + 1] = cell[1];
            }
        tamols_leg_job &j = h->sensor_jobs[k];
        memset(&j, 0, sizeof j);
        if (int rc = tamols_fit_grid(dense.data(), l.rows, l.cols, j.origin, j.row_axis, j.col_axis)) return rc;
        j.map = k; j.r0 = 0; j.c0 = 0; j.nr = l.rows; j.nc = l.cols; j.flags = l.flags;
        memcpy(j.seed, l.seed, sizeof j.seed);
        memcpy(j.hip, l.hip, sizeof j.hip);
        memcpy(j.diag_foot, l.diag_foot, sizeof j.diag_foot);
        memcpy(j.base_xy, l.base_xy, sizeof j.base_xy);
        memcpy(j.vel, l.vel, sizeof j.vel);
    }
    return tamols_score_host(h, h->sensor_heights.data(), n_legs, R, Cc, h->sensor_xy.data(), h->sensor_jobs.data(), n_legs,
                             results, nullptr, 0);
}
