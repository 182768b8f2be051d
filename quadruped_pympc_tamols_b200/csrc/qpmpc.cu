// qpmpc.cu -- sampling-MPC solve for sm_100a behind the C ABI of include/qpmpc.h.
//
// One solve = ONE launch on the handle's stream:
//   k_rollout       grid (ceil(N/128), robots): thread = one rollout (throughput regime, batched robots).
//   k_rollout_quad  grid (ceil(N/QB), robots): four lanes = one rollout, lane = leg (latency regime, one robot).
//               Either kernel draws its noise rows with Philox into shared-memory columns (or copies reference-
//               supplied rows), integrates the SRBD over the horizon, saturates the cost, and the block reduces
//               {min cost, argmin, sum exp(-(c - min_b)), sum exp(-(c - min_b)) * par[P]} from the parameters still
//               resident in shared memory.  The block that arrives LAST for its robot (arrival ticket) then runs
//   finalize_robot  in the same launch: merges the per-block partials with an online-softmax rescale, applies the
//               random-sampling / MPPI / CEM-MPPI update, evaluates the first-step GRF and the one-step
//               prediction, and (CEM) selects the elites and their per-parameter spread.  k_finalize (grid
//               (robots)) runs the same function as a second launch when the scratch does not fit or QPMPC_FUSED=0.
// qpmpc_solve (host buffers) sends its inputs in the kernel parameter block and receives the results as
// {word, sequence} pairs in mapped pinned memory: one stream command per solve, no copies, no synchronise.
// Reference semantics: quadruped_pympc/controllers/sampling/centroidal_nmpc_jax.py:629-1094 and
// centroidal_nmpc_jax_gait_adaptive.py:634-1131 (see include/qpmpc.h for the entry-point map).
// No CPU fallback: every entry point fails with QPMPC_E_NODEVICE / QPMPC_E_CUDA when the GPU path cannot run.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>

#include "../../include/qpmpc.h"
#include "philox.cuh"
#include "srbd.cuh"
#include "noise.cuh"
#include "rollout.cuh"
#include "rollout_pair.cuh"

namespace qp {

constexpr int PART_HDR = 4;  // per-block partial: [min cost, sum w, argmin cost, argmin index bits] + V[P]
constexpr int FIN_THREADS_MAX = 512;

// Everything one solve needs on the device; the rollout kernels and the finalize step share it.
struct SolveArgs {
    const float *state, *reference, *contact, *best, *sigma, *timing, *nominal;
    const int32_t *oswing;
    const float *eps, *freqs;  // reference-supplied noise (or the parameter matrix in costs_only mode)
    float *costs, *partials;
    qpmpc_result *result;
    float *best_out, *sigma_out;
    unsigned int *tickets;  // [robots] arrival counters of the fused finalize (zero between launches)
    // host-buffer solve (one robot): mapped host block the finalize step stores {word, ll_seq} pairs into, 8 bytes
    // per store (result record | best_out | sigma_out).  Every pair carries its own arrival flag, so no system-wide
    // fence and no separate done flag are needed: the host waits until each pair shows the current sequence number.
    uint2 *ll_out;
    unsigned int ll_seq;
    NoiseKey key;
    int n;            // rollouts per robot in this launch
    int costs_only;   // 1: jit_vectorized_rollout semantics (eps = parameters, no saturation, no partials)
    int nblk;         // rollout blocks per robot
    int robot_offset; // global id of robot 0 of this launch (Philox subsequence = robot_offset + r)
    int fused;        // 1: the last rollout block of a robot runs finalize_robot itself (one launch per solve)
    int chunk;        // rollouts per block of this launch
    int inline_in;    // bit 0: state / reference / contact / timing / nominal / oswing, bit 1: best, bit 2: sigma hold
                      // BYTE OFFSETS into the kernel's InlineIn parameter instead of device addresses
};
typedef SolveArgs RolloutArgs;
typedef SolveArgs FinalizeArgs;

// qpmpc_solve's ~1 KB of inputs (state | reference | contact | best | sigma | timing | nominal | oswing) ride in the
// kernel parameter block instead of a separate H2D copy: one command on the stream, no copy-engine -> compute
// dependency in front of the kernel.  Batched launches pass device pointers and ignore this block.
constexpr int INLINE_FLOATS = 608;
struct InlineIn {
    float v[INLINE_FLOATS];
};
__device__ __forceinline__ SolveArgs resolve_inputs(const SolveArgs &a_, const InlineIn &in) {
    SolveArgs a = a_;
    const char *base = reinterpret_cast<const char *>(in.v);
    if (a.inline_in & 1) {
        a.state = reinterpret_cast<const float *>(base + reinterpret_cast<size_t>(a_.state));
        a.reference = reinterpret_cast<const float *>(base + reinterpret_cast<size_t>(a_.reference));
        a.contact = reinterpret_cast<const float *>(base + reinterpret_cast<size_t>(a_.contact));
        a.timing = reinterpret_cast<const float *>(base + reinterpret_cast<size_t>(a_.timing));
        a.nominal = reinterpret_cast<const float *>(base + reinterpret_cast<size_t>(a_.nominal));
        a.oswing = reinterpret_cast<const int32_t *>(base + reinterpret_cast<size_t>(a_.oswing));
    }
    if (a.inline_in & 2) a.best = reinterpret_cast<const float *>(base + reinterpret_cast<size_t>(a_.best));
    if (a.inline_in & 4) a.sigma = reinterpret_cast<const float *>(base + reinterpret_cast<size_t>(a_.sigma));
    return a;
}

// ---------------------------------------------------------------- noise: noise.cuh (blocks of 12 values) ---
// step_frequencies_vec[i]: ..._gait_adaptive.py:687-692 (random_sampling), :834-838 (mppi), :994-999 (cem)
__device__ __forceinline__ float freq_sample(const DevCfg &c, const NoiseKey &nk, int i, uint32_t robot,
                                             float nominal, int oswing) {
    const uint32_t x = draw(nk, STREAM_FREQ, 0, i, robot).x;
    if (c.method == QPMPC_METHOD_CEM) {
        const int k = (int)(((uint64_t)x * 3ull) >> 32);
        const float v = k == 0 ? 0.0f : (k == 1 ? 0.2f : 0.4f);
        return __fadd_rn(__fmul_rn(v, (float)oswing), nominal);
    }
    const int k = (int)(((uint64_t)x * (uint64_t)c.nfreq) >> 32);
    if (c.method == QPMPC_METHOD_RANDOM && !oswing) return nominal;
    return c.freqs[k];
}

// ---------------------------------------------------------------- reductions -------------------------------
__device__ __forceinline__ void argmin_combine(float &c, int &i, float oc, int oi) {
    if (oc < c || (oc == c && oi < i)) {
        c = oc;
        i = oi;
    }
}

// lexicographic (cost, index) minimum over the block; result broadcast to all threads.  s_red: >= 64 floats.
__device__ __forceinline__ void block_argmin(float &c, int &i, float *s_red) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float oc = __shfl_xor_sync(0xffffffffu, c, o);
        const int oi = __shfl_xor_sync(0xffffffffu, i, o);
        argmin_combine(c, i, oc, oi);
    }
    __syncthreads();
    if (lane == 0) {
        s_red[wid] = c;
        s_red[32 + wid] = __int_as_float(i);
    }
    __syncthreads();
    c = s_red[0];
    i = __float_as_int(s_red[32]);
    for (int w = 1; w < nw; ++w) argmin_combine(c, i, s_red[w], __float_as_int(s_red[32 + w]));
}

__device__ __forceinline__ float block_sum(float v, float *s_red) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if (lane == 0) s_red[wid] = v;
    __syncthreads();
    float s = s_red[0];
    for (int w = 1; w < nw; ++w) s += s_red[w];
    return s;
}

// One SRBD step of ONE rollout shared by the four lanes of a quad (lane = leg): the lane shapes its own leg's force
// (in place: f becomes the ground reaction force), the quad sums force and torque with two shuffle stages, lanes
// 0..2 evaluate sin/cos of roll, pitch, yaw and hand them round, and every lane integrates the same body state.
// Must be called by all 32 lanes of the warp.  Used by k_rollout_quad and by the fused finalize's one-step prediction.
__device__ __forceinline__ float quad_sum(float v) {
    v += __shfl_xor_sync(0xffffffffu, v, 1);
    v += __shfl_xor_sync(0xffffffffu, v, 2);
    return v;
}
__device__ __forceinline__ void quad_leg_step(const DevCfg &c, float f[3], float st, float ref_fz, float footx,
                                              float footy, float footz, float x[12], float dt, int leg, int qbase) {
    shape_leg(c, f, st, st * (1.0f / 3.0f), ref_fz);
    // this leg's share of the net force and of the torque about the centre of mass, summed over the quad
    const float fzs = f[2] * st;
    const float rx = footx - x[0], ry = footy - x[1], rz = footz - x[2];
    const float lx = quad_sum(f[0]), ly = quad_sum(f[1]), lz = quad_sum(fzs);
    const float tx = quad_sum(ry * fzs - rz * f[1]);
    const float ty = quad_sum(rz * f[0] - rx * fzs);
    const float tz = quad_sum(rx * f[1] - ry * f[0]);
    // lanes 0, 1, 2 of the quad evaluate sin/cos of roll, pitch, yaw and hand them round
    const float ang = leg == 0 ? x[6] : (leg == 1 ? x[7] : x[8]);
    float sa, ca;
    sincos_fast(ang, &sa, &ca);
    const float sr = __shfl_sync(0xffffffffu, sa, qbase), cr = __shfl_sync(0xffffffffu, ca, qbase);
    const float sp = __shfl_sync(0xffffffffu, sa, qbase + 1), cp = __shfl_sync(0xffffffffu, ca, qbase + 1);
    const float sy = __shfl_sync(0xffffffffu, sa, qbase + 2), cy = __shfl_sync(0xffffffffu, ca, qbase + 2);
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
    x[0] += x[3] * dt;
    x[1] += x[4] * dt;
    x[2] += x[5] * dt;
    x[3] += (c.inv_mass * lx) * dt;
    x[4] += (c.inv_mass * ly) * dt;
    x[5] += (c.inv_mass * lz - 9.81f) * dt;
    x[6] += roll_rate * dt;
    x[7] += pitch_rate * dt;
    x[8] += yaw_rate * dt;
    x[9] += (c.Iinv[0] * gx + c.Iinv[1] * gy + c.Iinv[2] * gz) * dt;
    x[10] += (c.Iinv[3] * gx + c.Iinv[4] * gy + c.Iinv[5] * gz) * dt;
    x[11] += (c.Iinv[6] * gx + c.Iinv[7] * gy + c.Iinv[8] * gz) * dt;
}

// ---------------------------------------------------------------- finalize kernel --------------------------
// Shared: best_out[P] | sel (elite indices) [cem_elite] | red[64] | misc
// Finalize step of robot r, run by ALL threads of one block (any multiple of 32): either the last rollout block of
// the robot (fused, scratch = that block's parameter columns, no longer needed) or the stand-alone k_finalize.
// Partials and costs were written by other blocks: they are read through L2 (ld.global.cg).
constexpr int CEM_SLICES_MAX = 16;  // elite slices of the CEM spread pass (one noise block per thread and elite)
__host__ __device__ inline int finalize_acc_rows(int P, int cem_elite, int threads) {
    const int G = threads / (P / 4) > 1 ? threads / (P / 4) : 1;
    int J = threads / (P / 12) < CEM_SLICES_MAX ? threads / (P / 12) : CEM_SLICES_MAX;
    if (J < 1) J = 1;
    return cem_elite > 0 && J > G ? J : G;
}
constexpr int CEM_HIST_INTS = 264;  // 256 radix bins + the scalars of cem_select_elites
__host__ __device__ inline int finalize_smem_floats(int P, int nblk, int cem_elite, int threads) {
    return finalize_acc_rows(P, cem_elite, threads) * P + P + 64 + 3 * nblk + 1 + (cem_elite > 0 ? 2 * cem_elite + CEM_HIST_INTS : 0) +
           (int)(sizeof(qpmpc_result) / 4) + 8;
}

// The K smallest (cost, index) pairs of costs[0..N) -- argsort(costs)[:K] as a SET, ties at the threshold resolved
// towards the lower index like a stable sort -- written to sel[0..K) in ascending INDEX order (a deterministic order
// for the sums that follow).  Radix select on the order-preserving bit pattern of the float: four passes of 8 bits
// with a 256-bin histogram in shared memory find the K-th smallest value, whatever K is; the K sequential block-wide
// argmin passes this replaces cost 5 us each (C3: 10 elites = 50 us of a 106 us solve, 200 elites in the gait-adaptive
// variant = 1 ms).  All threads of the block; hist: 264 ints, tmp: K ints, red: 64 floats.
__device__ __forceinline__ uint32_t ordered_key(float c) {
    const uint32_t b = __float_as_uint(c);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ void cem_select_elites(const float *costs, int N, int K, int *sel, int *tmp, int *hist, float *red) {
    const int t = threadIdx.x, T = blockDim.x;
    uint32_t prefix = 0u, mask = 0u;
    int need = K;  // how many of the elements matching `prefix` under `mask` belong to the K smallest
    for (int shift = 24; shift >= 0; shift -= 8) {
        for (int b = t; b < 256; b += T) hist[b] = 0;
        __syncthreads();
        for (int i = t; i < N; i += T) {
            const uint32_t u = ordered_key(__ldcg(costs + i));
            if ((u & mask) == prefix) atomicAdd(&hist[(u >> shift) & 255u], 1);
        }
        __syncthreads();
        if (t < 32) {  // the bin in which the cumulative count reaches `need`
            int loc[8], sum = 0;
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                loc[k] = hist[8 * t + k];
                sum += loc[k];
            }
            int incl = sum;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int v = __shfl_up_sync(0xffffffffu, incl, o);
                if (t >= o) incl += v;
            }
            const unsigned m = __ballot_sync(0xffffffffu, incl >= need);
            const int first = __ffs(m) - 1;
            if (t == first) {
                int below = incl - sum, k = 0;
                while (below + loc[k] < need) below += loc[k++];
                hist[256] = 8 * t + k;   // the bin
                hist[257] = below;       // elements of this pass strictly below it
                hist[258] = loc[k];      // elements in it
            }
        }
        __syncthreads();
        prefix |= (uint32_t)hist[256] << shift;
        mask |= 255u << shift;
        need -= hist[257];
        __syncthreads();
    }
    // prefix = key of the K-th smallest cost; `need` of the hist[258] elements equal to it are wanted: those of lowest
    // index.  All of them unless the threshold value is shared by more elements than fit (only then: bisection on the
    // index, each round one counting pass -- costs saturated at 1e6 are the one realistic source of such ties).
    const int equal = hist[258];
    int idx_max = N - 1;
    if (equal > need) {
        int lo = 0, hi = N - 1;  // smallest idx_max with count(key == prefix && i <= idx_max) >= need
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            int cnt = 0;
            for (int i = t; i <= mid; i += T) cnt += ordered_key(__ldcg(costs + i)) == prefix ? 1 : 0;
            const float tot = block_sum((float)cnt, red);  // exact: counts <= N < 2^24
            if ((int)tot >= need) hi = mid; else lo = mid + 1;
            __syncthreads();
        }
        idx_max = lo;
    }
    if (t == 0) hist[259] = 0;
    __syncthreads();
    for (int i = t; i < N; i += T) {
        const uint32_t u = ordered_key(__ldcg(costs + i));
        if (u < prefix || (u == prefix && i <= idx_max)) tmp[atomicAdd(&hist[259], 1)] = i;
    }
    __syncthreads();
    for (int e = t; e < K; e += T) {  // rank by index: the arrival order above is not deterministic
        const int me = tmp[e];
        int rank = 0;
        for (int f = 0; f < K; ++f) rank += tmp[f] < me ? 1 : 0;
        sel[rank] = me;
    }
    __syncthreads();
}

#ifdef QPMPC_TIMING
#define QP_TICK(k) do { if (threadIdx.x == 0) { unsigned long long g_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g_)); tick_[k] = g_; } } while (0)
#else
#define QP_TICK(k)
#endif
__device__ void finalize_robot(const DevCfg &c, const FinalizeArgs &a, const int r, float *smem) {
    const int P = c.P, N = c.N, t = threadIdx.x;
#ifdef QPMPC_TIMING
    unsigned long long tick_[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#endif
    QP_TICK(0);
    // scratch: acc[G][P] | best[P] | red[64] | sc[nblk] | sw[nblk] | lb[nblk + 1] | sel[cem_elite] | result record
    const int G_ = finalize_acc_rows(P, c.method == QPMPC_METHOD_CEM ? c.cem_elite : 0, (int)blockDim.x);
    float *s_acc = smem;
    float *s_best = s_acc + G_ * P;
    float *s_red = s_best + P;
    float *s_sc = s_red + 64;
    float *s_sw = s_sc + a.nblk;
    int *s_lb = reinterpret_cast<int *>(s_sw + a.nblk);
    int *s_sel = s_lb + a.nblk + 1;  // cem: elite indices [K] | selection scratch [K] | radix histogram
    int *s_hist = s_sel + 2 * c.cem_elite;
    qpmpc_result *s_res = reinterpret_cast<qpmpc_result *>(s_sel + (c.method == QPMPC_METHOD_CEM ? 2 * c.cem_elite + CEM_HIST_INTS : 0));
    __syncthreads();  // fused call: every thread is done with the columns this scratch overlays
    const float *part = a.partials + (size_t)r * a.nblk * (PART_HDR + P);
    const int stride = PART_HDR + P;
    const uint32_t robot = (uint32_t)(r + a.robot_offset);

    // Warp 0 alone walks the block headers {min_b, sum_t w, min_b, argmin} (all loads of a lane in flight at once, one
    // L2 round trip), finds the global minimum / first argmin (nanargmin: lowest index among equal minima), rescales
    // each block by exp(-(min_b - min)) (online softmax: each rollout block left {min_b, sum_t w, V[P]} relative to
    // its own minimum; best + sum_i w_i eps_i = sum_i w_i par_i / sum_i w_i with w = softmin(costs), :828-836), sums
    // the denominators and compacts the live blocks -- shuffles only, ONE block barrier at the end.  Blocks whose
    // minimum sits far above the global one have a scale of exactly 0 (upstream those samples contribute 0 * eps):
    // they are skipped; the live ones are kept in ascending order (deterministic summation order).
    float *s_hdr = s_red;  // [0] min, [1] argmin, [2] denominator, [3] live count
    float *s_pre = s_red + 8;  // state[24] | contact_sequence[:, 0] (4): staged now, used by the prediction step below
    if (t >= 32 && t < 56) s_pre[t - 32] = a.state[r * 24 + t - 32];
    if (t >= 56 && t < 60) s_pre[t - 32] = a.contact[(size_t)r * 4 * c.H + (t - 56) * c.H];
    if (blockDim.x < 64 && t < 28)  // 32-thread blocks: the one warp does it
        s_pre[t] = t < 24 ? a.state[r * 24 + t] : a.contact[(size_t)r * 4 * c.H + (t - 24) * c.H];
    if (t < 32) {
        float bc = INFINITY;
        int bi = 0x7fffffff;
        constexpr int HB = 8;  // headers per lane and trip, all in flight before the first shared-memory store
        for (int b0 = t; b0 < a.nblk; b0 += 32 * HB) {
            float4 hd[HB];
#pragma unroll
            for (int j = 0; j < HB; ++j) {
                const int b = b0 + 32 * j;
                hd[j] = __ldcg(reinterpret_cast<const float4 *>(part + (size_t)(b < a.nblk ? b : 0) * stride));
            }
#pragma unroll
            for (int j = 0; j < HB; ++j) {
                const int b = b0 + 32 * j;
                if (b < a.nblk) {
                    s_sc[b] = hd[j].x;
                    s_sw[b] = hd[j].y;
                    argmin_combine(bc, bi, hd[j].z, __float_as_int(hd[j].w));
                }
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float oc = __shfl_xor_sync(0xffffffffu, bc, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            argmin_combine(bc, bi, oc, oi);
        }
        float dpart = 0.f;
        int nl = 0;
        if (c.method != QPMPC_METHOD_RANDOM) {
            for (int b0 = 0; b0 < a.nblk; b0 += 32) {
                const int b = b0 + t;
                float sc = 0.f;
                if (b < a.nblk) {
                    sc = __expf(-(s_sc[b] - bc));
                    s_sc[b] = sc;
                    dpart += s_sw[b] * sc;
                }
                const bool live = sc != 0.f;
                const unsigned m = __ballot_sync(0xffffffffu, live);
                if (live) s_lb[nl + __popc(m & ((1u << t) - 1u))] = b;
                nl += __popc(m);
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) dpart += __shfl_xor_sync(0xffffffffu, dpart, o);
        }
        if (t == 0) {
            s_hdr[0] = bc;
            s_hdr[1] = __int_as_float(bi);
            s_hdr[2] = dpart;
            s_hdr[3] = __int_as_float(nl);
        }
    }
    __syncthreads();
    const float bc = s_hdr[0];
    const int bi = __float_as_int(s_hdr[1]);
    QP_TICK(1);

    if (c.method == QPMPC_METHOD_RANDOM) {  // best = control_parameters_vec[best_index], :692
        if (a.eps) {
            for (int p = t; p < P; p += blockDim.x)
                s_best[p] = __fadd_rn(a.best[(size_t)r * P + p], a.eps[((size_t)r * N + bi) * P + p]);
        } else {  // the winner's row is drawn again, one noise block per thread
            const RowNoise rn = row_noise(c, bi);
            for (int b = t; b < P / NOISE_BLOCK; b += blockDim.x) {
                float e[NOISE_BLOCK];
                noise_block(rn, a.key, b, robot, e, [](int) { return 1.0f; });
#pragma unroll
                for (int k = 0; k < NOISE_BLOCK; ++k) {
                    const int p = param_of(c, b, k);
                    s_best[p] = __fadd_rn(a.best[(size_t)r * P + p], e[k]);
                }
            }
        }
    } else {
        const float den = s_hdr[2];
        QP_TICK(2);
        const int nlive = __float_as_int(s_hdr[3]);
        // sum_b scale_b V_b[p]: the block splits into G row groups of P/4 threads, each thread streams float4 slices of
        // every G-th live row through L2 with 16 loads in flight; the G group sums are then added in group order.
        const int P4 = P >> 2;
        const int G = max(1, (int)blockDim.x / P4);
        for (int q = t; q < G * P4; q += blockDim.x) {
            const int g = q / P4, p4 = q - g * P4;
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
            // fixed batches of MB predicated loads: a short row list must not fall into a one-load-at-a-time
            // remainder loop (12 rows per group at 139 blocks = 12 serial L2 round trips, 4.5 us, before this)
            constexpr int MB = 12;
            for (int k0 = g; k0 < nlive; k0 += MB * G) {
                float4 v[MB];
                float sc[MB];
#pragma unroll
                for (int j = 0; j < MB; ++j) {
                    const int k = k0 + j * G;
                    const bool ok = k < nlive;
                    const int b = s_lb[ok ? k : 0];
                    sc[j] = ok ? s_sc[b] : 0.f;
                    v[j] = __ldcg(reinterpret_cast<const float4 *>(part + (size_t)b * stride + PART_HDR) + p4);
                }
#pragma unroll
                for (int j = 0; j < MB; ++j) {  // ascending row order within the group, as before
                    if (k0 + j * G < nlive) {
                        acc.x = fmaf(sc[j], v[j].x, acc.x);
                        acc.y = fmaf(sc[j], v[j].y, acc.y);
                        acc.z = fmaf(sc[j], v[j].z, acc.z);
                        acc.w = fmaf(sc[j], v[j].w, acc.w);
                    }
                }
            }
            *reinterpret_cast<float4 *>(s_acc + g * P + 4 * p4) = acc;
        }
        __syncthreads();
        for (int p = t; p < P; p += blockDim.x) {
            float num = 0.f;
            for (int g = 0; g < G; ++g) num += s_acc[g * P + p];
            s_best[p] = num / den;
        }
    }
    // best rollout's step frequency: one lane of the last warp, off thread 0's critical path
    if (t == (int)blockDim.x - 1) {
        float bf;
        if (c.ga)
            bf = a.freqs ? a.freqs[(size_t)r * N + bi] : freq_sample(c, a.key, bi, robot, a.nominal[r], a.oswing[r]);
        else
            bf = (c.method == QPMPC_METHOD_CEM) ? 1.65f : 1.4f;  // centroidal_nmpc_jax.py:786,930,1083
        s_res->best_freq = bf;
    }
    __syncthreads();
    QP_TICK(3);
    constexpr int RW = (int)(sizeof(qpmpc_result) / 4);
    const bool ll = a.ll_out != nullptr && r == 0;
    auto ll_store = [&](int k, float v) {
        uint2 w = make_uint2(__float_as_uint(v), a.ll_seq);
        asm volatile("st.volatile.global.v2.u32 [%0], {%1, %2};" ::"l"(a.ll_out + k), "r"(w.x), "r"(w.y) : "memory");
    };
    for (int p = t; p < P; p += blockDim.x) {
        if (ll) ll_store(RW + p, s_best[p]);
        else a.best_out[(size_t)r * P + p] = s_best[p];
    }

    // first GRFs and one-step prediction from the new mean (centroidal_nmpc_jax.py:707-784: spline at step 0.0 with
    // horizon_leg = 1, the passed-in contact_sequence[:, 0]): warp 0 runs it as one quad step, lane = leg (the eight
    // quads of the warp compute the same thing; quad 0 writes).  The record is assembled in shared memory, then one
    // warp-wide store moves it out (it may live in mapped host memory: two 128-byte bursts).
    if (t < 32) {
        const int leg = t & 3, qbase = t & 28;
        const float *state = s_pre;
        const float st = s_pre[24 + leg];
        const float nst = quad_sum(st);
        const int index = chunk_index(c, 0.0f);
        const float q = chunk_tau(c, 0.0f, 1.0f, index);
        const CubicBasis B = cubic_basis(q);
        const float *pl = s_best + leg * c.np;
        float f[3];
        if (c.param == QPMPC_PARAM_ZERO_ORDER)
            spline_eval<QPMPC_PARAM_ZERO_ORDER>(c, [&](int j) { return pl[j]; }, 0, index, q, B, f);
        else if (c.param == QPMPC_PARAM_LINEAR)
            spline_eval<QPMPC_PARAM_LINEAR>(c, [&](int j) { return pl[j]; }, 0, index, q, B, f);
        else
            spline_eval<QPMPC_PARAM_CUBIC>(c, [&](int j) { return pl[j]; }, 0, index, q, B, f);
        float x[12];
#pragma unroll
        for (int k = 0; k < 12; ++k) x[k] = state[k];
        quad_leg_step(c, f, st, c.mg / nst, state[12 + 3 * leg], state[13 + 3 * leg], state[14 + 3 * leg], x, c.dts[0],
                      leg, qbase);
        if (t < 4) {
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                s_res->grf[3 * leg + k] = f[k];
                s_res->footholds[3 * leg + k] = 0.f;
                s_res->predicted_state[3 * leg + k] = leg == 0 ? x[k] : leg == 1 ? x[3 + k] : leg == 2 ? x[6 + k] : x[9 + k];
                s_res->predicted_state[12 + 3 * leg + k] = state[12 + 3 * leg + k];
            }
        }
        if (t == 0) {
            s_res->best_cost = bc;
            s_res->best_index = bi;
            s_res->reserved = 0;
        }
        QP_TICK(4);
#ifdef QPMPC_TIMING
        __syncwarp();
        if (t == 0)
            for (int k = 1; k < 5; ++k) s_res->footholds[k] = (float)(tick_[k] - tick_[k - 1]);
#endif
    }
    __syncthreads();
    {
        float *dst = reinterpret_cast<float *>(a.result + r);
        for (int k = t; k < RW; k += blockDim.x) {
            const float v = reinterpret_cast<const float *>(s_res)[k];
            if (ll) ll_store(k, v);
            else dst[k] = v;
        }
    }

    if (c.method == QPMPC_METHOD_CEM && a.sigma_out != nullptr) {
    // ---- CEM: elites = argsort(costs)[:K] (stable: ties by index), new sigma from their noise rows,
    // centroidal_nmpc_jax.py:1075-1081 (K=10, sqrt, clip [0.2,5]) / ..._gait_adaptive.py:1115-1121 (K=200, no sqrt, [0.5,3])
    const int K = min(c.cem_elite, N);
    const float *costs = a.costs + (size_t)r * N;
    // Shortcut for the radix select (18 us at C3: one block walking 10,000 costs through L2 four times): the K-th
    // smallest of the rollout blocks' MINIMA (their partial headers) is an upper bound tau of the K-th smallest cost
    // -- those K minima are K distinct costs <= tau -- so the elites are among the costs <= tau, about K of them, since
    // the K best rollouts of a solve mostly sit in K different blocks.  One pass collects them (12 loads in flight per
    // thread), a rank count by (cost, index) -- argsort's stable order -- picks the K.  Not applicable with fewer
    // blocks than elites, or when more costs pass than the candidate buffer holds (saturated costs): radix select.
    bool selected = false;
    if (a.nblk >= K && K <= 64) {
        float *s_bmin = s_sw;                      // the block weights are consumed by now
        float *cand_c = s_acc;                     // candidates: costs [CAP] | indices [CAP] (s_acc is free until the spread)
        int *cand_i = reinterpret_cast<int *>(s_acc) + (G_ * P) / 2;
        const int CAP = (G_ * P) / 2;
        for (int b = t; b < a.nblk; b += blockDim.x) s_bmin[b] = __ldcg(part + (size_t)b * stride);
        if (t == 0) s_hist[259] = 0;
        __syncthreads();
        for (int b = t; b < a.nblk; b += blockDim.x) {  // rank of this block's minimum among all (ties by block index)
            const float v = s_bmin[b];
            int rank = 0;
            for (int o = 0; o < a.nblk; ++o) {
                const float w = s_bmin[o];
                rank += (w < v || (w == v && o < b)) ? 1 : 0;
            }
            if (rank == K - 1) s_red[0] = v;
        }
        __syncthreads();
        const float tau = s_red[0];
        constexpr int SB = 12;
        for (int i0 = t; i0 < N; i0 += SB * blockDim.x) {
            float v[SB];
#pragma unroll
            for (int j = 0; j < SB; ++j) {
                const int i = i0 + j * blockDim.x;
                v[j] = i < N ? __ldcg(costs + i) : INFINITY;
            }
#pragma unroll
            for (int j = 0; j < SB; ++j) {
                const int i = i0 + j * blockDim.x;
                if (i < N && v[j] <= tau) {
                    const int slot = atomicAdd(&s_hist[259], 1);
                    if (slot < CAP) {
                        cand_c[slot] = v[j];
                        cand_i[slot] = i;
                    }
                }
            }
        }
        __syncthreads();
        const int M = s_hist[259];
        if (M >= K && M <= CAP) {  // (M < K only with NaN costs in the headers: leave it to the radix path)
            int *s_tmp = s_sel + K;
            for (int e = t; e < M; e += blockDim.x) {
                const float v = cand_c[e];
                const int vi = cand_i[e];
                int rank = 0;
                for (int o = 0; o < M; ++o) {
                    const float w = cand_c[o];
                    rank += (w < v || (w == v && cand_i[o] < vi)) ? 1 : 0;
                }
                if (rank < K) s_tmp[rank] = vi;
            }
            __syncthreads();
            for (int e = t; e < K; e += blockDim.x) {  // ascending index: the deterministic order of the sums below
                const int me = s_tmp[e];
                int rank = 0;
                for (int f = 0; f < K; ++f) rank += s_tmp[f] < me ? 1 : 0;
                s_sel[rank] = me;
            }
            selected = true;
        }
        __syncthreads();
    }
    if (!selected) cem_select_elites(costs, N, K, s_sel, s_sel + K, s_hist, s_red);
    // Per-parameter mean and spread of the elites' noise rows, two passes (jnp.cov: mean, then centred squares).  The
    // rows are drawn again: thread (j, b) owns noise block b of the elites k = j, j + J, ...; the J slice sums are then
    // added in slice order.  s_acc: [J][P] (order-index layout), s_best is free by now and keeps the means.
    {
        const int B = P / NOISE_BLOCK;
        const int J = max(1, min((int)blockDim.x / B, (int)CEM_SLICES_MAX));
        const float *sig = a.sigma + (size_t)r * P;
        float *s_mean = s_best;
        for (int pass = 0; pass < 2; ++pass) {
            for (int q = t; q < J * B; q += blockDim.x) {
                const int j = q / B, b = q - j * B;
                float acc[NOISE_BLOCK];
#pragma unroll
                for (int k = 0; k < NOISE_BLOCK; ++k) acc[k] = 0.f;
                for (int k0 = j; k0 < K; k0 += J) {
                    const int i = s_sel[k0];
                    float e[NOISE_BLOCK];
                    if (a.eps) {
#pragma unroll
                        for (int k = 0; k < NOISE_BLOCK; ++k) e[k] = a.eps[((size_t)r * N + i) * P + param_of(c, b, k)];
                    } else {
                        noise_block(row_noise(c, i), a.key, b, robot, e, [&](int k) { return sig[param_of(c, b, k)]; });
                    }
#pragma unroll
                    for (int k = 0; k < NOISE_BLOCK; ++k) {
                        const float d = e[k] - (pass == 0 ? 0.f : s_mean[param_of(c, b, k)]);
                        acc[k] += pass == 0 ? d : d * d;
                    }
                }
#pragma unroll
                for (int k = 0; k < NOISE_BLOCK; ++k) s_acc[j * P + b * NOISE_BLOCK + k] = acc[k];
            }
            __syncthreads();
            for (int o = t; o < P; o += blockDim.x) {
                float sum = 0.f;
                for (int j = 0; j < J; ++j) sum += s_acc[j * P + o];
                const int p = param_of(c, o / NOISE_BLOCK, o % NOISE_BLOCK);
                if (pass == 0) {
                    s_mean[p] = sum / (float)K;
                } else {
                    float v = sum / (float)(K - 1) + 1e-8f;
                    if (c.cem_sqrt) v = sqrtf(v);
                    v = v > c.cem_hi ? c.cem_hi : v;
                    v = v < c.cem_lo ? c.cem_lo : v;
                    if (ll) ll_store(RW + P + p, v);
                    else a.sigma_out[(size_t)r * P + p] = v;
                }
            }
            __syncthreads();
        }
    }
    }
}

__global__ void __launch_bounds__(FIN_THREADS_MAX) k_finalize(const __grid_constant__ DevCfg c, const FinalizeArgs a_,
                                                              const __grid_constant__ InlineIn in) {
    const FinalizeArgs a = resolve_inputs(a_, in);
    extern __shared__ float smem[];
    finalize_robot(c, a, blockIdx.x, smem);
}

// Arrival ticket of the fused finalize: true in the block that arrives last for its robot (all partials and costs of
// the robot are then visible through L2); that block also re-arms the counter for the next launch on the stream.
__device__ __forceinline__ bool last_block_of_robot(unsigned int *ticket, int nblk) {
    __shared__ unsigned int s_ticket;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_ticket = atomicAdd(ticket, 1u);
    __syncthreads();
    if (s_ticket != (unsigned)(nblk - 1)) return false;
    if (threadIdx.x == 0) *ticket = 0u;
    __threadfence();
    return true;
}

// ---------------------------------------------------------------- rollout kernel (throughput regime) -------
// Shared memory layout (floats):
//   w[T] | live[T] (u16) | par[P][T+1] | best[P] | sigma[P] (cem only) | state[24] | ref[12] | timing[4] | st[8H] |
//   basis[4H] | ref_fz[H] | index[H] (int) | mgk[8] | red[64]
// par[p][t] = best[p] + eps[t][p] is the parameter vector of thread t's rollout, one COLUMN per thread: the rollout's
// own reads are conflict-free, and with the row stride T+1 so are the row-wise reads of the softmin-weighted sum.
// (sigma only exists for cem_mppi: at P = 144, T = 128 leaving it out is what lets three blocks share an SM.)
__host__ __device__ inline int rollout_smem_floats(int P, int H, int T, bool cem) {
    return T + T / 2 + P * (T + 1) + (cem ? 2 : 1) * P + 24 + 12 + 4 + 8 * H + H + 4 * H + H + 8 + 40;
}

// V[p] = sum_t w_t par[p][t] over the block, written to part[PART_HDR + p].  At temperature 1 with costs of 1e3..1e5
// almost every weight underflows to exactly 0 (the reference multiplies those by eps and adds 0): the live (t, w)
// pairs are compacted in ascending t (deterministic summation order) and only they are visited.
template <int STRIDE>
__device__ __forceinline__ void weighted_rows(const float *s_par, float *s_w, unsigned short *s_live, float *s_red, float w, int P,
                                              float *part) {
    const int t = threadIdx.x, T = blockDim.x, lane = t & 31, wid = t >> 5, nw = T >> 5;
    const bool live = w > 0.f;
    const unsigned m = __ballot_sync(0xffffffffu, live);
    __syncthreads();  // s_red free again
    if (lane == 0) s_red[wid] = __int_as_float(__popc(m));
    __syncthreads();
    int base = 0, total = 0;
    for (int k = 0; k < nw; ++k) {
        const int cnt = __float_as_int(s_red[k]);
        base += k < wid ? cnt : 0;
        total += cnt;
    }
    if (live) {
        const int k = base + __popc(m & ((1u << lane) - 1u));
        s_live[k] = (unsigned short)t;
        s_w[k] = w;
    }
    __syncthreads();
    for (int p = t; p < P; p += T) {
        const float *row = s_par + p * STRIDE;
        float acc = 0.f;
        for (int k = 0; k < total; ++k) acc = fmaf(s_w[k], row[s_live[k]], acc);
        part[PART_HDR + p] = acc;
    }
}

// Block partials of the thread-per-rollout kernel with the fewest barriers: the same arithmetic in the same order as
// block_argmin + block_sum + weighted_rows (bit-identical partials), but each reduction owns its slice of s_red, so no
// barrier is needed to recycle it, and the softmin sum travels with the live-sample count: 3 barriers instead of 7
// in front of the arrival ticket.  s_red: [0, 8) cost minima | [8, 16) their indices | [16, 24) sums | [24, 32) counts.
template <int STRIDE>
__device__ __forceinline__ void rollout_block_partials(const DevCfg &c, const float *s_par, float *s_w, unsigned short *s_live,
                                                       float *s_red, float cost, bool valid, int i, int P, float *part) {
    const int t = threadIdx.x, T = blockDim.x, lane = t & 31, wid = t >> 5, nw = T >> 5;
    float bc = cost;
    int bi = valid ? i : 0x7fffffff;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float oc = __shfl_xor_sync(0xffffffffu, bc, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        argmin_combine(bc, bi, oc, oi);
    }
    if (lane == 0) {
        s_red[wid] = bc;
        s_red[8 + wid] = __int_as_float(bi);
    }
    __syncthreads();
    bc = s_red[0];
    bi = __float_as_int(s_red[8]);
    for (int w = 1; w < nw; ++w) argmin_combine(bc, bi, s_red[w], __float_as_int(s_red[8 + w]));
    if (c.method == QPMPC_METHOD_RANDOM) {
        if (t == 0) {
            part[0] = bc;
            part[1] = 0.f;
            part[2] = bc;
            part[3] = __int_as_float(bi);
        }
        return;
    }
    // softmin numerators relative to the block minimum (merged exactly by finalize_robot), :828-836
    const float w = valid ? __expf(-(cost - bc)) : 0.f;
    float v = w;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    const bool live = w > 0.f;
    const unsigned m = __ballot_sync(0xffffffffu, live);
    if (lane == 0) {
        s_red[16 + wid] = v;
        s_red[24 + wid] = __int_as_float(__popc(m));
    }
    __syncthreads();
    float sw = s_red[16];
    for (int k = 1; k < nw; ++k) sw += s_red[16 + k];
    int base = 0, total = 0;
    for (int k = 0; k < nw; ++k) {
        const int cnt = __float_as_int(s_red[24 + k]);
        base += k < wid ? cnt : 0;
        total += cnt;
    }
    if (t == 0) {
        part[0] = bc;
        part[1] = sw;
        part[2] = bc;
        part[3] = __int_as_float(bi);
    }
    if (live) {
        const int k = base + __popc(m & ((1u << lane) - 1u));
        s_live[k] = (unsigned short)t;
        s_w[k] = w;
    }
    __syncthreads();
    // finalize_robot turns sum_b scale_b V_b / sum_b scale_b sw_b into the updated parameters
    for (int p = t; p < P; p += T) {
        const float *row = s_par + p * STRIDE;
        float acc = 0.f;
        for (int k = 0; k < total; ++k) acc = fmaf(s_w[k], row[s_live[k]], acc);
        part[PART_HDR + p] = acc;
    }
}

template <int PARAM, bool GA, int T>
__global__ void __launch_bounds__(T, T == 128 ? 3 : T == 64 ? 6 : 1) k_rollout(const __grid_constant__ DevCfg c, const RolloutArgs a_,
                                                                  const __grid_constant__ InlineIn in) {
    const RolloutArgs a = resolve_inputs(a_, in);
    extern __shared__ __align__(16) float smem[];
    constexpr int STRIDE = T + 1;
    const int P = c.P, H = c.H, t = threadIdx.x;
    float *s_w = smem;
    unsigned short *s_live = reinterpret_cast<unsigned short *>(s_w + T);
    float *s_par = s_w + T + T / 2;
    float *s_best = s_par + P * STRIDE;
    float *s_sigma = s_best + P;
    float *s_state = s_sigma + (c.method == QPMPC_METHOD_CEM ? P : 0);
    float *s_ref = s_state + 24;
    float *s_timing = s_ref + 12;
    float *s_st = s_timing + 4;
    float *s_basis = s_st + 8 * H;  // float4 tables first: 16-byte aligned for every horizon
    float *s_reffz = s_basis + 4 * H;
    int *s_index = reinterpret_cast<int *>(s_reffz + H);
    float *s_mgk = s_reffz + 2 * H;
    float *s_red = s_mgk + 8;

    const int r = blockIdx.y;
    const int i = blockIdx.x * T + t;
    const bool valid = i < a.n;

    // ---- stage the per-robot problem and its step tables.  Every global load of the thread is issued before its
    // first store, and the small jobs (state, reference, step tables) go to different warps: with all of them on warp
    // 0 behind the parameter copy, that warp chained four global round trips while the others sat at the barrier
    // (9 % of the C5 kernel's stall samples on this one barrier, profiles/r4a_ncu_source_c5.txt).
    {
        const bool wide = T >= 128;
        const int role = wide ? (t >> 5) : 0, tl = wide ? (t & 31) : t;
        const bool cem = c.method == QPMPC_METHOD_CEM;
        const bool j_best = t < P, j_state = (!wide || role == 1) && tl < 24, j_ref = (!wide || role == 2) && tl < 12;
        const bool j_tab = !GA && (!wide || role == 3) && tl < H;
        const float *cs = a.contact + (size_t)r * 4 * H;
        float bv = 0.f, sv = 0.f, stv = 0.f, rfv = 0.f, s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
        if (j_best && !a.costs_only) bv = a.best[(size_t)r * P + t];
        if (j_best && cem && a.sigma) sv = a.sigma[(size_t)r * P + t];
        if (j_state) stv = a.state[r * 24 + tl];
        if (j_ref) rfv = a.reference[r * 24 + tl];
        if (j_tab) { s0 = cs[tl]; s1 = cs[H + tl]; s2 = cs[2 * H + tl]; s3 = cs[3 * H + tl]; }
        if (j_best) s_best[t] = bv;
        if (j_best && cem) s_sigma[t] = sv;
        if (j_state) s_state[tl] = stv;
        if (j_ref) s_ref[tl] = rfv;
        if (j_tab) {
            const int n = tl;
            s_reffz[n] = c.mg / (((s0 + s1) + s2) + s3);
            s_st[8 * n + 0] = s0; s_st[8 * n + 1] = s1; s_st[8 * n + 2] = s2; s_st[8 * n + 3] = s3;
            s_st[8 * n + 4] = s0 * (1.0f / 3.0f); s_st[8 * n + 5] = s1 * (1.0f / 3.0f);
            s_st[8 * n + 6] = s2 * (1.0f / 3.0f); s_st[8 * n + 7] = s3 * (1.0f / 3.0f);
            const int idx = chunk_index(c, (float)n);
            s_index[n] = idx;
            const float4 B = chunk_basis<PARAM>(chunk_tau(c, (float)n, (float)H, idx));
            *reinterpret_cast<float4 *>(s_basis + 4 * n) = B;
        }
        for (int p = t + T; p < P; p += T) {
            s_best[p] = a.costs_only ? 0.f : a.best[(size_t)r * P + p];
            if (cem) s_sigma[p] = a.sigma ? a.sigma[(size_t)r * P + p] : 0.f;
        }
        if (GA) {
            if (t < 5) s_mgk[t] = c.mg / (float)t;
            if (t < 4) s_timing[t] = a.timing[r * 4 + t];
        }
    }
    __syncthreads();

    // ---- this thread's parameter vector -> shared column t
    const uint32_t robot = (uint32_t)(r + a.robot_offset);
    if (a.eps) {
        const float *src = a.eps + ((size_t)r * a.n + (size_t)blockIdx.x * T) * P;
        const int rows = min(T, a.n - blockIdx.x * T);
        for (int e = t; e < T * P; e += T) {
            const int row = e / P, p = e - row * P;
            s_par[p * STRIDE + row] = row < rows ? s_best[p] + src[e] : 0.f;
        }
        __syncthreads();
    } else {
        float *dst = s_par + t;
        if (!valid) {  // tail block: zero columns keep everything finite
            for (int p = 0; p < P; ++p) dst[p * STRIDE] = 0.f;
        } else {
            const RowNoise rn = row_noise(c, i);
            for (int b = 0; b < P / NOISE_BLOCK; ++b) {
                float e[NOISE_BLOCK];
                noise_block(rn, a.key, b, robot, e, [&](int k) { return s_sigma[param_of(c, b, k)]; });
#pragma unroll
                for (int k = 0; k < NOISE_BLOCK; ++k) {
                    const int p = param_of(c, b, k);
                    dst[p * STRIDE] = s_best[p] + e[k];
                }
            }
        }
    }

    // ---- rollout
    float cost = INFINITY;
    if (valid) {
        const float *col = s_par + t;
        if (GA) {
            const float freq = a.freqs ? a.freqs[(size_t)r * a.n + i]
                                       : freq_sample(c, a.key, i, robot, a.nominal[r], a.oswing[r]);
            cost = rollout_gait_adaptive<PARAM, STRIDE>(c, s_state, s_ref, s_mgk, s_timing, freq, col);
        } else {
            const FixedTables tb = {s_st, s_reffz, s_basis, s_index};
            cost = rollout_fixed<PARAM, STRIDE>(c, s_state, s_ref, tb, col);
        }
        if (!a.costs_only) {  // centroidal_nmpc_jax.py:686-687
            cost = isnan(cost) ? 1000000.0f : cost;
            cost = isinf(cost) ? 1000000.0f : cost;
        }
        a.costs[(size_t)r * a.n + i] = cost;
    }
    if (a.costs_only) return;

    // ---- block partials
    float *part = a.partials + ((size_t)r * a.nblk + blockIdx.x) * (PART_HDR + P);
    static_assert(T <= 256, "rollout_block_partials keeps one s_red slot per warp in slices of 8");
    rollout_block_partials<STRIDE>(c, s_par, s_w, s_live, s_red, cost, valid, i, P, part);
    if (a.fused && last_block_of_robot(a.tickets + r, a.nblk)) finalize_robot(c, resolve_inputs(a_, in), r, s_par);
}

// ---------------------------------------------------------------- pair rollout kernel (throughput regime) --
// Zero-order parametrisation, fixed gait: TWO rollouts per thread in packed FP32 (rollout_pair.cuh), and NO parameter
// matrix anywhere.  The twelve forces of horizon step n are noise block n of the rollout's row (noise.cuh): the thread
// draws them in the same basic block in which it integrates the step, uses them and forgets them.  A block owns
// `chunk` = 2 T passes rollouts of one robot (all N of them when robots alone fill the machine: one block per robot,
// the robot's problem staged once, no ticket) and keeps their costs in shared memory.  The softmin-weighted parameter
// sum of the update, sum_i w_i (best + eps_i), visits the LIVE samples only -- at temperature 1 with costs of
// 1e3..1e5 all but a handful of weights underflow to exactly 0 -- and draws their rows again from the counter
// (thread (s, b) = noise block b of the live samples s, s + S, ...; slice sums added in slice order: deterministic).
// Compared with k_rollout<zero_order>: no 74 KB of parameter columns per 128 rollouts (occupancy no longer bound by
// shared memory), no 144 shared-memory stores and loads per rollout, half the issue slots for the FP32 work.
// Shared memory (floats): best_t[H][12] | sigma_t[H][12] (cem) | state[24] | ref[12] | st[H][8] | ref_fz[H] (padded to
//   4) | cost[chunk] | w[chunk] | live[chunk] (u16) | red[64] | acc[S][P]; the fused finalize overlays cost.. onwards.
// MODE 0: in-kernel noise, scalar-sigma normals (mppi); 1: in-kernel noise, any method; 2: rows supplied in global
// memory (reference noise for the equivalence tests, or the parameter matrix of jit_vectorized_rollout).
constexpr int PAIR_T = 128;
constexpr float PAIR_LIVE_CUTOFF = 24.0f;  // cost distance from the minimum beyond which a sample's weight is dropped
__host__ __device__ inline int pair_fixed_floats(int P, int H, bool cem) {
    return (cem ? 2 : 1) * P + 24 + 12 + 8 * H + ((H + 3) & ~3);
}
__host__ __device__ inline int pair_smem_floats(int P, int H, int chunk, int T, bool cem) {
    const int S = T / (P / NOISE_BLOCK) > 1 ? T / (P / NOISE_BLOCK) : 1;
    return pair_fixed_floats(P, H, cem) + chunk + chunk + chunk / 2 + 64 + S * P;
}

// Block partials of `cn` rollouts base .. base + cn - 1 of robot r whose costs sit in s_cost: {min, sum w, min, argmin}
// + V[P] = sum over the live samples of w (best + eps), the rows drawn again.  All threads of the block (T = PAIR_T).
struct PairSmem {
    float *s_bt, *s_sg, *s_cost, *s_w, *s_red, *s_acc;
    unsigned short *s_live;
};
template <int MODE>
__device__ __forceinline__ void pair_block_partials(const DevCfg &c, const RolloutArgs &a, const PairSmem &sm, int r,
                                                    int base, int cn, int chunk, float *part) {
    constexpr int T = PAIR_T;
    const int P = c.P, H = c.H, t = threadIdx.x;
    float *s_bt = sm.s_bt, *s_sg = sm.s_sg, *s_cost = sm.s_cost, *s_w = sm.s_w, *s_red = sm.s_red, *s_acc = sm.s_acc;
    unsigned short *s_live = sm.s_live;
    const uint32_t robot = (uint32_t)(r + a.robot_offset);
    float bc = INFINITY;
    int bi = 0x7fffffff;
    for (int j = t; j < cn; j += T) argmin_combine(bc, bi, s_cost[j], base + j);
    block_argmin(bc, bi, s_red);
    if (c.method == QPMPC_METHOD_RANDOM) {
        if (t == 0) {
            part[0] = bc;
            part[1] = 0.f;
            part[2] = bc;
            part[3] = __int_as_float(bi);
        }
    } else {
        // softmin numerators relative to the block minimum (merged exactly by finalize_robot), :828-836.  Thread t owns
        // the contiguous costs [E t, E t + E): live samples are compacted in ascending order with one block scan.
        // LIVE = within PAIR_LIVE_CUTOFF of the block minimum, i.e. weight >= e^-24 = 3.8e-11 of the largest: the
        // rest is left out of numerator AND denominator, which moves the update by at most 2 N e^-24 max|eps| (1e-6
        // at N = 1024), below the float32 round-off of the sum itself.  Without the cutoff a converged controller
        // (cost spread of tens, not thousands) keeps hundreds of samples at weights of 1e-30 whose rows would all
        // have to be drawn again for nothing (tools/debug_nlive.py: 757 of 1024 with w > 0, 206 within the cutoff).
        const int E = (chunk + T - 1) / T, j0 = E * t, j1 = min(j0 + E, cn);
        const int lane = t & 31, wid = t >> 5;
        float sw = 0.f;
        int cnt = 0;
        for (int j = j0; j < j1; ++j) {
            const float d = s_cost[j] - bc;
            if (d <= PAIR_LIVE_CUTOFF) {
                sw += __expf(-d);
                ++cnt;
            }
        }
        int incl = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) sw += __shfl_xor_sync(0xffffffffu, sw, o);
        __syncthreads();  // s_red free again
        if (lane == 31) s_red[wid] = __int_as_float(incl);
        if (lane == 0) s_red[32 + wid] = sw;
        __syncthreads();
        int off = incl - cnt, nlive = 0;
        float swb = 0.f;
        for (int k = 0; k < T / 32; ++k) {
            const int wc = __float_as_int(s_red[k]);
            off += k < wid ? wc : 0;
            nlive += wc;
            swb += s_red[32 + k];
        }
        for (int j = j0; j < j1; ++j) {
            const float d = s_cost[j] - bc;
            if (d <= PAIR_LIVE_CUTOFF) {
                s_live[off] = (unsigned short)j;
                s_w[off] = __expf(-d);
                ++off;
            }
        }
        if (t == 0) {
            part[0] = bc;
            part[1] = swb;
            part[2] = bc;
            part[3] = __int_as_float(bi);
        }
        __syncthreads();
        // V[p] = sum over the live samples of w (best[p] + eps[p]): thread (sl, b) draws noise block b of the live
        // samples 2 sl, 2 sl + 1, then 2 (sl + S), ... again, two rows at a time through the packed generator
        const int B = P / NOISE_BLOCK, S = max(1, T / B);
        for (int u = t; u < S * B; u += T) {
            const int sl = u / B, b = u - sl * B;
            float acc[NOISE_BLOCK];
            if (MODE == 2) {
#pragma unroll
                for (int k = 0; k < NOISE_BLOCK; ++k) acc[k] = 0.f;
                for (int l = sl; l < nlive; l += S) {
                    const float *src = a.eps + ((size_t)r * a.n + base + s_live[l]) * P;
                    const float w = s_w[l];
#pragma unroll
                    for (int k = 0; k < NOISE_BLOCK; ++k)
                        acc[k] = fmaf(w, __fadd_rn(src[k * H + b], s_bt[NOISE_BLOCK * b + k]), acc[k]);
                }
            } else {
                v2 acc2[NOISE_BLOCK];
                float bst[NOISE_BLOCK];
#pragma unroll
                for (int k = 0; k < NOISE_BLOCK; ++k) {
                    acc2[k] = f2(0.f);
                    bst[k] = s_bt[NOISE_BLOCK * b + k];
                }
                for (int l = 2 * sl; l < nlive; l += 2 * S) {
                    const bool two = l + 1 < nlive;
                    const int ia = base + s_live[l], ib = base + s_live[two ? l + 1 : l];
                    const v2 w2 = f2(s_w[l], two ? s_w[l + 1] : 0.f);
                    v2 par[NOISE_BLOCK];
                    noise_block_pair<MODE == 0>(row_noise(c, ia), row_noise(c, ib), a.key, b, robot, bst, par,
                                                [&](int k) { return s_sg[NOISE_BLOCK * b + k]; });
#pragma unroll
                    for (int k = 0; k < NOISE_BLOCK; ++k) acc2[k] = vfma(par[k], w2, acc2[k]);
                }
#pragma unroll
                for (int k = 0; k < NOISE_BLOCK; ++k) acc[k] = acc2[k].x + acc2[k].y;
            }
#pragma unroll
            for (int k = 0; k < NOISE_BLOCK; ++k) s_acc[sl * P + NOISE_BLOCK * b + k] = acc[k];
        }
        __syncthreads();
        for (int o = t; o < P; o += T) {
            float v = 0.f;
            for (int sl = 0; sl < S; ++sl) v += s_acc[sl * P + o];
            part[PART_HDR + param_of(c, o / NOISE_BLOCK, o % NOISE_BLOCK)] = v;
        }
    }
}

__host__ __device__ inline int pair_update_scratch_floats(int P, int N, int T) {
    const int S = T / (P / NOISE_BLOCK) > 1 ? T / (P / NOISE_BLOCK) : 1;
    return N + N + N / 2 + 1 + 64 + S * P;
}
__host__ __device__ inline int pair_update_smem_floats(int P, int N, int T, bool cem) {
    return (cem ? 2 : 1) * P + pair_update_scratch_floats(P, N, T);
}

template <int MODE, int T, bool QSPARSE>
__global__ void __launch_bounds__(T, 3) k_rollout_pair(const __grid_constant__ DevCfg c, const RolloutArgs a_,
                                                       const __grid_constant__ InlineIn in) {
    const RolloutArgs a = resolve_inputs(a_, in);
    extern __shared__ __align__(16) float smem[];
    const int P = c.P, H = c.H, t = threadIdx.x;
    const bool cem = c.method == QPMPC_METHOD_CEM;
    float *s_bt = smem;
    float *s_sg = s_bt + P;
    float *s_state = s_sg + (cem ? P : 0);
    float *s_ref = s_state + 24;
    float *s_st = s_ref + 12;
    float *s_reffz = s_st + 8 * H;
    float *s_cost = s_reffz + ((H + 3) & ~3);
    const int chunk = a.chunk;
    float *s_w = s_cost + chunk;
    unsigned short *s_live = reinterpret_cast<unsigned short *>(s_w + chunk);
    float *s_red = s_w + chunk + chunk / 2;
    float *s_acc = s_red + 64;

    const int r = blockIdx.y;
    const int base = blockIdx.x * chunk;
    const int cn = min(chunk, a.n - base);  // rollouts of this block
    const uint32_t robot = (uint32_t)(r + a.robot_offset);

    // ---- stage the robot's problem: best / sigma transposed to step-major order (value k of step n at [12 n + k]),
    // state, reference, contact tables.  All global loads of a thread are issued before its first store.
    {
        const int role = t >> 5, tl = t & 31;
        const bool j_state = role == 1 && tl < 24, j_ref = role == 2 && tl < 12, j_tab = role == 3 && tl < H;
        const float *cs = a.contact + (size_t)r * 4 * H;
        float stv = 0.f, rfv = 0.f, s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
        if (j_state) stv = a.state[r * 24 + tl];
        if (j_ref) rfv = a.reference[r * 24 + tl];
        if (j_tab) { s0 = cs[tl]; s1 = cs[H + tl]; s2 = cs[2 * H + tl]; s3 = cs[3 * H + tl]; }
        for (int p = t; p < P; p += T) {
            const int k = p / H, n = p - k * H;
            s_bt[NOISE_BLOCK * n + k] = a.costs_only ? 0.f : a.best[(size_t)r * P + p];
            if (cem) s_sg[NOISE_BLOCK * n + k] = a.sigma ? a.sigma[(size_t)r * P + p] : 0.f;
        }
        if (j_state) s_state[tl] = stv;
        if (j_ref) s_ref[tl] = rfv;
        if (j_tab) {
            const int n = tl;
            s_reffz[n] = c.mg / (((s0 + s1) + s2) + s3);
            s_st[8 * n + 0] = s0; s_st[8 * n + 1] = s1; s_st[8 * n + 2] = s2; s_st[8 * n + 3] = s3;
            s_st[8 * n + 4] = s0 * (1.0f / 3.0f); s_st[8 * n + 5] = s1 * (1.0f / 3.0f);
            s_st[8 * n + 6] = s2 * (1.0f / 3.0f); s_st[8 * n + 7] = s3 * (1.0f / 3.0f);
        }
    }
    __syncthreads();

    float feet[12], ref[12];
#pragma unroll
    for (int k = 0; k < 12; ++k) {
        feet[k] = s_state[12 + k];
        ref[k] = s_ref[k];
    }

    // ---- rollouts: pass q integrates rollouts base + 2 T q + t (A) and base + 2 T q + T + t (B)
    for (int q0 = 0; q0 < cn; q0 += 2 * T) {
        const int ja = q0 + t, jb = q0 + T + t;
        const bool va = ja < cn, vb = jb < cn;
        const int ia = base + ja, ib = base + jb;
        v2 x[12];
#pragma unroll
        for (int k = 0; k < 12; ++k) x[k] = f2(s_state[k]);
        v2 cost = f2(0.f);
        const RowNoise rna = row_noise(c, va ? ia : 0), rnb = row_noise(c, vb ? ib : 0);
        const float *ea = nullptr, *eb = nullptr;
        if (MODE == 2) {
            ea = a.eps + ((size_t)r * a.n + (va ? ia : base)) * P;
            eb = a.eps + ((size_t)r * a.n + (vb ? ib : base)) * P;
        }
        BlockRng carry[2] = {{0u, 0u}, {0u, 0u}};  // seeds of the odd block of the current pair (noise.cuh)
#pragma unroll 1
        for (int n = 0; n < H; ++n) {
            v2 f[NOISE_BLOCK];
            const float4 b0 = *reinterpret_cast<const float4 *>(s_bt + NOISE_BLOCK * n);
            const float4 b1 = *reinterpret_cast<const float4 *>(s_bt + NOISE_BLOCK * n + 4);
            const float4 b2 = *reinterpret_cast<const float4 *>(s_bt + NOISE_BLOCK * n + 8);
            const float bst[NOISE_BLOCK] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w, b2.x, b2.y, b2.z, b2.w};
            if (MODE == 2) {
#pragma unroll
                for (int k = 0; k < NOISE_BLOCK; ++k) f[k] = vadd(f2(ea[k * H + n], eb[k * H + n]), bst[k]);
            } else {
                noise_block_pair<MODE == 0>(rna, rnb, a.key, n, robot, bst, f, [&](int k) { return s_sg[NOISE_BLOCK * n + k]; }, carry);
            }
            const float4 st4 = *reinterpret_cast<const float4 *>(s_st + 8 * n);
            const float4 st34 = *reinterpret_cast<const float4 *>(s_st + 8 * n + 4);
            const float st[4] = {st4.x, st4.y, st4.z, st4.w}, st3[4] = {st34.x, st34.y, st34.z, st34.w};
            pair_step<QSPARSE>(c, x, feet, ref, f, st, st3, s_reffz[n], c.dts[n], cost);
        }
        if (!a.costs_only) {  // centroidal_nmpc_jax.py:686-687
            cost.x = isnan(cost.x) ? 1000000.0f : cost.x;
            cost.x = isinf(cost.x) ? 1000000.0f : cost.x;
            cost.y = isnan(cost.y) ? 1000000.0f : cost.y;
            cost.y = isinf(cost.y) ? 1000000.0f : cost.y;
        }
        if (va) {
            a.costs[(size_t)r * a.n + ia] = cost.x;
            s_cost[ja] = cost.x;
        }
        if (vb) {
            a.costs[(size_t)r * a.n + ib] = cost.y;
            s_cost[jb] = cost.y;
        }
    }
    if (a.costs_only || !a.fused) return;  // not fused: k_pair_update (one block per robot) does the rest from the costs
    __syncthreads();
    const PairSmem sm = {s_bt, s_sg, s_cost, s_w, s_red, s_acc, s_live};
    pair_block_partials<MODE>(c, a, sm, r, base, cn, chunk, a.partials + ((size_t)r * a.nblk + blockIdx.x) * (PART_HDR + P));
    if (a.nblk > 1) {
        if (!last_block_of_robot(a.tickets + r, a.nblk)) return;
    } else {
        __syncthreads();  // the block's own partial, written above, is read back through L2 by finalize_robot
    }
    finalize_robot(c, resolve_inputs(a_, in), r, s_cost);
}

// Second launch of the throughput regime (launches many waves deep): ONE block per robot turns the robot's N costs into
// the update -- block partials over all N rollouts, then finalize_robot with a single partial.  Splitting this off the
// rollout kernel keeps three blocks per SM in the step loop all the time: with the fused epilogue a block spent ~30 %
// of its residency between barriers (argmin, scan, live rows, merge, prediction) while holding a third of the SM's
// warps (profiles/r5c_ncu_source_c5.txt: 31 % of the stall samples on 16 % of the instructions, outside the loop).
template <int MODE>
__global__ void __launch_bounds__(PAIR_T) k_pair_update(const __grid_constant__ DevCfg c, const RolloutArgs a_,
                                                        const __grid_constant__ InlineIn in) {
    RolloutArgs a = resolve_inputs(a_, in);
    a.nblk = 1;  // this block is the robot's only partial
    extern __shared__ __align__(16) float smem[];
    const int P = c.P, H = c.H, t = threadIdx.x, r = blockIdx.x, N = a.n;
    const bool cem = c.method == QPMPC_METHOD_CEM;
    float *s_bt = smem;
    float *s_sg = s_bt + P;
    float *s_cost = s_sg + (cem ? P : 0);  // finalize_robot's scratch overlays everything from here on
    float *s_w = s_cost + N;
    unsigned short *s_live = reinterpret_cast<unsigned short *>(s_w + N);
    float *s_red = s_w + N + N / 2 + 1;
    float *s_acc = s_red + 64;
    for (int p = t; p < P; p += PAIR_T) {
        const int k = p / H, n = p - k * H;
        s_bt[NOISE_BLOCK * n + k] = a.best[(size_t)r * P + p];
        if (cem) s_sg[NOISE_BLOCK * n + k] = a.sigma ? a.sigma[(size_t)r * P + p] : 0.f;
    }
    const float *costs = a.costs + (size_t)r * N;
    for (int j = t; j < N; j += PAIR_T) s_cost[j] = __ldcg(costs + j);
    __syncthreads();
    const PairSmem sm = {s_bt, s_sg, s_cost, s_w, s_red, s_acc, s_live};
    pair_block_partials<MODE>(c, a, sm, r, 0, N, N, a.partials + (size_t)r * (PART_HDR + P));
    __syncthreads();  // the partial, written above, is read back through L2 by finalize_robot
    a = resolve_inputs(a_, in);
    a.nblk = 1;
    finalize_robot(c, a, r, s_cost);
}

// ---------------------------------------------------------------- quad rollout kernel (latency regime) ------
// One solve of ONE robot is 10,000 rollouts = 313 warps for 592 warp schedulers: with a thread per rollout most of
// the machine idles and every warp runs a ~9,000-instruction dependent chain.  Here FOUR lanes share a rollout, one
// per leg: a lane generates a quarter of the rollout's noise, evaluates its own leg's spline, force limits, lever arm
// and torque, the quad sums the six force/torque components with two shuffle stages, the three sin/cos pairs are
// computed by three different lanes and exchanged, and each lane accumulates a quarter of the stage cost.  4x the
// warps, a ~3x shorter chain per warp, and a looped (not unrolled) body that stays in the instruction cache on a
// first and only wave.  Runtime horizon / spline count: this is also the kernel for any non-default horizon when
// the launch is small.
// Shared memory (floats): w[QB] | par[P][QB+1] | best[P] | sigma[P] | state[24] | ref[12] | timing[4] | stance[4H] |
//   ref_fz[H] | q[H] | index[H] | mgk[8] | cost[QB] | red[64]
// QB = rollouts per block (4 QB threads): 128 for a full-size solve (79 partials to merge at N = 10,000), 32 for
// small N (more blocks, more SMs).
__host__ __device__ inline int quad_smem_floats(int P, int H, int QB) {
    return QB + P * (QB + 1) + 2 * P + 24 + 12 + 4 + 4 * H + 3 * H + 8 + QB + 64;
}

template <int PARAM, bool GA, int QB>
__global__ void __launch_bounds__(4 * QB) k_rollout_quad(const __grid_constant__ DevCfg c, const RolloutArgs a_,
                                                       const __grid_constant__ InlineIn in) {
    const RolloutArgs a = resolve_inputs(a_, in);
    extern __shared__ __align__(16) float smem[];
#ifdef QPMPC_TIMING
    unsigned long long g0_;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g0_));
#endif
    constexpr int T = 4 * QB, stride = QB + 1;
    const int P = c.P, H = c.H, t = threadIdx.x;
    const int leg = t & 3, rb = t >> 2, lane = t & 31, qbase = lane & ~3;
    float *s_w = smem;
    float *s_par = s_w + QB;
    float *s_best = s_par + P * stride;
    float *s_sigma = s_best + P;
    float *s_state = s_sigma + P;
    float *s_ref = s_state + 24;
    float *s_timing = s_ref + 12;
    float *s_stance = s_timing + 4;
    float *s_reffz = s_stance + 4 * H;
    float *s_q = s_reffz + H;
    int *s_index = reinterpret_cast<int *>(s_q + H);
    float *s_mgk = s_q + 2 * H;
    float *s_cost = s_mgk + 8;

    const int r = blockIdx.y;
    const int i = blockIdx.x * QB + rb;
    const bool valid = i < a.n;

    // ---- stage the per-robot problem
    for (int p = t; p < P; p += T) {
        s_best[p] = a.costs_only ? 0.f : a.best[(size_t)r * P + p];
        s_sigma[p] = (c.method == QPMPC_METHOD_CEM && a.sigma) ? a.sigma[(size_t)r * P + p] : 0.f;
    }
    if (t < 24) s_state[t] = a.state[r * 24 + t];
    if (t >= 32 && t < 44) s_ref[t - 32] = a.reference[r * 24 + t - 32];
    if (!GA) {
        const float *cs = a.contact + (size_t)r * 4 * H;
        for (int k = t; k < 4 * H; k += T) s_stance[k] = cs[k];
        for (int n = t; n < H; n += T) {
            s_reffz[n] = c.mg / (((cs[n] + cs[H + n]) + cs[2 * H + n]) + cs[3 * H + n]);
            const int idx = chunk_index(c, (float)n);
            s_index[n] = idx;
            s_q[n] = chunk_tau(c, (float)n, (float)H, idx);
        }
    } else {
        if (t >= 64 && t < 69) s_mgk[t - 64] = c.mg / (float)(t - 64);
        if (t >= 96 && t < 100) s_timing[t - 96] = a.timing[r * 4 + t - 96];
    }
    __syncthreads();
#ifdef QPMPC_TIMING
    unsigned long long gs_, gn_, gl_;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gs_));
#endif

    // ---- parameters of the block's 32 rollouts -> columns of s_par; lane `leg` draws groups leg, leg + 4, ...
    const uint32_t robot = (uint32_t)(r + a.robot_offset);
    if (a.eps) {
        const float *src = a.eps + ((size_t)r * a.n + (size_t)blockIdx.x * QB) * P;
        const int rows = min(QB, a.n - blockIdx.x * QB);
        for (int e = t; e < QB * P; e += T) {
            const int row = e / P, p = e - row * P;
            s_par[p * stride + row] = row < rows ? s_best[p] + src[e] : 0.f;
        }
        __syncthreads();
    } else {
        // lane `leg` of the quad draws noise blocks leg, leg + 4, ... of the rollout's row
        const RowNoise rn = row_noise(c, valid ? i : 0);
        for (int b = leg; b < P / NOISE_BLOCK; b += 4) {
            float e[NOISE_BLOCK];
            if (valid) {
                noise_block(rn, a.key, b, robot, e, [&](int k) { return s_sigma[param_of(c, b, k)]; });
            } else {
#pragma unroll
                for (int k = 0; k < NOISE_BLOCK; ++k) e[k] = 0.f;
            }
#pragma unroll
            for (int k = 0; k < NOISE_BLOCK; ++k) {
                const int p = param_of(c, b, k);
                s_par[p * stride + rb] = valid ? s_best[p] + e[k] : 0.f;
            }
        }
        __syncwarp();
    }

#ifdef QPMPC_TIMING
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gn_));
#endif
    // ---- rollout: lane = leg
    float x[12], ref[12];
#pragma unroll
    for (int k = 0; k < 12; ++k) {
        x[k] = s_state[k];
        ref[k] = s_ref[k];
    }
    const float footx = s_state[12 + 3 * leg], footy = s_state[13 + 3 * leg], footz = s_state[14 + 3 * leg];
    const float qa = c.q[3 * leg], qb = c.q[3 * leg + 1], qc = c.q[3 * leg + 2];
    const float *col = s_par + rb;
    float freq = 0.f, hl = (float)H;
    uint32_t bits = 0;
    if (GA) {
        // step frequency of this rollout, then the leg's own trot clock (periodic_gait_generator_jax.py:68-89)
        freq = valid ? (a.freqs ? a.freqs[(size_t)r * a.n + i] : freq_sample(c, a.key, i, robot, a.nominal[r], a.oswing[r]))
                     : 1.0f;
        const float inc = __fmul_rn(c.mpc_dt, freq);
        float tt = s_timing[leg];
        for (int n = 0; n < H; ++n) {
            tt = (tt >= 1.0f) ? 0.0f : tt;
            tt = __fadd_rn(tt, inc);
            bits |= (tt < c.duty ? 1u : 0u) << n;
        }
        hl = (float)(__popc(bits) + 1);  // horizon_leg = sum(contact) + 1, ..._gait_adaptive.py:345-348
    }
    // the three reference values of this lane's share of the stage cost (states 3 leg .. 3 leg + 2)
    const float ra = leg == 0 ? ref[0] : leg == 1 ? ref[3] : leg == 2 ? ref[6] : ref[9];
    const float rb_ = leg == 0 ? ref[1] : leg == 1 ? ref[4] : leg == 2 ? ref[7] : ref[10];
    const float rc = leg == 0 ? ref[2] : leg == 1 ? ref[5] : leg == 2 ? ref[8] : ref[11];
    const float rq = __fdividef((float)c.S, hl);  // position in chunk = step / (horizon_leg / S) - index
    const float b1 = c.bounds[1], b2 = c.bounds[2];
    const int hs = H * stride;
    float cost = 0.f;
    float kc[3][4];
    int ncnt = -1, cur = -1;
#pragma unroll 1
    for (int n = 0; n < H; ++n) {
        float st, ref_fz;
        int istep;
        float4 B = make_float4(0.f, 0.f, 0.f, 0.f);
        if (GA) {
            const int on = (bits >> n) & 1;
            const unsigned bal = __ballot_sync(0xffffffffu, on);
            st = (float)on;
            ref_fz = s_mgk[__popc((bal >> qbase) & 0xFu)];
            ncnt += on;  // the leg's own step counter advances only in stance, :353-356
            istep = max(ncnt, 0);  // a swing leg's value is multiplied by 0: any finite stand-in does
            if (PARAM != QPMPC_PARAM_ZERO_ORDER) {
                const float stepf = (float)istep;
                const int idx = chunk_of(c, stepf, b1, b2);
                if (idx != cur) {
                    load_chunk<PARAM, stride>(c, col, leg, idx, kc);
                    cur = idx;
                }
                B = chunk_basis<PARAM>(stepf * rq - (float)idx);
            }
        } else {
            st = s_stance[leg * H + n];
            ref_fz = s_reffz[n];
            istep = n;
            if (PARAM != QPMPC_PARAM_ZERO_ORDER) {
                const int idx = s_index[n];
                if (idx != cur) {
                    load_chunk<PARAM, stride>(c, col, leg, idx, kc);
                    cur = idx;
                }
                B = chunk_basis<PARAM>(s_q[n]);
            }
        }
        float f[3];
        if (PARAM == QPMPC_PARAM_ZERO_ORDER) {
            const float *p = col + leg * 3 * hs + istep * stride;
            f[0] = p[0];
            f[1] = p[hs];
            f[2] = p[2 * hs];
        } else {
            eval_chunk<PARAM>(kc, B, f);
        }
        quad_leg_step(c, f, st, ref_fz, footx, footy, footz, x, c.dts[n], leg, qbase);
        // stage cost: lane `leg` owns the terms of states 3 leg .. 3 leg + 2
        const float ea = (leg == 0 ? x[0] : leg == 1 ? x[3] : leg == 2 ? x[6] : x[9]) - ra;
        const float eb = (leg == 0 ? x[1] : leg == 1 ? x[4] : leg == 2 ? x[7] : x[10]) - rb_;
        const float ec = (leg == 0 ? x[2] : leg == 1 ? x[5] : leg == 2 ? x[8] : x[11]) - rc;
        cost += (ea * qa) * ea;
        cost += (eb * qb) * eb;
        cost += (ec * qc) * ec;
    }
#ifdef QPMPC_TIMING
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gl_));
#endif
    cost = quad_sum(cost);
    if (GA) {
        const float df = freq - 1.3f;
        cost += df * 100.f * df;  // ..._gait_adaptive.py:500
    }
    if (!a.costs_only) {  // centroidal_nmpc_jax.py:686-687
        cost = isnan(cost) ? 1000000.0f : cost;
        cost = isinf(cost) ? 1000000.0f : cost;
    }
    if (leg == 0) {
        if (valid) a.costs[(size_t)r * a.n + i] = cost;
        s_cost[rb] = valid ? cost : INFINITY;
    }
    if (a.costs_only) return;
    __syncthreads();

    // ---- block partials: thread t < QB owns the cost of rollout t of the block
    float *part = a.partials + ((size_t)r * a.nblk + blockIdx.x) * (PART_HDR + P);
    {
        float *s_red = s_cost + QB;
        const int it = blockIdx.x * QB + t;
        const bool own = t < QB && it < a.n;
        const float ct = own ? s_cost[t] : INFINITY;
        float bc = ct;
        int bi = own ? it : 0x7fffffff;
        block_argmin(bc, bi, s_red);
        const float w = own ? __expf(-(ct - bc)) : 0.f;
        if (t < QB) s_w[t] = w;
        const float sw = block_sum(w, s_red);
        if (t == 0) {
            part[0] = bc;
            part[1] = c.method == QPMPC_METHOD_RANDOM ? 0.f : sw;
            part[2] = bc;
            part[3] = __int_as_float(bi);
        }
    }
    if (c.method != QPMPC_METHOD_RANDOM) {
        __syncthreads();
        for (int p = t; p < P; p += T) {
            const float *row = s_par + p * stride;
            float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
            for (int k = 0; k < QB; k += 4) {
                const float4 w4 = *reinterpret_cast<const float4 *>(s_w + k);
                a0 = fmaf(w4.x, row[k], a0);
                a1 = fmaf(w4.y, row[k + 1], a1);
                a2 = fmaf(w4.z, row[k + 2], a2);
                a3 = fmaf(w4.w, row[k + 3], a3);
            }
            part[PART_HDR + p] = (a0 + a1) + (a2 + a3);
        }
    }
#ifdef QPMPC_TIMING
    unsigned long long g1_, g2_, g3_;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g1_));
    if (a.fused && last_block_of_robot(a.tickets + r, a.nblk)) {
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g2_));
        finalize_robot(c, resolve_inputs(a_, in), r, s_par);
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g3_));
        if (t == 0) {
            a.result[r].footholds[6] = (float)(g1_ - g0_);
            a.result[r].footholds[7] = (float)(g2_ - g1_);
            a.result[r].footholds[8] = (float)(g3_ - g2_);
            a.result[r].footholds[9] = (float)(gs_ - g0_);
            a.result[r].footholds[10] = (float)(gn_ - gs_);
            a.result[r].footholds[11] = (float)(gl_ - gn_);
        }
    }
#else
    if (a.fused && last_block_of_robot(a.tickets + r, a.nblk)) finalize_robot(c, resolve_inputs(a_, in), r, s_par);
#endif
}

// ---------------------------------------------------------------- noise dump -------------------------------
__global__ void k_dump_noise(const __grid_constant__ DevCfg c, NoiseKey key, uint32_t robot, const float *sigma,
                             float nominal, int oswing, float *eps_out, float *freqs_out) {
    const int B = c.P / NOISE_BLOCK;
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx < (long long)c.N * B) {
        const int i = (int)(idx / B), b = (int)(idx % B);
        float e[NOISE_BLOCK];
        noise_block(row_noise(c, i), key, b, robot, e, [&](int k) { return sigma[param_of(c, b, k)]; });
#pragma unroll
        for (int k = 0; k < NOISE_BLOCK; ++k) eps_out[(size_t)i * c.P + param_of(c, b, k)] = e[k];
    }
    if (freqs_out && c.ga && idx < c.N) freqs_out[idx] = freq_sample(c, key, (int)idx, robot, nominal, oswing);
}

// The same dump through the PACKED generator of k_rollout_pair (rows i and i + half of a thread as its halves A and
// B): the test hook behind "the two generators are bit-identical" (QPMPC_DUMP=pair, tests/test_sampling_parity_gpu.py).
__global__ void k_dump_noise_pair(const __grid_constant__ DevCfg c, NoiseKey key, uint32_t robot, const float *sigma,
                                  float *eps_out) {
    const int B = c.P / NOISE_BLOCK, half = (c.N + 1) / 2;
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)half * B) return;
    const int ia = (int)(idx / B), b = (int)(idx % B), ib = ia + half;
    const bool vb = ib < c.N;
    v2 e[NOISE_BLOCK];
    auto sig = [&](int k) { return sigma ? sigma[param_of(c, b, k)] : 1.0f; };
    if (c.method == QPMPC_METHOD_MPPI) noise_block_pair_eps<true>(row_noise(c, ia), row_noise(c, vb ? ib : 0), key, b, robot, e, sig);
    else noise_block_pair_eps<false>(row_noise(c, ia), row_noise(c, vb ? ib : 0), key, b, robot, e, sig);
#pragma unroll
    for (int k = 0; k < NOISE_BLOCK; ++k) {
        eps_out[(size_t)ia * c.P + param_of(c, b, k)] = e[k].x;
        if (vb) eps_out[(size_t)ib * c.P + param_of(c, b, k)] = e[k].y;
    }
}

// ---------------------------------------------------------------- gait clock --------------
// One thread per (robot, leg).  periodic_gait_generator.py:47-76: phase += dt * f, wrap to [0, 1), a leg in the init
// hold stays in contact until its phase passes its offset (then restarts at 0), otherwise contact = phase < duty.
// Separately rounded product and sum (Python floats), exact fmod.
__device__ __forceinline__ float gait_run(double &phase, bool &init, double offset, double duty, double dt, double f) {
    phase = __dadd_rn(phase, __dmul_rn(dt, f));
    phase = fmod(phase, 1.0);
    if (phase < 0.0) phase = __dadd_rn(phase, 1.0);  // Python's % takes the divisor's sign
    if (init) {
        if (!(phase <= offset)) {
            init = false;
            phase = 0.0;
        }
        return 1.0f;
    }
    return phase < duty ? 1.0f : 0.0f;
}

// steps > 0: contact sequence of `steps` columns, clock state NOT stored; steps == 0: one run(dt), state stored
__global__ void k_gait(const __grid_constant__ qpmpc_gait_batch g, int steps, double dt) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= g.num_robots * 4) return;
    const int r = idx >> 2;
    double phase = g.d_phase[idx];
    bool init = g.d_init[idx] != 0;
    const double offset = g.d_phase_offset[idx], duty = g.d_duty_factor[r], f = g.d_step_freq[r];
    if (steps == 0) {
        g.d_contact[idx] = gait_run(phase, init, offset, duty, dt, f);
        g.d_phase[idx] = phase;
        g.d_init[idx] = init ? 1 : 0;
        return;
    }
    const bool full = g.d_full_stance != nullptr && g.d_full_stance[r] != 0;
    float *out = g.d_contact + (size_t)idx * steps;
    for (int n = 0; n < steps; ++n) {
        const float c = gait_run(phase, init, offset, duty, n == 0 ? 0.0 : g.step_dts[n], f);
        out[n] = full ? 1.0f : c;
    }
}

}  // namespace qp

// ================================================================ host side ================================
using namespace qp;

static thread_local std::string g_err;
static int fail(int code, const std::string &msg) {
    g_err = msg;
    return code;
}
#define CU(call)                                                                                         \
    do {                                                                                                 \
        cudaError_t e_ = (call);                                                                         \
        if (e_ != cudaSuccess)                                                                           \
            return fail(QPMPC_E_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_));               \
    } while (0)

typedef void (*rollout_fn)(const qp::DevCfg, const qp::RolloutArgs, const qp::InlineIn);

// Experiment switches, read from the environment ONCE per handle (qpmpc_create): five getenv calls per solve were
// measurable on a 30 us call.  QPMPC_KERNEL=quad|thread, QPMPC_QB=32|72|128, QPMPC_FUSED=0, QPMPC_ZEROCOPY=0|1,
// QPMPC_INPUTS=copy|mapped, QPMPC_PAIR=0.
struct EnvSwitches {
    int kernel;    // 0 auto, 1 quad, 2 thread
    int qb;        // 0 auto, else 32 / 72 / 128
    int fused;     // -1 auto (fused unless the launch is many waves deep), 0: stand-alone k_finalize, 1: fused
    int zerocopy;  // -1 unset, 0 classic copies, 1 inputs mapped too
    int inputs;    // 0 auto (parameter block), 1 copy, 2 mapped
    int threads;   // 0 auto (128 rollouts per block of the thread-per-rollout kernel), 64 (QPMPC_T=64)
    int pair;      // 0: QPMPC_PAIR=0 keeps the thread-per-rollout kernel where the packed pair kernel would run
    int dump_pair; // 1: QPMPC_DUMP=pair makes qpmpc_generate_noise use the packed generator (test hook)
};
static EnvSwitches read_env() {
    EnvSwitches e = {0, 0, -1, -1, 0, 0, 1, 0};
    if (const char *k = getenv("QPMPC_KERNEL")) e.kernel = !strcmp(k, "quad") ? 1 : !strcmp(k, "thread") ? 2 : 0;
    if (const char *q = getenv("QPMPC_QB")) e.qb = atoi(q) == 128 ? 128 : atoi(q) == 72 ? 72 : 32;
    if (const char *f = getenv("QPMPC_FUSED")) e.fused = f[0] == '0' ? 0 : 1;
    if (const char *z = getenv("QPMPC_ZEROCOPY")) e.zerocopy = z[0] == '0' ? 0 : z[0] == '1' ? 1 : -1;
    if (const char *i = getenv("QPMPC_INPUTS")) e.inputs = !strcmp(i, "copy") ? 1 : !strcmp(i, "mapped") ? 2 : 0;
    if (const char *q = getenv("QPMPC_T")) e.threads = atoi(q) == 64 ? 64 : 0;
    if (const char *q = getenv("QPMPC_PAIR")) e.pair = q[0] == '0' ? 0 : 1;
    if (const char *q = getenv("QPMPC_DUMP")) e.dump_pair = !strcmp(q, "pair") ? 1 : 0;
    return e;
}

struct qpmpc_handle {
    qpmpc_config cfg;
    DevCfg dc;
    int device;
    cudaStream_t stream;
    int num_sms;
    rollout_fn fn128, fn64, fn32, fn_quad32, fn_quad72, fn_quad128, fn_pair[3], fn_pair_update[3];
    EnvSwitches env;
    size_t max_smem;
    // single-robot staging: one pinned input block, one pinned output block
    float *h_in, *d_in;
    float *h_out, *d_out;
    float *d_iter;        // [2][2P]: mean | sigma carried between chained iterations (qpmpc_solve_iterated)
    float *m_io, *md_io;  // mapped (zero-copy) host block of qpmpc_solve: inputs | {word, seq} output pairs
    unsigned int seq;     // value the next solve stores to the done flag
    size_t in_floats, out_floats;
    float *d_costs;     // [max_robots * N]
    float *d_partials;  // [max_robots * nblk_max * (4 + P)]
    unsigned int *d_tickets;  // [max_robots], zero between launches
    size_t partial_floats;
    int64_t launches;
    const char *last_kernel;       // rollout kernel of the most recent launch: "pair", "quad" or "thread"
    int profiling;                  // qpmpc_set_profiling: record events around each kernel of a solve
    cudaEvent_t ev[3];
};

static int num_params_of(const qpmpc_config &c) {
    if (c.parametrization == QPMPC_PARAM_LINEAR) return (c.num_splines + 1) * 3 * 4;
    if (c.parametrization == QPMPC_PARAM_CUBIC) return 4 * 3 * c.num_splines * 4;
    return c.horizon * 3 * 4;
}

// centroidal_model_jax.py:67-91 evaluated in float32 (same cofactor order as the reference's inertia inverse)
static void inverse3(const float A[9], float out[9]) {
    const float a11 = A[0], a12 = A[1], a13 = A[2], a21 = A[3], a22 = A[4], a23 = A[5], a31 = A[6], a32 = A[7],
                a33 = A[8];
    const float det = a11 * (a33 * a22 - a32 * a23) - a21 * (a33 * a12 - a32 * a13) + a31 * (a23 * a12 - a22 * a13);
    out[0] = (a33 * a22 - a32 * a23) / det;
    out[1] = -(a33 * a12 - a32 * a13) / det;
    out[2] = (a23 * a12 - a22 * a13) / det;
    out[3] = -(a33 * a21 - a31 * a23) / det;
    out[4] = (a33 * a11 - a31 * a13) / det;
    out[5] = -(a23 * a11 - a21 * a13) / det;
    out[6] = (a32 * a21 - a31 * a22) / det;
    out[7] = -(a32 * a11 - a31 * a12) / det;
    out[8] = (a22 * a11 - a21 * a12) / det;
}

// thread-per-rollout kernels: 128-thread blocks, or 32-thread blocks when P is too large for 128 columns of shared memory
template <int T>
static rollout_fn pick_rollout(int param, int ga) {
    switch (param * 2 + (ga ? 1 : 0)) {
        case 0: return k_rollout<QPMPC_PARAM_ZERO_ORDER, false, T>;
        case 1: return k_rollout<QPMPC_PARAM_ZERO_ORDER, true, T>;
        case 2: return k_rollout<QPMPC_PARAM_LINEAR, false, T>;
        case 3: return k_rollout<QPMPC_PARAM_LINEAR, true, T>;
        case 4: return k_rollout<QPMPC_PARAM_CUBIC, false, T>;
        default: return k_rollout<QPMPC_PARAM_CUBIC, true, T>;
    }
}
template <int QB>
static rollout_fn pick_quad(int param, int ga) {
    switch (param * 2 + (ga ? 1 : 0)) {
        case 0: return k_rollout_quad<QPMPC_PARAM_ZERO_ORDER, false, QB>;
        case 1: return k_rollout_quad<QPMPC_PARAM_ZERO_ORDER, true, QB>;
        case 2: return k_rollout_quad<QPMPC_PARAM_LINEAR, false, QB>;
        case 3: return k_rollout_quad<QPMPC_PARAM_LINEAR, true, QB>;
        case 4: return k_rollout_quad<QPMPC_PARAM_CUBIC, false, QB>;
        default: return k_rollout_quad<QPMPC_PARAM_CUBIC, true, QB>;
    }
}
struct LaunchPlan {
    bool quad;    // k_rollout_quad (4 lanes per rollout) instead of k_rollout (1 thread per rollout)
    bool pair;    // k_rollout_pair (2 rollouts per thread, packed FP32; zero-order fixed gait, throughput regime)
    int T, nblk;  // threads per block, blocks per robot
    int chunk;    // rollouts per block
    size_t smem;
};
// Latency regime (fewer than two rollouts-as-threads warps per SM sub-partition, e.g. one robot x 10,000 rollouts):
// the quad kernel.  Throughput regime (batched robots): one thread per rollout, 128-thread blocks.
static LaunchPlan plan_launch(const qpmpc_handle *h, int n, long long total_rollouts) {
    const DevCfg &d = h->dc;
    const long long subparts = (long long)h->num_sms * 4;
    LaunchPlan pl;
    pl.pair = false;
    pl.quad = total_rollouts < subparts * 32 * 2;
    if (h->env.kernel == 1) pl.quad = true;   // QPMPC_KERNEL pins the choice (the tests run both kernels at every size)
    if (h->env.kernel == 2) pl.quad = false;
    // rollouts per quad block: 128 keeps the number of partials the last block merges small; 32 spreads a small solve
    // (72 x 4 threads = 9 warps: N = 10,000 becomes 139 blocks, one per SM in a single wave)
    int QB = n < 4096 ? 32 : ((n + 71) / 72 <= h->num_sms ? 72 : 128);
    if (h->env.qb) QB = h->env.qb;
    if (pl.quad && QB == 72 && (size_t)quad_smem_floats(d.P, d.H, 72) * 4 > h->max_smem) QB = 32;
    if (pl.quad && QB == 128 && (size_t)quad_smem_floats(d.P, d.H, 128) * 4 > h->max_smem) QB = 32;
    if (pl.quad && (size_t)quad_smem_floats(d.P, d.H, QB) * 4 > h->max_smem) pl.quad = false;
    const bool cem = d.method == QPMPC_METHOD_CEM;
    if (pl.quad) {
        pl.T = 4 * QB;
        pl.nblk = (n + QB - 1) / QB;
        pl.chunk = QB;
        pl.smem = (size_t)quad_smem_floats(d.P, d.H, QB) * 4;
    } else if (d.param == QPMPC_PARAM_ZERO_ORDER && !d.ga && h->env.pair != 0) {
        // one block per robot (all passes in one block: the problem is staged once, no ticket) as long as the robots
        // alone give every SM its three blocks; fewer passes per block, more blocks per robot, otherwise
        const long long R = total_rollouts / (n > 0 ? n : 1);
        const int per_pass = 2 * PAIR_T;
        int passes = (n + per_pass - 1) / per_pass;
        auto blocks_of = [&](int ps) { return (n + per_pass * ps - 1) / (per_pass * ps); };
        while (passes > 1 && (R * blocks_of(passes) < 3LL * h->num_sms || per_pass * passes > 32768 ||
                              (size_t)pair_smem_floats(d.P, d.H, per_pass * passes, PAIR_T, cem) * 4 > h->max_smem))
            passes = (passes + 1) / 2;
        pl.pair = true;
        pl.T = PAIR_T;
        pl.chunk = per_pass * passes;
        pl.nblk = blocks_of(passes);
        pl.smem = (size_t)pair_smem_floats(d.P, d.H, pl.chunk, PAIR_T, cem) * 4;
    } else {
        int T = (size_t)rollout_smem_floats(d.P, d.H, 128, d.method == QPMPC_METHOD_CEM) * 4 <= h->max_smem ? 128 : 32;
        if (T == 128 && h->env.threads == 64) T = 64;
        pl.T = T;
        pl.nblk = (n + T - 1) / T;
        pl.chunk = T;
        pl.smem = (size_t)rollout_smem_floats(d.P, d.H, T, d.method == QPMPC_METHOD_CEM) * 4;
    }
    return pl;
}

// eps: rows supplied in global memory (reference noise or the parameter matrix of jit_vectorized_rollout)
static rollout_fn pick_fn(qpmpc_handle *h, const LaunchPlan &pl, bool eps) {
    h->last_kernel = pl.pair ? "pair" : pl.quad ? "quad" : "thread";
    if (pl.pair) return eps ? h->fn_pair[2] : h->dc.method == QPMPC_METHOD_MPPI ? h->fn_pair[0] : h->fn_pair[1];
    if (pl.quad) return pl.T == 512 ? h->fn_quad128 : pl.T == 288 ? h->fn_quad72 : h->fn_quad32;
    return pl.T == 128 ? h->fn128 : pl.T == 64 ? h->fn64 : h->fn32;
}

extern "C" const char *qpmpc_last_error(void) { return g_err.c_str(); }
#define QPMPC_STR2(x) #x
#define QPMPC_STR(x) QPMPC_STR2(x)
extern "C" const char *qpmpc_version(void) { return "qpmpc-b200 0.2 noise-v3 (sm_100a, cudart " QPMPC_STR(CUDART_VERSION) ")"; }

extern "C" int qpmpc_create(const qpmpc_config *cfg, qpmpc_handle **out) {
    if (!cfg || !out) return fail(QPMPC_E_ARG, "null argument");
    *out = nullptr;
    const qpmpc_config &c = *cfg;
    if (c.horizon < 1 || c.horizon > QPMPC_MAX_H) return fail(QPMPC_E_ARG, "horizon out of range [1, 32]");
    if (c.parametrization < 0 || c.parametrization > 2) return fail(QPMPC_E_ARG, "unknown control_parametrization");
    if (c.sampling_method < 0 || c.sampling_method > 2) return fail(QPMPC_E_ARG, "unknown sampling_method");
    if (c.num_splines < 1 || c.num_splines > MAX_SPLINES) return fail(QPMPC_E_ARG, "num_splines out of range [1, 16]");
    if (c.num_parallel_computations < 1) return fail(QPMPC_E_ARG, "num_parallel_computations must be >= 1");
    if (c.max_robots < 1) return fail(QPMPC_E_ARG, "max_robots must be >= 1");
    const int P = num_params_of(c);
    if (P > QPMPC_MAX_P) return fail(QPMPC_E_ARG, "too many control parameters");
    if (c.gait_adaptive && (c.num_step_freq < 1 || c.num_step_freq > QPMPC_MAX_FREQ))
        return fail(QPMPC_E_ARG, "num_step_freq out of range [1, 8]");
    if (c.sampling_method == QPMPC_METHOD_CEM && c.cem_elite < 2) return fail(QPMPC_E_ARG, "cem_elite must be >= 2");

    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(QPMPC_E_NODEVICE, "no CUDA device: this library has no CPU fallback");
    }
    if (c.device < 0 || c.device >= ndev) return fail(QPMPC_E_ARG, "device ordinal out of range");
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, c.device));
    if (prop.major != 10) return fail(QPMPC_E_NODEVICE, std::string("device is ") + prop.name + " (sm_" + std::to_string(prop.major) + std::to_string(prop.minor) + "); kernels are built for sm_100a only");
    CU(cudaSetDevice(c.device));

    qpmpc_handle *h = new (std::nothrow) qpmpc_handle();
    if (!h) return fail(QPMPC_E_ARG, "out of host memory");
    h->cfg = c;
    h->env = read_env();
    h->device = c.device;
    h->launches = 0;
    h->last_kernel = "none";
    h->profiling = 0;
    DevCfg &d = h->dc;
    memset(&d, 0, sizeof d);
    d.H = c.horizon; d.param = c.parametrization; d.S = c.num_splines; d.ga = c.gait_adaptive ? 1 : 0;
    d.method = c.sampling_method; d.N = c.num_parallel_computations; d.P = P; d.np = P / 4;
    d.mass = c.mass; d.inv_mass = 1.0f / c.mass; d.mg = (float)((double)c.mass * 9.81);
    memcpy(d.I, c.inertia, sizeof d.I);
    inverse3(c.inertia, d.Iinv);
    d.mu = c.mu; d.fzmin = c.f_z_min; d.fzmax = c.f_z_max;
    memcpy(d.dts, c.dts, sizeof d.dts);
    memcpy(d.q, c.q_diag, sizeof d.q);
    d.duty = c.duty_factor; d.mpc_dt = c.mpc_dt;
    memcpy(d.sig_rs, c.sigma_random_sampling, sizeof d.sig_rs);
    d.sig_mppi = c.sigma_mppi;
    d.nfreq = c.num_step_freq;
    memcpy(d.freqs, c.step_freq_available, sizeof d.freqs);
    d.cem_elite = c.cem_elite; d.cem_sqrt = c.cem_sqrt; d.cem_lo = c.cem_lo; d.cem_hi = c.cem_hi;
    for (int k = 0; k <= d.S; ++k) d.bounds[k] = (float)((double)d.H * (double)k / (double)d.S);

    h->max_smem = prop.sharedMemPerBlockOptin;
    h->num_sms = prop.multiProcessorCount;
    if ((size_t)rollout_smem_floats(P, d.H, 32, d.method == QPMPC_METHOD_CEM) * 4 > h->max_smem) {
        delete h;
        return fail(QPMPC_E_ARG, "control parameter vector too large for shared memory");
    }
    h->fn128 = pick_rollout<128>(d.param, d.ga);
    h->fn32 = pick_rollout<32>(d.param, d.ga);
    h->fn64 = pick_rollout<64>(d.param, d.ga);
    h->fn_quad32 = pick_quad<32>(d.param, d.ga);
    h->fn_quad72 = pick_quad<72>(d.param, d.ga);
    h->fn_quad128 = pick_quad<128>(d.param, d.ga);
    // stage cost without its zero-weight terms when Q has the reference's pattern (rollout_pair.cuh: QSPARSE)
    const bool qsparse = c.q_diag[0] == 0.f && c.q_diag[1] == 0.f && c.q_diag[8] == 0.f && c.q_diag[3] > 0.f &&
                         c.q_diag[4] > 0.f && c.q_diag[6] > 0.f;
    h->fn_pair[0] = qsparse ? k_rollout_pair<0, PAIR_T, true> : k_rollout_pair<0, PAIR_T, false>;
    h->fn_pair[1] = qsparse ? k_rollout_pair<1, PAIR_T, true> : k_rollout_pair<1, PAIR_T, false>;
    h->fn_pair[2] = qsparse ? k_rollout_pair<2, PAIR_T, true> : k_rollout_pair<2, PAIR_T, false>;
    h->fn_pair_update[0] = k_pair_update<0>;
    h->fn_pair_update[1] = k_pair_update<1>;
    h->fn_pair_update[2] = k_pair_update<2>;
    cudaError_t e = cudaSuccess;
    size_t static_smem = 0;  // the kernels' own __shared__ variables come out of the same per-block budget
    for (rollout_fn f : {h->fn128, h->fn64, h->fn32, h->fn_quad32, h->fn_quad72, h->fn_quad128, h->fn_pair[0], h->fn_pair[1], h->fn_pair[2], h->fn_pair_update[0], h->fn_pair_update[1], h->fn_pair_update[2]}) {
        cudaFuncAttributes fa;
        if (e == cudaSuccess) e = cudaFuncGetAttributes(&fa, (const void *)f);
        if (e == cudaSuccess && fa.sharedSizeBytes > static_smem) static_smem = fa.sharedSizeBytes;
    }
    h->max_smem -= static_smem;
    for (rollout_fn f : {h->fn128, h->fn64, h->fn32, h->fn_quad32, h->fn_quad72, h->fn_quad128, h->fn_pair[0], h->fn_pair[1], h->fn_pair[2], h->fn_pair_update[0], h->fn_pair_update[1], h->fn_pair_update[2]})
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute((const void *)f, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->max_smem);
    if (e != cudaSuccess) { delete h; return fail(QPMPC_E_CUDA, std::string("cudaFuncSetAttribute: ") + cudaGetErrorString(e)); }

    const int H = d.H;
    h->in_floats = 24 + 24 + 4 * H + P + P + 4 + 2;  // state ref contact best sigma timing nominal oswing
    h->out_floats = sizeof(qpmpc_result) / 4 + 2 * P;
    const int nblk_max = (d.N + 31) / 32;
    h->partial_floats = (size_t)c.max_robots * nblk_max * (PART_HDR + P);
    bool ok = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking) == cudaSuccess &&
              cudaEventCreate(&h->ev[0]) == cudaSuccess && cudaEventCreate(&h->ev[1]) == cudaSuccess &&
              cudaEventCreate(&h->ev[2]) == cudaSuccess &&
              cudaMallocHost(&h->h_in, h->in_floats * 4) == cudaSuccess &&
              cudaMallocHost(&h->h_out, h->out_floats * 4 + (size_t)d.N * 4) == cudaSuccess &&
              cudaHostAlloc(&h->m_io, (h->in_floats + 2 * h->out_floats + 16) * 4, cudaHostAllocMapped) == cudaSuccess &&
              cudaHostGetDevicePointer(&h->md_io, h->m_io, 0) == cudaSuccess &&
              cudaMalloc(&h->d_in, h->in_floats * 4) == cudaSuccess &&
              cudaMalloc(&h->d_out, h->out_floats * 4) == cudaSuccess &&
              cudaMalloc(&h->d_iter, (size_t)4 * P * 4) == cudaSuccess &&
              cudaMalloc(&h->d_costs, (size_t)c.max_robots * d.N * 4) == cudaSuccess &&
              cudaMalloc(&h->d_partials, h->partial_floats * 4) == cudaSuccess &&
              cudaMalloc(&h->d_tickets, (size_t)c.max_robots * 4) == cudaSuccess &&
              cudaMemset(h->d_tickets, 0, (size_t)c.max_robots * 4) == cudaSuccess;
    if (!ok) {
        std::string m = std::string("allocation failed: ") + cudaGetErrorString(cudaGetLastError());
        qpmpc_destroy(h);
        return fail(QPMPC_E_CUDA, m);
    }
    memset(h->m_io, 0, (h->in_floats + 2 * h->out_floats + 16) * 4);  // sequence number 0 = nothing published yet
    h->seq = 0;
    *out = h;
    return QPMPC_OK;
}

extern "C" int qpmpc_destroy(qpmpc_handle *h) {
    if (!h) return QPMPC_OK;
    cudaSetDevice(h->device);
    if (h->stream) { cudaStreamSynchronize(h->stream); cudaStreamDestroy(h->stream); }
    for (int k = 0; k < 3; ++k)
        if (h->ev[k]) cudaEventDestroy(h->ev[k]);
    if (h->h_in) cudaFreeHost(h->h_in);
    if (h->h_out) cudaFreeHost(h->h_out);
    if (h->m_io) cudaFreeHost(h->m_io);
    if (h->d_in) cudaFree(h->d_in);
    if (h->d_out) cudaFree(h->d_out);
    if (h->d_iter) cudaFree(h->d_iter);
    if (h->d_costs) cudaFree(h->d_costs);
    if (h->d_partials) cudaFree(h->d_partials);
    if (h->d_tickets) cudaFree(h->d_tickets);
    delete h;
    return QPMPC_OK;
}

extern "C" int qpmpc_num_params(const qpmpc_handle *h) { return h ? h->dc.P : QPMPC_E_ARG; }
extern "C" void *qpmpc_stream(qpmpc_handle *h) { return h ? (void *)h->stream : nullptr; }
extern "C" int64_t qpmpc_launch_count(const qpmpc_handle *h) { return h ? h->launches : 0; }
extern "C" const char *qpmpc_last_kernel(const qpmpc_handle *h) { return h ? h->last_kernel : "none"; }

// launches k_rollout + k_finalize for `b` on the handle's stream (no synchronisation)
static const InlineIn g_no_inline = {};

static int launch_solve(qpmpc_handle *h, const qpmpc_batch &b, uint64_t seed, uint64_t offset,
                        uint2 *ll_out = nullptr, unsigned int ll_seq = 0,
                        const InlineIn *inl = nullptr, int inline_mask = 7) {
    const DevCfg &d = h->dc;
    const int R = b.num_robots, N = d.N, P = d.P;
    const LaunchPlan pl = plan_launch(h, N, (long long)N * R);
    const int nblk = pl.nblk;
    if ((size_t)R * nblk * (PART_HDR + P) > h->partial_floats) return fail(QPMPC_E_ARG, "internal: partial buffer too small");
    SolveArgs ra;
    memset(&ra, 0, sizeof ra);
    ra.state = b.d_state; ra.reference = b.d_reference; ra.contact = b.d_contact_sequence; ra.best = b.d_best_in;
    ra.sigma = b.d_sigma_in; ra.timing = b.d_timing; ra.nominal = b.d_nominal_freq; ra.oswing = b.d_optimize_swing;
    ra.eps = b.d_eps; ra.freqs = b.d_freqs;
    ra.costs = b.d_costs ? b.d_costs : h->d_costs; ra.partials = h->d_partials;
    ra.result = b.d_result; ra.best_out = b.d_best_out; ra.sigma_out = b.d_sigma_out; ra.tickets = h->d_tickets;
    ra.key = NoiseKey{seed, offset}; ra.n = N; ra.costs_only = 0; ra.nblk = nblk; ra.robot_offset = b.robot_offset;
    ra.ll_out = ll_out; ra.ll_seq = ll_seq;
    ra.inline_in = inl ? inline_mask : 0;  // the caller then filled those input pointers of `b` with byte offsets
    const InlineIn &in = inl ? *inl : g_no_inline;
    // One launch per solve: the last rollout block of each robot runs the finalize step in the shared memory its
    // parameter columns occupied.  Falls back to the stand-alone k_finalize when that scratch is too small
    // (huge N with tiny P) or when QPMPC_FUSED=0 asks for the two-kernel path (tests cover both).
    const int K = d.method == QPMPC_METHOD_CEM ? d.cem_elite : 0;
    const int scratch = pl.pair ? pair_smem_floats(P, d.H, pl.chunk, pl.T, d.method == QPMPC_METHOD_CEM) -
                                      pair_fixed_floats(P, d.H, d.method == QPMPC_METHOD_CEM)
                        : pl.quad ? P * (pl.T / 4 + 1) : P * (pl.T + 1);
    // Throughput regime (batched robots, many waves of blocks): the stand-alone k_finalize wins -- the arrival ticket
    // (fence + atomic round trip with the block's warps parked) and a last block that finalizes while two of its
    // SM's three slots wait cost more than one extra launch (C5: 3.09 -> 2.93 ms).  Latency regime: one launch.
    // (the pair kernel with one block per robot has no ticket and overlaps its epilogue with the other blocks' step
    // loops: fused 1.49 ms against 1.41 + 0.11 ms for rollouts + k_pair_update at C5)
    const bool deep = !pl.quad && !(pl.pair && nblk == 1) && (long long)R * nblk > 8LL * h->num_sms;
    const bool want_fused = h->env.fused < 0 ? !deep : h->env.fused != 0;
    ra.fused = finalize_smem_floats(P, nblk, K, pl.T) <= scratch && want_fused;
    // pair kernel, not fused: k_pair_update (one block per robot over all N costs) needs them in shared memory
    const bool cem_ = d.method == QPMPC_METHOD_CEM;
    if (pl.pair && !ra.fused &&
        ((size_t)pair_update_smem_floats(P, N, PAIR_T, cem_) * 4 > h->max_smem || N > 32768 ||
         finalize_smem_floats(P, 1, K, PAIR_T) > pair_update_scratch_floats(P, N, PAIR_T)))
        ra.fused = 1;
    ra.chunk = pl.chunk;
    if (h->profiling) CU(cudaEventRecord(h->ev[0], h->stream));
    pick_fn(h, pl, b.d_eps != nullptr)<<<dim3(nblk, R), pl.T, pl.smem, h->stream>>>(d, ra, in);
    CU(cudaGetLastError());
    if (h->profiling) CU(cudaEventRecord(h->ev[1], h->stream));
    h->launches += 1;
    if (!ra.fused && pl.pair) {
        const int mode = b.d_eps != nullptr ? 2 : d.method == QPMPC_METHOD_MPPI ? 0 : 1;
        h->fn_pair_update[mode]<<<R, PAIR_T, (size_t)pair_update_smem_floats(P, N, PAIR_T, cem_) * 4, h->stream>>>(d, ra, in);
        CU(cudaGetLastError());
        h->launches += 1;
    } else if (!ra.fused) {
        // one block per robot: 512 threads when there are many rollout blocks to merge (single robot), 128 otherwise
        const int fin_threads = nblk > 16 ? FIN_THREADS_MAX : 128;
        const size_t fsmem = (size_t)finalize_smem_floats(P, nblk, K, fin_threads) * 4;
        k_finalize<<<R, fin_threads, fsmem, h->stream>>>(d, ra, in);
        CU(cudaGetLastError());
        h->launches += 1;
    }
    if (h->profiling) CU(cudaEventRecord(h->ev[2], h->stream));
    return QPMPC_OK;
}

extern "C" int qpmpc_solve_batched(qpmpc_handle *h, const qpmpc_batch *batch, uint64_t seed, uint64_t offset) {
    if (!h || !batch) return fail(QPMPC_E_ARG, "null argument");
    const qpmpc_batch &b = *batch;
    if (b.num_robots < 1 || b.num_robots > h->cfg.max_robots) return fail(QPMPC_E_ARG, "num_robots exceeds max_robots");
    if (!b.d_state || !b.d_reference || !b.d_contact_sequence || !b.d_best_in || !b.d_result || !b.d_best_out)
        return fail(QPMPC_E_ARG, "missing device pointer in qpmpc_batch");
    if (h->dc.method == QPMPC_METHOD_CEM && !b.d_sigma_in) return fail(QPMPC_E_ARG, "cem_mppi needs d_sigma_in");
    if (h->dc.ga && !b.d_timing) return fail(QPMPC_E_ARG, "gait adaptive needs d_timing");
    if (h->dc.ga && !b.d_freqs && (!b.d_nominal_freq || !b.d_optimize_swing))
        return fail(QPMPC_E_ARG, "gait adaptive needs d_nominal_freq and d_optimize_swing");
    CU(cudaSetDevice(h->device));
    return launch_solve(h, b, seed, offset);
}

// qpmpc_solve and qpmpc_solve_iterated: `iterations` launches back to back, one wait.  Iteration i reads the mean (and
// sigma) iteration i - 1 left in h->d_iter[(i - 1) & 1]; only the last one publishes to the host.
static int solve_impl(qpmpc_handle *h, const float *state, const float *reference, const float *contact_sequence,
                      const float *best_in, uint64_t seed, uint64_t offset, int iterations, const float *timing,
                      float nominal_step_frequency, int32_t optimize_swing, const float *sigma_in, const float *d_eps,
                      const float *d_freqs, qpmpc_result *out, float *best_out, float *sigma_out, float *costs_out) {
    if (!h || !state || !reference || !contact_sequence || !best_in || !out || !best_out)
        return fail(QPMPC_E_ARG, "null argument");
    const DevCfg &d = h->dc;
    const int H = d.H, P = d.P, N = d.N;
    if (d.method == QPMPC_METHOD_CEM && !sigma_in) return fail(QPMPC_E_ARG, "cem_mppi needs sigma_in");
    if (d.ga && !timing) return fail(QPMPC_E_ARG, "gait adaptive needs timing");
    if (iterations < 1) return fail(QPMPC_E_ARG, "iterations must be >= 1");
    if (iterations > 1 && (d_eps || d_freqs)) return fail(QPMPC_E_ARG, "chained iterations draw their own noise");
    CU(cudaSetDevice(h->device));
    // Default I/O: the ~1 KB of inputs rides in the kernel's parameter block (InlineIn) and the finalize step stores the
    // results straight into a mapped pinned block and raises a flag there that the host spins on -- ONE command on
    // the stream, no copy commands and no stream-synchronise wake-up on the critical path.
    //   QPMPC_INPUTS=copy    one H2D copy from pinned memory in front of the kernel (also used when the inputs do not
    //                        fit the parameter block);   QPMPC_INPUTS=mapped   the kernel reads the inputs over PCIe
    //                        (every block does: slowest, kept as the measured negative result)
    //   QPMPC_ZEROCOPY=0     classic H2D + kernel + D2H + cudaStreamSynchronize
    const EnvSwitches &ev = h->env;
    const bool zero_copy = ev.zerocopy != 0;
    const bool mapped_in = zero_copy && (ev.zerocopy == 1 || ev.inputs == 2);
    const bool inline_in = !mapped_in && h->in_floats <= (size_t)INLINE_FLOATS && ev.inputs != 1 && ev.zerocopy != 0;
    const size_t o_state = 0, o_ref = 24, o_contact = 48, o_best = o_contact + 4 * H, o_sigma = o_best + P,
                 o_timing = o_sigma + P, o_nominal = o_timing + 4, o_oswing = o_nominal + 1;
    const size_t res_floats = sizeof(qpmpc_result) / 4;
    InlineIn inl;
    float *p = inline_in ? inl.v : mapped_in ? h->m_io : h->h_in;  // host view of the input block
    // device view of the input block (byte offsets into the parameter block when the inputs ride inline)
    float *din = inline_in ? static_cast<float *>(nullptr) : mapped_in ? h->md_io : h->d_in;
    const size_t ll_off = (h->in_floats + 3) & ~(size_t)3;  // 16-byte aligned start of the output pairs
    float *hout = h->h_out;
    float *dout = h->d_out;  // unused by the kernel when the outputs go out as pairs
    memcpy(p + o_state, state, 24 * 4);
    memcpy(p + o_ref, reference, 24 * 4);
    memcpy(p + o_contact, contact_sequence, (size_t)4 * H * 4);
    memcpy(p + o_best, best_in, (size_t)P * 4);
    if (sigma_in) memcpy(p + o_sigma, sigma_in, (size_t)P * 4); else memset(p + o_sigma, 0, (size_t)P * 4);
    if (timing) memcpy(p + o_timing, timing, 16); else memset(p + o_timing, 0, 16);
    p[o_nominal] = nominal_step_frequency;
    memcpy(p + o_oswing, &optimize_swing, 4);
    if (!mapped_in && !inline_in)
        CU(cudaMemcpyAsync(h->d_in, h->h_in, h->in_floats * 4, cudaMemcpyHostToDevice, h->stream));
    qpmpc_batch b;
    memset(&b, 0, sizeof b);
    b.num_robots = 1;
    b.d_state = din + o_state; b.d_reference = din + o_ref; b.d_contact_sequence = din + o_contact;
    b.d_best_in = din + o_best; b.d_sigma_in = din + o_sigma; b.d_timing = din + o_timing;
    b.d_nominal_freq = din + o_nominal; b.d_optimize_swing = reinterpret_cast<const int32_t *>(din + o_oswing);
    b.d_eps = d_eps; b.d_freqs = d_freqs;
    b.d_result = reinterpret_cast<qpmpc_result *>(dout);
    b.d_best_out = dout + res_floats;
    b.d_sigma_out = (d.method == QPMPC_METHOD_CEM) ? dout + res_floats + P : nullptr;
    b.d_costs = h->d_costs;
    // iterations 0 .. K-2: outputs stay on the device (ping-pong mean / sigma blocks, scratch result record)
    for (int it = 0; it + 1 < iterations; ++it) {
        qpmpc_batch bi = b;
        float *io = h->d_iter + (size_t)(it & 1) * 2 * P;
        bi.d_best_out = io;
        bi.d_sigma_out = (d.method == QPMPC_METHOD_CEM) ? io + P : nullptr;
        bi.d_result = reinterpret_cast<qpmpc_result *>(h->d_out);
        int rc = launch_solve(h, bi, seed, offset + (uint64_t)it, nullptr, 0, inline_in ? &inl : nullptr, it == 0 ? 7 : 1);
        if (rc) return rc;
        b.d_best_in = io;  // the next iteration starts from this mean (and sigma): device addresses from now on
        b.d_sigma_in = io + P;
    }
    const int last_mask = iterations > 1 ? 1 : 7;
    const uint64_t last_offset = offset + (uint64_t)(iterations - 1);
    if (zero_copy) {
        volatile unsigned long long *pairs = reinterpret_cast<volatile unsigned long long *>(h->m_io + ll_off);
        unsigned int seq = ++h->seq;
        if (seq == 0) seq = ++h->seq;  // the block starts zeroed: 0 never names a solve
        int rc = launch_solve(h, b, seed, last_offset, reinterpret_cast<uint2 *>(h->md_io + ll_off), seq, inline_in ? &inl : nullptr,
                              last_mask);
        if (rc) return rc;
        if (costs_out) {  // 4 N bytes: a regular copy behind the kernel; its completion implies the pairs too
            CU(cudaMemcpyAsync(h->h_out + h->out_floats, h->d_costs, (size_t)N * 4, cudaMemcpyDeviceToHost, h->stream));
            CU(cudaStreamSynchronize(h->stream));
        }
        // wait for every {word, seq} pair of this solve and unpack the words
        const size_t words = res_floats + P + (d.method == QPMPC_METHOD_CEM ? P : 0);
        uint32_t *dst = reinterpret_cast<uint32_t *>(hout);
        unsigned long long spins = 0;
        for (size_t k = 0; k < words; ++k) {
            unsigned long long v;
            while ((unsigned int)((v = pairs[k]) >> 32) != seq) {
                __builtin_ia32_pause();
                if ((++spins & 0xFFFFF) == 0) {  // ~ms: make sure the kernel is still alive rather than spin forever
                    const cudaError_t q = cudaStreamQuery(h->stream);
                    if (q != cudaErrorNotReady && (unsigned int)(pairs[k] >> 32) != seq)
                        return fail(QPMPC_E_CUDA, q != cudaSuccess ? std::string("solve kernel: ") + cudaGetErrorString(q)
                                                                   : std::string("solve finished without publishing its results"));
                }
            }
            dst[k] = (uint32_t)v;
        }
    } else {
        int rc = launch_solve(h, b, seed, last_offset, nullptr, 0, inline_in ? &inl : nullptr, last_mask);
        if (rc) return rc;
        CU(cudaMemcpyAsync(h->h_out, h->d_out, h->out_floats * 4, cudaMemcpyDeviceToHost, h->stream));
        if (costs_out) CU(cudaMemcpyAsync(h->h_out + h->out_floats, h->d_costs, (size_t)N * 4, cudaMemcpyDeviceToHost, h->stream));
        CU(cudaStreamSynchronize(h->stream));
    }
    memcpy(out, hout, sizeof(qpmpc_result));
    memcpy(best_out, hout + res_floats, (size_t)P * 4);
    if (sigma_out && d.method == QPMPC_METHOD_CEM) memcpy(sigma_out, hout + res_floats + P, (size_t)P * 4);
    if (costs_out) memcpy(costs_out, h->h_out + h->out_floats, (size_t)N * 4);
    return QPMPC_OK;
}

extern "C" int qpmpc_solve(qpmpc_handle *h, const float *state, const float *reference, const float *contact_sequence,
                           const float *best_in, uint64_t seed, uint64_t offset, const float *timing,
                           float nominal_step_frequency, int32_t optimize_swing, const float *sigma_in,
                           const float *d_eps, const float *d_freqs, qpmpc_result *out, float *best_out,
                           float *sigma_out, float *costs_out) {
    return solve_impl(h, state, reference, contact_sequence, best_in, seed, offset, 1, timing, nominal_step_frequency,
                      optimize_swing, sigma_in, d_eps, d_freqs, out, best_out, sigma_out, costs_out);
}

extern "C" int qpmpc_solve_iterated(qpmpc_handle *h, const float *state, const float *reference,
                                    const float *contact_sequence, const float *best_in, uint64_t seed,
                                    uint64_t first_offset, int32_t iterations, const float *timing,
                                    float nominal_step_frequency, int32_t optimize_swing, const float *sigma_in,
                                    qpmpc_result *out, float *best_out, float *sigma_out, float *costs_out) {
    return solve_impl(h, state, reference, contact_sequence, best_in, seed, first_offset, iterations, timing,
                      nominal_step_frequency, optimize_swing, sigma_in, nullptr, nullptr, out, best_out, sigma_out,
                      costs_out);
}

extern "C" int qpmpc_copy_costs(qpmpc_handle *h, float *costs_out) {
    if (!h || !costs_out) return fail(QPMPC_E_ARG, "null argument");
    CU(cudaSetDevice(h->device));
    const size_t bytes = (size_t)h->dc.N * 4;
    CU(cudaMemcpyAsync(h->h_out + h->out_floats, h->d_costs, bytes, cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    memcpy(costs_out, h->h_out + h->out_floats, bytes);
    return QPMPC_OK;
}

extern "C" int qpmpc_rollout_costs(qpmpc_handle *h, const float *state, const float *reference,
                                   const float *contact_sequence, const float *timing, const float *d_params,
                                   const float *d_freqs, int32_t n, float *d_costs) {
    if (!h || !state || !reference || !contact_sequence || !d_params || !d_costs) return fail(QPMPC_E_ARG, "null argument");
    if (n < 1) return fail(QPMPC_E_ARG, "n must be >= 1");
    const DevCfg &d = h->dc;
    if (d.ga && (!timing || !d_freqs)) return fail(QPMPC_E_ARG, "gait adaptive needs timing and d_freqs");
    CU(cudaSetDevice(h->device));
    const int H = d.H, P = d.P;
    const bool inline_in = h->in_floats <= (size_t)INLINE_FLOATS;
    InlineIn inl;
    float *p = inline_in ? inl.v : h->h_in;
    float *din = inline_in ? static_cast<float *>(nullptr) : h->d_in;
    memcpy(p, state, 24 * 4);
    memcpy(p + 24, reference, 24 * 4);
    memcpy(p + 48, contact_sequence, (size_t)4 * H * 4);
    const size_t o_timing = 48 + 4 * H + 2 * P;
    if (timing) memcpy(p + o_timing, timing, 16); else memset(p + o_timing, 0, 16);
    if (!inline_in) CU(cudaMemcpyAsync(h->d_in, h->h_in, h->in_floats * 4, cudaMemcpyHostToDevice, h->stream));
    RolloutArgs ra;
    memset(&ra, 0, sizeof ra);
    ra.state = din; ra.reference = din + 24; ra.contact = din + 48; ra.timing = din + o_timing;
    ra.best = din + 48 + 4 * H; ra.sigma = din + 48 + 4 * H + P;  // unused in costs-only mode; kept inside the block
    ra.nominal = din + o_timing + 4; ra.oswing = reinterpret_cast<const int32_t *>(din + o_timing + 5);
    ra.inline_in = inline_in ? 7 : 0;
    ra.eps = d_params; ra.freqs = d_freqs; ra.costs = d_costs; ra.n = n; ra.costs_only = 1;
    const LaunchPlan pl = plan_launch(h, n, n);
    ra.nblk = pl.nblk;
    ra.chunk = pl.chunk;
    pick_fn(h, pl, true)<<<dim3(pl.nblk, 1), pl.T, pl.smem, h->stream>>>(d, ra, inline_in ? inl : g_no_inline);
    CU(cudaGetLastError());
    h->launches += 1;
    CU(cudaStreamSynchronize(h->stream));
    return QPMPC_OK;
}

extern "C" int qpmpc_generate_noise(qpmpc_handle *h, uint64_t seed, uint64_t offset, int32_t robot, const float *sigma_in,
                                    float nominal_step_frequency, int32_t optimize_swing, float *d_eps_out,
                                    float *d_freqs_out) {
    if (!h || !d_eps_out) return fail(QPMPC_E_ARG, "null argument");
    const DevCfg &d = h->dc;
    if (d.method == QPMPC_METHOD_CEM && !sigma_in) return fail(QPMPC_E_ARG, "cem_mppi needs sigma_in");
    CU(cudaSetDevice(h->device));
    const float *d_sigma = nullptr;
    if (d.method == QPMPC_METHOD_CEM) {
        const size_t o_sigma = 48 + 4 * d.H + d.P;
        memcpy(h->h_in + o_sigma, sigma_in, (size_t)d.P * 4);
        CU(cudaMemcpyAsync(h->d_in + o_sigma, h->h_in + o_sigma, (size_t)d.P * 4, cudaMemcpyHostToDevice, h->stream));
        d_sigma = h->d_in + o_sigma;
    }
    const long long total = (long long)d.N * (d.P / NOISE_BLOCK);
    const int threads = 256;
    const long long blocks = (total + threads - 1) / threads;
    k_dump_noise<<<(unsigned)blocks, threads, 0, h->stream>>>(d, NoiseKey{seed, offset}, (uint32_t)robot, d_sigma,
                                                             nominal_step_frequency, optimize_swing, d_eps_out,
                                                             d_freqs_out);
    CU(cudaGetLastError());
    h->launches += 1;
    if (h->env.dump_pair) {  // test hook: overwrite the rows with the packed generator's (must be bit-identical)
        k_dump_noise_pair<<<(unsigned)blocks, threads, 0, h->stream>>>(d, NoiseKey{seed, offset}, (uint32_t)robot, d_sigma,
                                                                      d_eps_out);
        CU(cudaGetLastError());
        h->launches += 1;
    }
    CU(cudaStreamSynchronize(h->stream));
    return QPMPC_OK;
}

extern "C" int qpmpc_set_profiling(qpmpc_handle *h, int32_t on) {
    if (!h) return fail(QPMPC_E_ARG, "null argument");
    h->profiling = on ? 1 : 0;
    return QPMPC_OK;
}

extern "C" int qpmpc_profile_last(qpmpc_handle *h, float *rollout_ms, float *finalize_ms) {
    if (!h || !rollout_ms || !finalize_ms) return fail(QPMPC_E_ARG, "null argument");
    if (!h->profiling) return fail(QPMPC_E_ARG, "profiling is off (qpmpc_set_profiling)");
    CU(cudaSetDevice(h->device));
    CU(cudaEventSynchronize(h->ev[2]));
    CU(cudaEventElapsedTime(rollout_ms, h->ev[0], h->ev[1]));
    CU(cudaEventElapsedTime(finalize_ms, h->ev[1], h->ev[2]));
    return QPMPC_OK;
}

static int gait_check(qpmpc_handle *h, const qpmpc_gait_batch *g) {
    if (!h || !g) return fail(QPMPC_E_ARG, "null argument");
    if (g->num_robots < 1) return fail(QPMPC_E_ARG, "num_robots must be >= 1");
    if (!g->d_phase || !g->d_init || !g->d_phase_offset || !g->d_step_freq || !g->d_duty_factor || !g->d_contact)
        return fail(QPMPC_E_ARG, "missing device pointer in qpmpc_gait_batch");
    return QPMPC_OK;
}

extern "C" int qpmpc_contact_sequences(qpmpc_handle *h, const qpmpc_gait_batch *g) {
    if (int rc = gait_check(h, g)) return rc;
    if (g->horizon < 1 || g->horizon > QPMPC_MAX_H) return fail(QPMPC_E_ARG, "horizon out of range [1, 32]");
    CU(cudaSetDevice(h->device));
    const int threads = 128, blocks = (g->num_robots * 4 + threads - 1) / threads;
    k_gait<<<blocks, threads, 0, h->stream>>>(*g, g->horizon, 0.0);
    CU(cudaGetLastError());
    h->launches += 1;
    return QPMPC_OK;
}

// tau[leg] = -J_leg^T f_leg (interfaces/wb_interface.py:368-371) for a batch of robots, straight
// from the result records the solve left on the device.  One thread per (robot, leg); float64 like numpy, the float32
// forces widened first (numpy promotes the jax float32 array the same way).
__global__ void k_stance_torques(int num_robots, const qpmpc_result *__restrict__ res, const double *__restrict__ jac,
                                 double *__restrict__ tau) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= num_robots * 4) return;
    const int r = idx >> 2, leg = idx & 3;
    const double *J = jac + (size_t)idx * 9;  // feet_jac[leg][:, legs_qvel_idx[leg]], row-major 3 x 3
    const double f0 = (double)res[r].grf[3 * leg], f1 = (double)res[r].grf[3 * leg + 1], f2 = (double)res[r].grf[3 * leg + 2];
#pragma unroll
    for (int j = 0; j < 3; ++j)
        tau[(size_t)idx * 3 + j] = -__dadd_rn(__dadd_rn(__dmul_rn(J[j], f0), __dmul_rn(J[3 + j], f1)), __dmul_rn(J[6 + j], f2));
}

extern "C" int qpmpc_stance_torques(qpmpc_handle *h, int32_t num_robots, const qpmpc_result *d_result,
                                    const double *d_feet_jac, double *d_tau) {
    if (!h) return fail(QPMPC_E_ARG, "null handle");
    if (num_robots < 1 || !d_result || !d_feet_jac || !d_tau) return fail(QPMPC_E_ARG, "stance_torques: null pointer or no robots");
    CU(cudaSetDevice(h->device));
    const int threads = 128, blocks = (num_robots * 4 + threads - 1) / threads;
    k_stance_torques<<<blocks, threads, 0, h->stream>>>(num_robots, d_result, d_feet_jac, d_tau);
    CU(cudaGetLastError());
    h->launches += 1;
    return QPMPC_OK;
}

// dataset rows (MatLogger.record_step, simulation/simulation.py:241-304) for a batch of robots.  One
// thread per (robot, column): column 0 is the time, the others are gathered from the observables' device arrays and
// widened to float64 (exact for every source type).
namespace qp {
__global__ void k_pack_observables(const __grid_constant__ qpmpc_obs_layout L, int num_robots, double time, double *rows) {
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)num_robots * L.num_columns) return;
    const int r = (int)(idx / L.num_columns);
    int col = (int)(idx - (long long)r * L.num_columns);
    double v = time;
    if (col > 0) {
        --col;
        int s = 0;
        while (col >= L.seg[s].count) col -= L.seg[s++].count;
        const qpmpc_obs_segment &g = L.seg[s];
        const long long e = (long long)r * g.robot_stride + col;
        v = g.dtype == QPMPC_OBS_F32   ? (double)static_cast<const float *>(g.d_src)[e]
            : g.dtype == QPMPC_OBS_F64 ? static_cast<const double *>(g.d_src)[e]
            : g.dtype == QPMPC_OBS_I32 ? (double)static_cast<const int32_t *>(g.d_src)[e]
                                       : (double)static_cast<const uint8_t *>(g.d_src)[e];
    }
    rows[idx] = v;
}
}  // namespace qp

extern "C" int qpmpc_pack_observables(qpmpc_handle *h, const qpmpc_obs_layout *layout, int32_t num_robots, double time,
                                      double *d_rows) {
    if (!h || !layout || !d_rows) return fail(QPMPC_E_ARG, "null argument");
    if (num_robots < 1) return fail(QPMPC_E_ARG, "num_robots must be >= 1");
    if (layout->num_segments < 0 || layout->num_segments > QPMPC_MAX_OBS_SEGMENTS)
        return fail(QPMPC_E_ARG, "num_segments out of range [0, 24]");
    int cols = 1;
    for (int k = 0; k < layout->num_segments; ++k) {
        const qpmpc_obs_segment &g = layout->seg[k];
        if (!g.d_src || g.count < 1 || g.dtype < 0 || g.dtype > QPMPC_OBS_U8)
            return fail(QPMPC_E_ARG, "observable segment " + std::to_string(k) + ": null source, empty or unknown dtype");
        cols += g.count;
    }
    if (cols != layout->num_columns) return fail(QPMPC_E_ARG, "num_columns must be 1 + the sum of the segment counts");
    CU(cudaSetDevice(h->device));
    const long long total = (long long)num_robots * cols;
    const int threads = 256;
    k_pack_observables<<<(unsigned)((total + threads - 1) / threads), threads, 0, h->stream>>>(*layout, num_robots, time, d_rows);
    CU(cudaGetLastError());
    h->launches += 1;
    return QPMPC_OK;
}

extern "C" int qpmpc_gait_run(qpmpc_handle *h, const qpmpc_gait_batch *g, double dt) {
    if (int rc = gait_check(h, g)) return rc;
    CU(cudaSetDevice(h->device));
    const int threads = 128, blocks = (g->num_robots * 4 + threads - 1) / threads;
    k_gait<<<blocks, threads, 0, h->stream>>>(*g, 0, dt);
    CU(cudaGetLastError());
    h->launches += 1;
    return QPMPC_OK;
}

// FP32 FMA-chain microbenchmark: the measured denominator of the rollout kernel's roofline.
namespace qp {
__global__ void k_fma_peak(float *out, int iters, float a, float b) {
    float x0 = threadIdx.x, x1 = x0 + 1.f, x2 = x0 + 2.f, x3 = x0 + 3.f, x4 = x0 + 4.f, x5 = x0 + 5.f, x6 = x0 + 6.f,
          x7 = x0 + 7.f;
#pragma unroll 4
    for (int i = 0; i < iters; ++i) {
        x0 = fmaf(x0, a, b); x1 = fmaf(x1, a, b); x2 = fmaf(x2, a, b); x3 = fmaf(x3, a, b);
        x4 = fmaf(x4, a, b); x5 = fmaf(x5, a, b); x6 = fmaf(x6, a, b); x7 = fmaf(x7, a, b);
    }
    const float s = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
    if (s == 123.456f) out[0] = s;
}
}  // namespace qp

extern "C" int qpmpc_measure_fp32_peak(int32_t device, float *tflops) {
    if (!tflops) return fail(QPMPC_E_ARG, "null argument");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) {
        cudaGetLastError();
        return fail(QPMPC_E_NODEVICE, "no such CUDA device");
    }
    CU(cudaSetDevice(device));
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, device));
    float *d_out = nullptr;
    CU(cudaMalloc(&d_out, 4));
    cudaEvent_t e0, e1;
    CU(cudaEventCreate(&e0));
    CU(cudaEventCreate(&e1));
    const int blocks = prop.multiProcessorCount * 8, threads = 256, iters = 1 << 15;
    float best = 0.f;
    for (int rep = 0; rep < 6; ++rep) {
        CU(cudaEventRecord(e0, 0));
        k_fma_peak<<<blocks, threads>>>(d_out, iters, 0.999f, 0.001f);
        CU(cudaEventRecord(e1, 0));
        CU(cudaEventSynchronize(e1));
        float ms = 0.f;
        CU(cudaEventElapsedTime(&ms, e0, e1));
        const double fl = 2.0 * 8.0 * (double)iters * (double)blocks * threads;
        const float tf = (float)(fl / (ms * 1e-3) / 1e12);
        if (rep > 0 && tf > best) best = tf;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(d_out);
    *tflops = best;
    return QPMPC_OK;
}
