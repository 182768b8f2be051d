// Issue-rate microbenchmarks for the instruction classes the rollout kernels are made of (sm_100a).
// Each kernel runs NCH independent dependency chains per thread of one instruction class for `iters` trips;
// the host reports warp-instructions per clock per SM (4 schedulers: 4.0 is the issue ceiling) and, for
// floating-point classes, TFLOP/s.  Build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a
// -o tools/ubench tools/ubench.cu ; run on the GPU box: tools/ubench > gpurun_out/ubench.txt
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

constexpr int NCH = 8;

__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
    unsigned long long ra = *reinterpret_cast<unsigned long long *>(&a), rb = *reinterpret_cast<unsigned long long *>(&b),
                       rc = *reinterpret_cast<unsigned long long *>(&c), rd;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(ra), "l"(rb), "l"(rc));
    return *reinterpret_cast<float2 *>(&rd);
}
__device__ __forceinline__ float2 fmul2(float2 a, float2 b) {
    unsigned long long ra = *reinterpret_cast<unsigned long long *>(&a), rb = *reinterpret_cast<unsigned long long *>(&b), rd;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(rd) : "l"(ra), "l"(rb));
    return *reinterpret_cast<float2 *>(&rd);
}
__device__ __forceinline__ float2 fadd2(float2 a, float2 b) {
    unsigned long long ra = *reinterpret_cast<unsigned long long *>(&a), rb = *reinterpret_cast<unsigned long long *>(&b), rd;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(rd) : "l"(ra), "l"(rb));
    return *reinterpret_cast<float2 *>(&rd);
}

enum Kind {
    K_FFMA_RRR, K_FFMA_RRC, K_FFMA_SHARED_OPS, K_FMUL_RR, K_FADD_RR, K_FFMA2, K_FMUL2, K_FADD2, K_IMADW, K_IMAD, K_LOP3,
    K_IADD3, K_FMNMX, K_FSEL, K_MUFU_LG2, K_MUFU_SIN, K_MUFU_RCP, K_MUFU_SQRT, K_I2F, K_F2I, K_SHFL, K_MIX_FFMA_LOP,
    K_MIX_FFMA2_LOP, K_MIX_FFMA2_IMADW, K_MIX_FFMA_MUFU, K_MIX_ALL, K_LDS, K_FFMA2_BCAST, K_FFMA2_SHARED, K_FFMA2_CONST, K_FFMA2_ACC, K_FADD2_BCAST, K_SHF, K_FFMA_ACC, K_COUNT
};
static const char *names[K_COUNT] = {
    "FFMA r,r,r (distinct regs)", "FFMA r,r,c (const bank)", "FFMA r,r,r (2 shared operands)", "FMUL r,r", "FADD r,r",
    "FFMA2 (fma.rn.f32x2)", "FMUL2 (mul.rn.f32x2)", "FADD2 (add.rn.f32x2)", "IMAD.WIDE (umulhi+lo)", "IMAD (32-bit)", "LOP3",
    "IADD3", "FMNMX", "FSEL/compare-select", "MUFU.LG2", "MUFU.SIN", "MUFU.RCP", "MUFU.SQRT", "I2FP (u32->f32)", "F2I",
    "SHFL.BFLY", "mix 1 FFMA : 1 LOP3", "mix 1 FFMA2 : 1 LOP3", "mix 1 FFMA2 : 1 IMAD.WIDE", "mix 4 FFMA : 1 MUFU",
    "mix FFMA2+IMAD.WIDE+LOP3+MUFU (4:2:2:1)", "LDS.32 conflict-free",
    "FFMA2 p,scalar-bcast,r (R.F32)", "FFMA2 p,q0,r0 (2 shared pairs)", "FFMA2 p,const,const-ish (c-bank/imm)", "FFMA2 acc: p=fma(q,r,p)", "FADD2 p,scalar-bcast", "SHF.L.W (rotate)", "FFMA acc: x=fma(y,z,x)"};

template <int KIND>
__global__ void __launch_bounds__(256) k_bench(float *out, const float *in, int iters, float ca, float cb) {
    __shared__ float s_buf[1024];
    const int t = threadIdx.x;
    s_buf[t] = in[t & 63];
    s_buf[t + 256] = in[(t + 1) & 63];
    s_buf[t + 512] = 0.f;
    s_buf[t + 768] = 0.f;
    __syncthreads();
    float x[NCH], y[NCH], z[NCH];
    float2 p[NCH], q[NCH], r[NCH];
    uint32_t u[NCH], v[NCH];
#pragma unroll
    for (int k = 0; k < NCH; ++k) {
        x[k] = in[(t + k) & 63];
        y[k] = in[(t + k + 7) & 63];
        z[k] = in[(t + k + 13) & 63];
        p[k] = make_float2(x[k], y[k]);
        q[k] = make_float2(y[k], z[k]);
        r[k] = make_float2(z[k], x[k]);
        u[k] = __float_as_uint(x[k]) * 2654435761u + k;
        v[k] = __float_as_uint(y[k]) ^ 0x9E3779B9u;
    }
    const float ya = y[0], za = z[0];
#pragma unroll 1
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int rep = 0; rep < 4; ++rep) {
#pragma unroll
            for (int k = 0; k < NCH; ++k) {
                if (KIND == K_FFMA_RRR) x[k] = fmaf(x[k], y[k], z[k]);
                if (KIND == K_FFMA_RRC) x[k] = fmaf(x[k], y[k], cb);
                if (KIND == K_FFMA_SHARED_OPS) x[k] = fmaf(x[k], ya, za);
                if (KIND == K_FMUL_RR) x[k] = x[k] * y[k];
                if (KIND == K_FADD_RR) x[k] = x[k] + y[k];
                if (KIND == K_FFMA2) p[k] = ffma2(p[k], q[k], r[k]);
                if (KIND == K_FMUL2) p[k] = fmul2(p[k], q[k]);
                if (KIND == K_FADD2) p[k] = fadd2(p[k], q[k]);
                if (KIND == K_IMADW) {
                    const uint32_t hi = __umulhi(0xD2511F53u, u[k]), lo = 0xD2511F53u * u[k];
                    u[k] = hi ^ v[k];
                    v[k] = lo;
                }
                if (KIND == K_IMAD) u[k] = u[k] * v[k] + 12345u;
                if (KIND == K_LOP3) u[k] = (u[k] ^ v[k]) ^ (uint32_t)(i + rep);
                if (KIND == K_IADD3) u[k] = u[k] + v[k] + (uint32_t)i;
                if (KIND == K_FMNMX) x[k] = fminf(fmaxf(x[k], y[k]), z[k]);
                if (KIND == K_FSEL) x[k] = x[k] > y[k] ? z[k] : x[k] + 1.0f;
                if (KIND == K_MUFU_LG2) asm volatile("lg2.approx.ftz.f32 %0, %0;" : "+f"(x[k]));
                if (KIND == K_MUFU_SIN) asm volatile("sin.approx.ftz.f32 %0, %0;" : "+f"(x[k]));
                if (KIND == K_MUFU_RCP) asm volatile("rcp.approx.ftz.f32 %0, %0;" : "+f"(x[k]));
                if (KIND == K_MUFU_SQRT) asm volatile("sqrt.approx.ftz.f32 %0, %0;" : "+f"(x[k]));
                if (KIND == K_I2F) {
                    x[k] = __uint2float_rn(u[k]);
                    u[k] = __float_as_uint(x[k]) + (uint32_t)k;
                }
                if (KIND == K_F2I) {
                    u[k] = (uint32_t)__float2int_rn(x[k]);
                    x[k] = __uint_as_float((u[k] & 0x7FFFFFu) | 0x3F800000u);
                }
                if (KIND == K_SHFL) x[k] = __shfl_xor_sync(0xffffffffu, x[k], 1 + (k & 3));
                if (KIND == K_MIX_FFMA_LOP) {
                    x[k] = fmaf(x[k], y[k], z[k]);
                    u[k] = (u[k] ^ v[k]) ^ (uint32_t)(i + rep);
                }
                if (KIND == K_MIX_FFMA2_LOP) {
                    p[k] = ffma2(p[k], q[k], r[k]);
                    u[k] = (u[k] ^ v[k]) ^ (uint32_t)(i + rep);
                }
                if (KIND == K_MIX_FFMA2_IMADW) {
                    p[k] = ffma2(p[k], q[k], r[k]);
                    const uint32_t hi = __umulhi(0xD2511F53u, u[k]), lo = 0xD2511F53u * u[k];
                    u[k] = hi ^ v[k];
                    v[k] = lo;
                }
                if (KIND == K_MIX_FFMA_MUFU) {
                    x[k] = fmaf(x[k], y[k], z[k]);
                    y[k] = fmaf(y[k], x[k], z[k]);
                    z[k] = fmaf(z[k], y[k], x[k]);
                    x[k] = fmaf(x[k], z[k], y[k]);
                    if ((k & 1) == 0) asm volatile("lg2.approx.ftz.f32 %0, %0;" : "+f"(z[k]));
                    else asm volatile("sin.approx.ftz.f32 %0, %0;" : "+f"(z[k]));
                }
                if (KIND == K_MIX_ALL) {  // per chain: 4 FFMA2, 2 IMAD.WIDE (4 SASS), 2 LOP3, 1 MUFU
                    p[k] = ffma2(p[k], q[k], r[k]);
                    q[k] = ffma2(q[k], r[k], p[k]);
                    const uint32_t hi = __umulhi(0xD2511F53u, u[k]), lo = 0xD2511F53u * u[k];
                    r[k] = ffma2(r[k], p[k], q[k]);
                    u[k] = hi ^ v[k] ^ (uint32_t)i;
                    v[k] = lo ^ (uint32_t)rep;
                    p[k] = ffma2(p[k], r[k], q[k]);
                    asm volatile("lg2.approx.ftz.f32 %0, %0;" : "+f"(x[k]));
                }
                if (KIND == K_LDS) x[k] += s_buf[(t + 32 * k + (i & 1) * 256) & 1023];
                if (KIND == K_FFMA2_BCAST) p[k] = ffma2(p[k], make_float2(y[k], y[k]), r[k]);
                if (KIND == K_FFMA2_SHARED) p[k] = ffma2(p[k], q[0], r[0]);
                if (KIND == K_FFMA2_CONST) p[k] = ffma2(p[k], make_float2(ca, ca), make_float2(cb, cb));
                if (KIND == K_FFMA2_ACC) p[k] = ffma2(q[k], r[k], p[k]);
                if (KIND == K_FADD2_BCAST) p[k] = fadd2(p[k], make_float2(y[k], y[k]));
                if (KIND == K_SHF) u[k] = __funnelshift_l(u[k], u[k], 11) ^ v[k];
                if (KIND == K_FFMA_ACC) x[k] = fmaf(y[k], z[k], x[k]);
            }
        }
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < NCH; ++k) s += x[k] + y[k] + z[k] + p[k].x + p[k].y + q[k].x + r[k].y + __uint_as_float(u[k]) + __uint_as_float(v[k]);
    if (s == 123.456f) out[0] = s;
}

struct Spec {
    int kind;
    double inst_per_chain;  // SASS instructions of the class per chain per rep (for the issue-rate figure)
    double flop_per_chain;
};

template <int KIND>
static void run(float *d_out, float *d_in, int sms, double clk_ghz, double inst_per_chain, double flop_per_chain) {
    const int blocks = sms * 8, threads = 256, iters = 4096;
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep) {
        CK(cudaEventRecord(e0));
        k_bench<KIND><<<blocks, threads>>>(d_out, d_in, iters, 0.999f, 0.001f);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        if (rep > 0 && ms < best) best = ms;
    }
    CK(cudaGetLastError());
    const double chains = (double)blocks * threads * NCH * 4.0 * iters;
    const double winst = chains * inst_per_chain / 32.0;
    const double cycles = best * 1e-3 * clk_ghz * 1e9;
    printf("%-44s %8.3f ms  %6.3f warp-inst/clk/SM  %8.2f TFLOP/s\n", names[KIND], best, winst / cycles / sms,
           chains * flop_per_chain / (best * 1e-3) / 1e12);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
}

int main() {
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    int clk_khz = 0;
    CK(cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0));
    const double clk = clk_khz * 1e-6;
    printf("# %s, %d SMs, clock attr %.3f GHz (rates assume this clock)\n", prop.name, prop.multiProcessorCount, clk);
    float *d_out, *d_in;
    CK(cudaMalloc(&d_out, 64));
    CK(cudaMalloc(&d_in, 256));
    float h_in[64];
    for (int k = 0; k < 64; ++k) h_in[k] = 0.5f + 0.001f * k;
    CK(cudaMemcpy(d_in, h_in, 256, cudaMemcpyHostToDevice));
    const int sms = prop.multiProcessorCount;
    run<K_FFMA_RRR>(d_out, d_in, sms, clk, 1, 2);
    run<K_FFMA_RRC>(d_out, d_in, sms, clk, 1, 2);
    run<K_FFMA_SHARED_OPS>(d_out, d_in, sms, clk, 1, 2);
    run<K_FMUL_RR>(d_out, d_in, sms, clk, 1, 1);
    run<K_FADD_RR>(d_out, d_in, sms, clk, 1, 1);
    run<K_FFMA2>(d_out, d_in, sms, clk, 1, 4);
    run<K_FMUL2>(d_out, d_in, sms, clk, 1, 2);
    run<K_FADD2>(d_out, d_in, sms, clk, 1, 2);
    run<K_IMADW>(d_out, d_in, sms, clk, 2, 0);
    run<K_IMAD>(d_out, d_in, sms, clk, 1, 0);
    run<K_LOP3>(d_out, d_in, sms, clk, 1, 0);
    run<K_IADD3>(d_out, d_in, sms, clk, 1, 0);
    run<K_FMNMX>(d_out, d_in, sms, clk, 2, 2);
    run<K_FSEL>(d_out, d_in, sms, clk, 3, 1);
    run<K_MUFU_LG2>(d_out, d_in, sms, clk, 1, 1);
    run<K_MUFU_SIN>(d_out, d_in, sms, clk, 2, 1);
    run<K_MUFU_RCP>(d_out, d_in, sms, clk, 1, 1);
    run<K_MUFU_SQRT>(d_out, d_in, sms, clk, 1, 1);
    run<K_I2F>(d_out, d_in, sms, clk, 2, 0);
    run<K_F2I>(d_out, d_in, sms, clk, 2, 0);
    run<K_SHFL>(d_out, d_in, sms, clk, 1, 0);
    run<K_MIX_FFMA_LOP>(d_out, d_in, sms, clk, 2, 2);
    run<K_MIX_FFMA2_LOP>(d_out, d_in, sms, clk, 2, 4);
    run<K_MIX_FFMA2_IMADW>(d_out, d_in, sms, clk, 3, 4);
    run<K_MIX_FFMA_MUFU>(d_out, d_in, sms, clk, 5, 8);
    run<K_MIX_ALL>(d_out, d_in, sms, clk, 8, 16);
    run<K_LDS>(d_out, d_in, sms, clk, 2, 1);
    run<K_FFMA2_BCAST>(d_out, d_in, sms, clk, 1, 4);
    run<K_FFMA2_SHARED>(d_out, d_in, sms, clk, 1, 4);
    run<K_FFMA2_CONST>(d_out, d_in, sms, clk, 1, 4);
    run<K_FFMA2_ACC>(d_out, d_in, sms, clk, 1, 4);
    run<K_FADD2_BCAST>(d_out, d_in, sms, clk, 1, 2);
    run<K_SHF>(d_out, d_in, sms, clk, 2, 0);
    run<K_FFMA_ACC>(d_out, d_in, sms, clk, 1, 2);
    return 0;
}
