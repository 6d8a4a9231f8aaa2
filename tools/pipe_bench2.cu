// pipe_bench2.cu -- second microbenchmark: what do the PACKED FP32 instructions cost,
// per operand form, and what overlaps with them?  (pipe_bench measured 112 lanes/clk/SM
// for a 3-register FFMA2 stream against 125 for scalar FFMA; the force kernel sits at
// 28.2 cycles per interaction pair per SM sub-partition.)
// Every test is a fully unrolled register-only loop, 8 independent chains per thread,
// 256 threads x 4 CTAs per SM; result = cycles per warp instruction per sub-partition
// (= elapsed SM cycles x 4 sub-partitions / warp instructions issued per SM).
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>

#define CHECK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

using f32x2 = unsigned long long;
__device__ __forceinline__ f32x2 pack2(float lo, float hi) { f32x2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void unpack2(f32x2 v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) { f32x2 d; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) { f32x2 d; asm volatile("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) { f32x2 d; asm volatile("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ float rsq(float x) { float r; asm volatile("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float ffma(float a, float b, float c) { float d; asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c)); return d; }
__device__ __forceinline__ float fadd(float a, float b) { float d; asm volatile("add.rn.f32 %0, %1, %2;" : "=f"(d) : "f"(a), "f"(b)); return d; }

__device__ __forceinline__ f32x2 fma2n(f32x2 a, f32x2 b, f32x2 c) { f32x2 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ f32x2 add2n(f32x2 a, f32x2 b) { f32x2 d; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ f32x2 mul2n(f32x2 a, f32x2 b) { f32x2 d; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ float rsqn(float x) { float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }

constexpr int U = 8;
constexpr int INNER = 16;

#define PROLOGUE                                                                           \
    f32x2 a[U], b[U], c[U];                                                                \
    for (int u = 0; u < U; ++u) {                                                          \
        a[u] = pack2(1.0f + u + threadIdx.x * 1e-3f, 0.5f + u);                            \
        b[u] = pack2(1.0001f + u * 1e-4f, 0.9999f - u * 1e-4f);                            \
        c[u] = pack2(0.5f + u * 1e-2f, 0.25f);                                             \
    }
#define EPILOGUE                                                                           \
    float s = 0, lo, hi;                                                                   \
    for (int u = 0; u < U; ++u) { unpack2(a[u], lo, hi); s += lo + hi; unpack2(b[u], lo, hi); s += lo + hi; unpack2(c[u], lo, hi); s += lo + hi; } \
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;

// T0: FFMA2 a = a*b + c, three distinct registers per instruction, all per-chain
__global__ void t_ffma2_abc(float* out, int iters) { PROLOGUE
    for (int i = 0; i < iters; ++i)
#pragma unroll
        for (int k = 0; k < INNER; ++k)
#pragma unroll
            for (int u = 0; u < U; ++u) a[u] = fma2(a[u], b[u], c[u]);
    EPILOGUE }
// T1: FFMA2 a = b*b + a  (two distinct registers: the r2 form)
__global__ void t_ffma2_bba(float* out, int iters) { PROLOGUE
    for (int i = 0; i < iters; ++i)
#pragma unroll
        for (int k = 0; k < INNER; ++k)
#pragma unroll
            for (int u = 0; u < U; ++u) a[u] = fma2(b[u], b[u], a[u]);
    EPILOGUE }
// T2: FMUL2 a = a*b
__global__ void t_fmul2(float* out, int iters) { PROLOGUE
    for (int i = 0; i < iters; ++i)
#pragma unroll
        for (int k = 0; k < INNER; ++k)
#pragma unroll
            for (int u = 0; u < U; ++u) a[u] = mul2(a[u], b[u]);
    EPILOGUE }
// T3: FADD2 a = a+b
__global__ void t_fadd2(float* out, int iters) { PROLOGUE
    for (int i = 0; i < iters; ++i)
#pragma unroll
        for (int k = 0; k < INNER; ++k)
#pragma unroll
            for (int u = 0; u < U; ++u) a[u] = add2(a[u], b[u]);
    EPILOGUE }
// T4: FMUL2 a = a*a (one distinct register)
__global__ void t_fmul2_aa(float* out, int iters) { PROLOGUE
    for (int i = 0; i < iters; ++i)
#pragma unroll
        for (int k = 0; k < INNER; ++k)
#pragma unroll
            for (int u = 0; u < U; ++u) a[u] = mul2(a[u], a[u]);
    EPILOGUE }
// T5: FFMA2 with a shared (loop-invariant) multiplier: a = a*B + c, B one register pair for all chains
__global__ void t_ffma2_shared(float* out, int iters) { PROLOGUE
    const f32x2 B = b[0];
    for (int i = 0; i < iters; ++i)
#pragma unroll
        for (int k = 0; k < INNER; ++k)
#pragma unroll
            for (int u = 0; u < U; ++u) a[u] = fma2(a[u], B, c[u]);
    EPILOGUE }
// T6: 8 FFMA2 (abc) + 2 MUFU per round: does MUFU issue hide under the FMA pipe?
__global__ void t_ffma2_mufu(float* out, int iters) { PROLOGUE
    float m0 = 1.5f + threadIdx.x, m1 = 2.5f + threadIdx.x;
    for (int i = 0; i < iters; ++i)
#pragma unroll
        for (int k = 0; k < INNER; ++k) {
#pragma unroll
            for (int u = 0; u < U; ++u) a[u] = fma2(a[u], b[u], c[u]);
            m0 = rsq(m0); m1 = rsq(m1);
        }
    EPILOGUE out[0] = m0 + m1; }
// T7: 8 FFMA2 (bba, 2 distinct) + 2 MUFU per round
__global__ void t_ffma2bba_mufu(float* out, int iters) { PROLOGUE
    float m0 = 1.5f + threadIdx.x, m1 = 2.5f + threadIdx.x;
    for (int i = 0; i < iters; ++i)
#pragma unroll
        for (int k = 0; k < INNER; ++k) {
#pragma unroll
            for (int u = 0; u < U; ++u) a[u] = fma2(b[u], b[u], a[u]);
            m0 = rsq(m0); m1 = rsq(m1);
        }
    EPILOGUE out[0] = m0 + m1; }
// T8: scalar FFMA a = a*b + c (control: 1 cycle per warp instruction)
__global__ void t_ffma_scalar(float* out, int iters) {
    float a[U], b[U], c[U];
    for (int u = 0; u < U; ++u) { a[u] = 1.0f + u + threadIdx.x * 1e-3f; b[u] = 1.0001f + u * 1e-4f; c[u] = 0.5f + u * 1e-2f; }
    for (int i = 0; i < iters; ++i)
#pragma unroll
        for (int k = 0; k < INNER; ++k)
#pragma unroll
            for (int u = 0; u < U; ++u) a[u] = ffma(a[u], b[u], c[u]);
    float s = 0;
    for (int u = 0; u < U; ++u) s += a[u] + b[u] + c[u];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// T9: FADD2 with a lane-duplicated operand (the compiler may emit the scalar .F32 operand form)
__global__ void t_fadd2_dup(float* out, int iters) { PROLOGUE
    float xs[U];
    for (int u = 0; u < U; ++u) xs[u] = 1e-3f * (u + 1) + threadIdx.x * 1e-6f;
    for (int i = 0; i < iters; ++i)
#pragma unroll
        for (int k = 0; k < INNER; ++k)
#pragma unroll
            for (int u = 0; u < U; ++u) a[u] = add2(a[u], pack2(xs[u], xs[u]));
    EPILOGUE }
// T10: the interaction-pair mix from registers, 4 chains: 12 packed + 2 MUFU (+1 FADD2 that
// advances jx so nothing is loop-invariant) = 13 packed + 2 MUFU per chain per round
constexpr int CH = 4;
__global__ void t_mix(float* out, int iters) {
    f32x2 jx[CH], jy[CH], jz[CH], jm[CH], ax[CH], ay[CH], az[CH];
    const f32x2 e2 = pack2(0.01f, 0.01f), step = pack2(1e-3f, 2e-3f);
    float nxs[CH], nys[CH], nzs[CH];
    for (int u = 0; u < CH; ++u) {
        jx[u] = pack2(1.f + u, 2.f + u); jy[u] = pack2(0.5f * u, 0.25f); jz[u] = pack2(3.f, 4.f + u); jm[u] = pack2(1e-3f, 2e-3f);
        ax[u] = ay[u] = az[u] = 0ull;
        nxs[u] = -0.1f * threadIdx.x - u; nys[u] = -0.3f - u; nzs[u] = 0.7f + u;
    }
    for (int i = 0; i < iters; ++i)
#pragma unroll
        for (int k = 0; k < INNER; ++k)
#pragma unroll
            for (int u = 0; u < CH; ++u) {
                jx[u] = add2n(jx[u], step);
                const f32x2 dx = add2n(jx[u], pack2(nxs[u], nxs[u])), dy = add2n(jy[u], pack2(nys[u], nys[u])), dz = add2n(jz[u], pack2(nzs[u], nzs[u]));
                f32x2 r2 = fma2n(dx, dx, e2); r2 = fma2n(dy, dy, r2); r2 = fma2n(dz, dz, r2);
                float p, q; unpack2(r2, p, q);
                const f32x2 inv = pack2(rsqn(p), rsqn(q));
                const f32x2 inv2 = mul2n(inv, inv), mi = mul2n(jm[u], inv), s = mul2n(mi, inv2);
                ax[u] = fma2n(s, dx, ax[u]); ay[u] = fma2n(s, dy, ay[u]); az[u] = fma2n(s, dz, az[u]);
            }
    float s = 0, lo, hi;
    for (int u = 0; u < CH; ++u) { unpack2(ax[u], lo, hi); s += lo + hi; unpack2(ay[u], lo, hi); s += lo + hi; unpack2(az[u], lo, hi); s += lo + hi; unpack2(jx[u], lo, hi); s += lo + hi; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// T11: same mix without the MUFU (inv := r2): isolates the packed-pipe cost of the mix
__global__ void t_mix_nomufu(float* out, int iters) {
    f32x2 jx[CH], jy[CH], jz[CH], jm[CH], ax[CH], ay[CH], az[CH];
    const f32x2 e2 = pack2(0.01f, 0.01f), step = pack2(1e-3f, 2e-3f);
    float nxs[CH], nys[CH], nzs[CH];
    for (int u = 0; u < CH; ++u) {
        jx[u] = pack2(1.f + u, 2.f + u); jy[u] = pack2(0.5f * u, 0.25f); jz[u] = pack2(3.f, 4.f + u); jm[u] = pack2(1e-3f, 2e-3f);
        ax[u] = ay[u] = az[u] = 0ull;
        nxs[u] = -0.1f * threadIdx.x - u; nys[u] = -0.3f - u; nzs[u] = 0.7f + u;
    }
    for (int i = 0; i < iters; ++i)
#pragma unroll
        for (int k = 0; k < INNER; ++k)
#pragma unroll
            for (int u = 0; u < CH; ++u) {
                jx[u] = add2n(jx[u], step);
                const f32x2 dx = add2n(jx[u], pack2(nxs[u], nxs[u])), dy = add2n(jy[u], pack2(nys[u], nys[u])), dz = add2n(jz[u], pack2(nzs[u], nzs[u]));
                f32x2 r2 = fma2n(dx, dx, e2); r2 = fma2n(dy, dy, r2); r2 = fma2n(dz, dz, r2);
                const f32x2 inv = r2;
                const f32x2 inv2 = mul2n(inv, inv), mi = mul2n(jm[u], inv), s = mul2n(mi, inv2);
                ax[u] = fma2n(s, dx, ax[u]); ay[u] = fma2n(s, dy, ay[u]); az[u] = fma2n(s, dz, az[u]);
            }
    float s = 0, lo, hi;
    for (int u = 0; u < CH; ++u) { unpack2(ax[u], lo, hi); s += lo + hi; unpack2(ay[u], lo, hi); s += lo + hi; unpack2(az[u], lo, hi); s += lo + hi; unpack2(jx[u], lo, hi); s += lo + hi; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <typename K> void run(const char* name, K kernel, int sms, int iters, double packed_per_round, double other_per_round, int ctas = 4)
{
    float* out;
    CHECK(cudaMalloc(&out, sizeof(float) * 256 * ctas * sms));
    cudaEvent_t e0, e1;
    CHECK(cudaEventCreate(&e0));
    CHECK(cudaEventCreate(&e1));
    kernel<<<sms * ctas, 256>>>(out, iters / 8);
    CHECK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int r = 0; r < 3; ++r) {
        CHECK(cudaEventRecord(e0));
        kernel<<<sms * ctas, 256>>>(out, iters);
        CHECK(cudaEventRecord(e1));
        CHECK(cudaDeviceSynchronize());
        float ms;
        CHECK(cudaEventElapsedTime(&ms, e0, e1));
        best = ms < best ? ms : best;
    }
    // warp-rounds per sub-partition: (256/32 warps) * ctas / 4 sub-partitions * iters * INNER
    const double rounds = 8.0 * ctas / 4.0 * double(iters) * INNER;
    const double cyc = best * 1e-3 * 1.965e9;  // SM clock under load on this part (bench.py records it)
    printf("%-44s %8.3f ms  %7.2f cycles/round/SMSP  = %5.2f per packed op (%g packed + %g other per round)\n", name, best, cyc / rounds,
           cyc / rounds / packed_per_round, packed_per_round, other_per_round);
    fflush(stdout);
    cudaFree(out);
}

int main()
{
    cudaDeviceProp prop;
    CHECK(cudaGetDeviceProperties(&prop, 0));
    const int sms = prop.multiProcessorCount;
    printf("pipe_bench2: %s, %d SMs; cycles assume 1965 MHz\n", prop.name, sms);
    const int it = 4000;
    run("T8  scalar FFMA a*b+c (control)", t_ffma_scalar, sms, it, U, 0);
    run("T0  FFMA2 a*b+c (3 distinct regs)", t_ffma2_abc, sms, it, U, 0);
    run("T1  FFMA2 b*b+a (2 distinct regs)", t_ffma2_bba, sms, it, U, 0);
    run("T5  FFMA2 a*B+c (B shared by all chains)", t_ffma2_shared, sms, it, U, 0);
    run("T2  FMUL2 a*b", t_fmul2, sms, it, U, 0);
    run("T4  FMUL2 a*a", t_fmul2_aa, sms, it, U, 0);
    run("T3  FADD2 a+b", t_fadd2, sms, it, U, 0);
    run("T9  FADD2 a+dup(x)", t_fadd2_dup, sms, it, U, 0);
    run("T6  8 FFMA2(abc) + 2 MUFU", t_ffma2_mufu, sms, it, U, 2);
    run("T7  8 FFMA2(bba) + 2 MUFU", t_ffma2bba_mufu, sms, it, U, 2);
    run("T10 mix: 4 x (13 packed + 2 MUFU)", t_mix, sms, it / 2, 13 * CH, 2 * CH);
    run("T11 mix without MUFU: 4 x 13 packed", t_mix_nomufu, sms, it / 2, 13 * CH, 0);
    run("T10 mix, 2 CTAs/SM", t_mix, sms, it / 2, 13 * CH, 2 * CH, 2);
    run("T0  FFMA2 abc, 2 CTAs/SM", t_ffma2_abc, sms, it, U, 0, 2);
    return 0;
}
