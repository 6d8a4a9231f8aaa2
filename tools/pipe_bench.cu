// Microbenchmarks behind the roofline denominator (SURVEY.md 8d asks for the
// MUFU rate to be verified before the table is trusted): lane-throughput of
// scalar FFMA, packed FFMA2/FADD2/FMUL2, MUFU.RSQ, and of the bare
// interaction-pair instruction mix with operands in registers.
// Prints one line per test: lanes/clk/SM at the SM clock observed in-kernel.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>

#define CHECK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

using f32x2 = unsigned long long;
__device__ __forceinline__ f32x2 pack2(float lo, float hi) { f32x2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void unpack2(f32x2 v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) { f32x2 d; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) { f32x2 d; asm volatile("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) { f32x2 d; asm volatile("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ float rsq(float x) { float r; asm volatile("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float ffma(float a, float b, float c) { float d; asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c)); return d; }

constexpr int U = 8;

__global__ void k_ffma(float* out, int iters, long long* clk)
{
    float a[U], b = 1.0001f + threadIdx.x * 1e-7f, c = 0.5f;
    for (int u = 0; u < U; ++u) a[u] = u + threadIdx.x;
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i)
#pragma unroll
        for (int u = 0; u < U; ++u) a[u] = ffma(a[u], b, c);
    long long t1 = clock64();
    float s = 0;
    for (int u = 0; u < U; ++u) s += a[u];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *clk = t1 - t0;
}

__global__ void k_ffma2(float* out, int iters, long long* clk)
{
    f32x2 a[U], b = pack2(1.0001f + threadIdx.x * 1e-7f, 0.9999f), c = pack2(0.5f, 0.25f);
    for (int u = 0; u < U; ++u) a[u] = pack2(u + threadIdx.x, u);
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i)
#pragma unroll
        for (int u = 0; u < U; ++u) a[u] = fma2(a[u], b, c);
    long long t1 = clock64();
    float s = 0, lo, hi;
    for (int u = 0; u < U; ++u) { unpack2(a[u], lo, hi); s += lo + hi; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *clk = t1 - t0;
}

__global__ void k_mufu(float* out, int iters, long long* clk)
{
    float a[U];
    for (int u = 0; u < U; ++u) a[u] = 1.5f + u + threadIdx.x;
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i)
#pragma unroll
        for (int u = 0; u < U; ++u) a[u] = rsq(a[u]);
    long long t1 = clock64();
    float s = 0;
    for (int u = 0; u < U; ++u) s += a[u];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *clk = t1 - t0;
}

// the interaction-pair mix with register operands: 12 packed + 2 MUFU per iteration per chain
template <int CH> __global__ void k_mix2(float* out, int iters, long long* clk)
{
    f32x2 ax[CH], ay[CH], az[CH], jx[CH], jy[CH], jz[CH], jm[CH];
    const f32x2 nx = pack2(-0.1f * threadIdx.x, -0.1f * threadIdx.x), ny = pack2(-0.3f, -0.3f), nz = pack2(0.7f, 0.7f), e2 = pack2(0.01f, 0.01f);
    for (int u = 0; u < CH; ++u) {
        ax[u] = ay[u] = az[u] = 0ull;
        jx[u] = pack2(1.f + u, 2.f + u); jy[u] = pack2(0.5f * u, 0.25f); jz[u] = pack2(3.f, 4.f + u); jm[u] = pack2(1e-3f, 2e-3f);
    }
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < CH; ++u) {
            const unsigned long long k = (unsigned long long)(i & 0xff) * 0x100000001ull;  // loop-variant operands
            jx[u] ^= k; jy[u] ^= k << 1; jz[u] ^= k << 2;                                  // (uniform datapath)
            f32x2 dx = add2(jx[u], nx), dy = add2(jy[u], ny), dz = add2(jz[u], nz);
            f32x2 r2 = fma2(dx, dx, e2); r2 = fma2(dy, dy, r2); r2 = fma2(dz, dz, r2);
            float a, b; unpack2(r2, a, b);
            f32x2 inv = pack2(rsq(a), rsq(b));
            f32x2 inv2 = mul2(inv, inv), mi = mul2(jm[u], inv), s = mul2(mi, inv2);
            ax[u] = fma2(s, dx, ax[u]); ay[u] = fma2(s, dy, ay[u]); az[u] = fma2(s, dz, az[u]);
        }
    }
    long long t1 = clock64();
    float s = 0, lo, hi;
    for (int u = 0; u < CH; ++u) { unpack2(ax[u], lo, hi); s += lo + hi; unpack2(ay[u], lo, hi); s += lo + hi; unpack2(az[u], lo, hi); s += lo + hi; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *clk = t1 - t0;
}

// scalar mix: 12 FP32 + 1 MUFU per iteration per chain
template <int CH> __global__ void k_mix1(float* out, int iters, long long* clk)
{
    float ax[CH], ay[CH], az[CH], jx[CH], jy[CH], jz[CH], jm[CH];
    const float xi = 0.1f * threadIdx.x, yi = 0.3f, zi = -0.7f, e2 = 0.01f;
    for (int u = 0; u < CH; ++u) { ax[u] = ay[u] = az[u] = 0.f; jx[u] = 1.f + u; jy[u] = 0.5f * u; jz[u] = 3.f; jm[u] = 1e-3f; }
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < CH; ++u) {
            float dx, dy, dz;
            jx[u] = __int_as_float(__float_as_int(jx[u]) ^ (i & 0xff));  // loop-variant operands (uniform datapath)
            jy[u] = __int_as_float(__float_as_int(jy[u]) ^ ((i & 0xff) << 1));
            jz[u] = __int_as_float(__float_as_int(jz[u]) ^ ((i & 0xff) << 2));
            asm volatile("sub.rn.f32 %0, %1, %2;" : "=f"(dx) : "f"(jx[u]), "f"(xi));
            asm volatile("sub.rn.f32 %0, %1, %2;" : "=f"(dy) : "f"(jy[u]), "f"(yi));
            asm volatile("sub.rn.f32 %0, %1, %2;" : "=f"(dz) : "f"(jz[u]), "f"(zi));
            float r2 = ffma(dx, dx, e2); r2 = ffma(dy, dy, r2); r2 = ffma(dz, dz, r2);
            float inv = rsq(r2), inv2, mi, s;
            asm volatile("mul.rn.f32 %0, %1, %2;" : "=f"(inv2) : "f"(inv), "f"(inv));
            asm volatile("mul.rn.f32 %0, %1, %2;" : "=f"(mi) : "f"(jm[u]), "f"(inv));
            asm volatile("mul.rn.f32 %0, %1, %2;" : "=f"(s) : "f"(mi), "f"(inv2));
            ax[u] = ffma(s, dx, ax[u]); ay[u] = ffma(s, dy, ay[u]); az[u] = ffma(s, dz, az[u]);
        }
    }
    long long t1 = clock64();
    float s = 0;
    for (int u = 0; u < CH; ++u) s += ax[u] + ay[u] + az[u];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *clk = t1 - t0;
}

// FP64 pipe: dependent DFMA chains (U per thread), and the pairs-once interaction mix of force_f64_sym.cuh with register
// operands (per chain and iteration: 3 DADD + 3 DFMA + MUFU.RSQ64H + 3 DMUL + 3 DFMA + 6 DFMA = 18 DP instructions).
__global__ void k_dfma(float* out, int iters, long long* clk)
{
    double a[U], b = 1.0001 + threadIdx.x * 1e-9, c = 0.5;
    for (int u = 0; u < U; ++u) a[u] = u + threadIdx.x;
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i)
#pragma unroll
        for (int u = 0; u < U; ++u) asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(a[u]) : "d"(b), "d"(c));
    long long t1 = clock64();
    double s = 0;
    for (int u = 0; u < U; ++u) s += a[u];
    out[blockIdx.x * blockDim.x + threadIdx.x] = float(s);
    if (threadIdx.x == 0 && blockIdx.x == 0) *clk = t1 - t0;
}

template <int CH, int SEED> __global__ void k_dmix(float* out, int iters, long long* clk)
{
    double ax[CH], ay[CH], az[CH], gx = 0, gy = 0, gz = 0, jx[CH], jy[CH], jz[CH];
    const double xi = 0.1 * threadIdx.x, yi = 0.3, zi = -0.7, e2 = 0.01;
    for (int u = 0; u < CH; ++u) { ax[u] = ay[u] = az[u] = 0.0; jx[u] = 1.0 + u; jy[u] = 0.5 * u; jz[u] = 3.0; }
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < CH; ++u) {
            jx[u] = __longlong_as_double(__double_as_longlong(jx[u]) ^ (long long)(i & 0xff));  // loop-variant operands
            const double dx = __dsub_rn(jx[u], xi), dy = __dsub_rn(jy[u], yi), dz = __dsub_rn(jz[u], zi);
            const double s = __fma_rn(dz, dz, __fma_rn(dy, dy, __fma_rn(dx, dx, e2)));
            double y;
            if (SEED == 0) {
                asm volatile("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(s));   // MUFU.RSQ64H
            } else if (SEED == 1) {
                float sf, yf;                                                        // F2F + MUFU.RSQ + F2F
                asm volatile("cvt.rn.f32.f64 %0, %1;" : "=f"(sf) : "d"(s));
                asm volatile("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(yf) : "f"(sf));
                asm volatile("cvt.f64.f32 %0, %1;" : "=d"(y) : "f"(yf));
            } else {
                // no MUFU at all (how much of the DP peak the rest of the mix reaches): integer seed
                y = __longlong_as_double(0x5FE6EB50C7B537A9ll - (__double_as_longlong(s) >> 1));
            }
            const double c = __dmul_rn(y, y), e = __fma_rn(s, -c, 1.0), c3 = __dmul_rn(c, y), p = __fma_rn(e, 1.875, 1.5);
            const double w = __fma_rn(c3, __dmul_rn(p, e), c3);
            ax[u] = __fma_rn(w, dx, ax[u]); ay[u] = __fma_rn(w, dy, ay[u]); az[u] = __fma_rn(w, dz, az[u]);
            gx = __fma_rn(w, dx, gx); gy = __fma_rn(w, dy, gy); gz = __fma_rn(w, dz, gz);
        }
    }
    long long t1 = clock64();
    double s = gx + gy + gz;
    for (int u = 0; u < CH; ++u) s += ax[u] + ay[u] + az[u];
    out[blockIdx.x * blockDim.x + threadIdx.x] = float(s);
    if (threadIdx.x == 0 && blockIdx.x == 0) *clk = t1 - t0;
}

template <typename K> void run(const char* name, K kernel, int threads, int ctas_per_sm, int sms, int iters, double lanes_per_thread_iter, const char* unit)
{
    float* out;
    long long* clk;
    CHECK(cudaMalloc(&out, sizeof(float) * threads * ctas_per_sm * sms));
    CHECK(cudaMalloc(&clk, sizeof(long long)));
    cudaEvent_t e0, e1;
    CHECK(cudaEventCreate(&e0));
    CHECK(cudaEventCreate(&e1));
    kernel<<<sms * ctas_per_sm, threads>>>(out, iters / 10, clk);  // warm-up
    CHECK(cudaDeviceSynchronize());
    CHECK(cudaEventRecord(e0));
    kernel<<<sms * ctas_per_sm, threads>>>(out, iters, clk);
    CHECK(cudaEventRecord(e1));
    CHECK(cudaDeviceSynchronize());
    float ms;
    CHECK(cudaEventElapsedTime(&ms, e0, e1));
    long long cyc;
    CHECK(cudaMemcpy(&cyc, clk, sizeof cyc, cudaMemcpyDeviceToHost));
    const double total = lanes_per_thread_iter * double(iters) * threads * ctas_per_sm * sms;
    const double per_clk_sm = lanes_per_thread_iter * double(iters) * threads * ctas_per_sm / double(cyc);
    printf("%-34s thr=%4d cta/sm=%d  %9.3f ms  %9.1f G%s/s  %7.2f %s/clk/SM  (block0 %lld cyc, ~%.0f MHz)\n", name, threads, ctas_per_sm, ms,
           total / ms * 1e-6, unit, per_clk_sm, unit, cyc, double(cyc) / ms * 1e-3);
    cudaFree(out);
    cudaFree(clk);
}

int main()
{
    cudaDeviceProp prop;
    CHECK(cudaGetDeviceProperties(&prop, 0));
    const int sms = prop.multiProcessorCount;
    printf("device: %s, %d SMs, clockRate %d kHz\n", prop.name, sms, prop.clockRate);
    const int it = 200000;
    for (int cps : {1, 2, 4}) {
        run("FFMA scalar (lanes=FMA)", k_ffma, 256, cps, sms, it, U, "fma");
        run("FFMA2 packed (lanes=2 FMA)", k_ffma2, 256, cps, sms, it, 2 * U, "fma");
        run("MUFU.RSQ", k_mufu, 256, cps, sms, it, U, "rsq");
    }
    for (int cps : {1, 2, 3, 4}) {
        run("mix scalar 12FP+1MUFU CH=4", k_mix1<4>, 256, cps, sms, it / 4, 4, "int");
        run("mix packed 12x2+2MUFU CH=2", k_mix2<2>, 256, cps, sms, it / 4, 4, "int");
        run("mix packed 12x2+2MUFU CH=4", k_mix2<4>, 256, cps, sms, it / 4, 8, "int");
    }
    for (int cps : {1, 2, 4}) {
        run("DFMA (lanes=DP instr)", k_dfma, 256, cps, sms, it / 4, U, "dfma");
        run("fp64 pairs-once mix 18DP+1MUFU CH=4", k_dmix<4, 0>, 128, cps, sms, it / 16, 4 * 18, "dp");
    }
    run("fp64 pairs-once mix CH=4, 5 CTAs x 128", k_dmix<4, 0>, 128, 5, sms, it / 16, 4 * 18, "dp");
    run("  ... seed by F2F+MUFU.RSQ+F2F", k_dmix<4, 1>, 128, 5, sms, it / 16, 4 * 18, "dp");
    run("  ... integer seed, no MUFU", k_dmix<4, 2>, 128, 5, sms, it / 16, 4 * 18, "dp");
    run("mix packed CH=4 thr=512", k_mix2<4>, 512, 2, sms, it / 4, 8, "int");
    run("mix packed CH=4 thr=1024", k_mix2<4>, 1024, 1, sms, it / 4, 8, "int");
    return 0;
}
