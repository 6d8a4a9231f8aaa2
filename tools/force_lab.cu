// force_lab.cu -- kernel laboratory (a development tool, not part of libnbody.so).
// Times variants of the fp32 all-pairs inner loop on synthetic bodies and checks
// each against an fp64 reference computed on the GPU for a sample of bodies, so a
// sweep over tile shape / unroll / occupancy / formulation costs ONE gpurun call.
//   ./force_lab [N=262144] [reps=3]
// Variant axes: R (i-bodies per thread), UNROLL (j-pairs per inner iteration),
// WARPS (compute warps per CTA), STAGES x TILE_PAIRS (smem ring), MINB (min CTAs
// per SM for __launch_bounds__), FORM (12 = baseline op mix, 11 = mass-scaled
// coordinates: one FP32 op fewer per interaction).
#include <cuda_runtime.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "force_f32.cuh"
#include "tma.cuh"

#define CHECK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

using nb::mbar_init; using nb::fence_mbar_init; using nb::mbar_wait; using nb::mbar_expect_tx; using nb::bulk_g2s; using nb::mbar_arrive;
using f32x2 = unsigned long long;

__device__ __forceinline__ f32x2 pack2(float lo, float hi) { f32x2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void unpack2(f32x2 v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) { f32x2 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) { f32x2 d; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) { f32x2 d; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ float rsq(float x) { float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }

struct LabParams {
    const float* jdata;   // FORM 12: 32 B per pair (x0 x1 y0 y1 | z0 z1 m0 m1); FORM 11: 48 B per pair (+ e0 e1 - -)
    const float* ipos;    // float4 (x,y,z,m) AoS, i-bodies
    float4* acc;          // out
    int n;                // bodies (multiple of everything)
    float eps2;
    int chunks;           // j range split into this many chunks; grid = i-tiles x chunks
};

template <int FORM> struct PairBytes { static constexpr int v = FORM == 12 ? 32 : 48; };

// One j-pair against one i-body.
template <int FORM>
__device__ __forceinline__ void interact(const unsigned char* rec, f32x2 nx, f32x2 ny, f32x2 nz, f32x2 eps2, f32x2& ax, f32x2& ay,
                                         f32x2& az);

template <>
__device__ __forceinline__ void interact<12>(const unsigned char* rec, f32x2 nx, f32x2 ny, f32x2 nz, f32x2 eps2, f32x2& ax,
                                             f32x2& ay, f32x2& az)
{
    const float4 A = *reinterpret_cast<const float4*>(rec);
    const float4 B = *reinterpret_cast<const float4*>(rec + 16);
    const f32x2 jx = pack2(A.x, A.y), jy = pack2(A.z, A.w), jz = pack2(B.x, B.y), jm = pack2(B.z, B.w);
    const f32x2 dx = add2(jx, nx), dy = add2(jy, ny), dz = add2(jz, nz);
    f32x2 r2 = fma2(dx, dx, eps2);
    r2 = fma2(dy, dy, r2);
    r2 = fma2(dz, dz, r2);
    float a, b;
    unpack2(r2, a, b);
    const f32x2 inv = pack2(rsq(a), rsq(b));
    const f32x2 inv2 = mul2(inv, inv), mi = mul2(jm, inv), s = mul2(mi, inv2);
    ax = fma2(s, dx, ax);
    ay = fma2(s, dy, ay);
    az = fma2(s, dz, az);
}

// Mass-scaled coordinates: c = m^-1/2, record = (c x, c y, c z, c, eps^2 c^2).
//   d = c (xj - xi) = fma(-xi, c, c xj);  r2s = |d|^2 + eps^2 c^2 = c^2 (r^2 + eps^2)
//   inv = rsqrt(r2s) = 1/(c r);  inv^3 d = m^{3/2} m^{-1/2} (xj - xi) / r^3 = m (xj - xi)/r^3
// m = 0 bodies are stored as c = 0, e = 1: d = 0, inv = 1, contribution exactly 0.
template <>
__device__ __forceinline__ void interact<11>(const unsigned char* rec, f32x2 nx, f32x2 ny, f32x2 nz, f32x2, f32x2& ax, f32x2& ay,
                                             f32x2& az)
{
    const float4 A = *reinterpret_cast<const float4*>(rec);
    const float4 B = *reinterpret_cast<const float4*>(rec + 16);
    const float2 E = *reinterpret_cast<const float2*>(rec + 32);
    const f32x2 jx = pack2(A.x, A.y), jy = pack2(A.z, A.w), jz = pack2(B.x, B.y), jc = pack2(B.z, B.w), je = pack2(E.x, E.y);
    const f32x2 dx = fma2(nx, jc, jx), dy = fma2(ny, jc, jy), dz = fma2(nz, jc, jz);
    f32x2 r2 = fma2(dx, dx, je);
    r2 = fma2(dy, dy, r2);
    r2 = fma2(dz, dz, r2);
    float a, b;
    unpack2(r2, a, b);
    const f32x2 inv = pack2(rsq(a), rsq(b));
    const f32x2 inv2 = mul2(inv, inv), s = mul2(inv2, inv);
    ax = fma2(s, dx, ax);
    ay = fma2(s, dy, ay);
    az = fma2(s, dz, az);
}

template <int R, int UNROLL, int WARPS, int STAGES, int TILE_PAIRS, int MINB, int FORM>
__global__ void __launch_bounds__(WARPS * 32 + 32, MINB) lab_kernel(const LabParams p)
{
    constexpr int PB = PairBytes<FORM>::v;
    constexpr int TILE_BYTES = TILE_PAIRS * PB;
    constexpr int THREADS = WARPS * 32;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + size_t(STAGES) * TILE_BYTES);
    uint64_t* empty = full + STAGES;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int ntiles = (p.n / 2) / TILE_PAIRS / p.chunks;
    const int chunk = blockIdx.x % p.chunks, tile_i = blockIdx.x / p.chunks;
    const unsigned char* src = reinterpret_cast<const unsigned char*>(p.jdata) + size_t(chunk) * ntiles * TILE_BYTES;

    if (tid == 0) {
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], WARPS);
        }
        fence_mbar_init();
    }
    __syncthreads();

    if (warp == WARPS) {
        if (lane == 0) {
            for (int t = 0; t < ntiles; ++t) {
                const int slot = t % STAGES, k = t / STAGES;
                if (k > 0) mbar_wait(&empty[slot], (k - 1) & 1);
                mbar_expect_tx(&full[slot], TILE_BYTES);
                bulk_g2s(smem_raw + size_t(slot) * TILE_BYTES, src + size_t(t) * TILE_BYTES, TILE_BYTES, &full[slot]);
            }
        }
        return;
    }

    const size_t i0 = size_t(tile_i) * (THREADS * R) + tid;
    f32x2 nx[R], ny[R], nz[R], ax[R], ay[R], az[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const float4 me = reinterpret_cast<const float4*>(p.ipos)[i0 + size_t(r) * THREADS];
        nx[r] = pack2(-me.x, -me.x);
        ny[r] = pack2(-me.y, -me.y);
        nz[r] = pack2(-me.z, -me.z);
        ax[r] = ay[r] = az[r] = 0ull;
    }
    const f32x2 eps2 = pack2(p.eps2, p.eps2);

    for (int t = 0; t < ntiles; ++t) {
        const int slot = t % STAGES;
        mbar_wait(&full[slot], (t / STAGES) & 1);
        const unsigned char* tp = smem_raw + size_t(slot) * TILE_BYTES;
#pragma unroll UNROLL
        for (int q = 0; q < TILE_PAIRS; ++q) {
#pragma unroll
            for (int r = 0; r < R; ++r) interact<FORM>(tp + q * PB, nx[r], ny[r], nz[r], eps2, ax[r], ay[r], az[r]);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[slot]);
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
        float xl, xh, yl, yh, zl, zh;
        unpack2(ax[r], xl, xh);
        unpack2(ay[r], yl, yh);
        unpack2(az[r], zl, zh);
        float* o = reinterpret_cast<float*>(&p.acc[i0 + size_t(r) * THREADS]);
        if (p.chunks == 1) {
            o[0] = xl + xh; o[1] = yl + yh; o[2] = zl + zh;
        } else {
            atomicAdd(o, xl + xh); atomicAdd(o + 1, yl + yh); atomicAdd(o + 2, zl + zh);
        }
    }
}

// scalar control: the same record, two scalar interactions (13 issue slots each)
template <>
__device__ __forceinline__ void interact<1>(const unsigned char* rec, f32x2 nx, f32x2 ny, f32x2 nz, f32x2 eps2, f32x2& ax, f32x2& ay,
                                            f32x2& az)
{
    const float4 A = *reinterpret_cast<const float4*>(rec);
    const float4 B = *reinterpret_cast<const float4*>(rec + 16);
    float nxi, nyi, nzi, e2, t;
    unpack2(nx, nxi, t);
    unpack2(ny, nyi, t);
    unpack2(nz, nzi, t);
    unpack2(eps2, e2, t);
    float axl, axh, ayl, ayh, azl, azh;
    unpack2(ax, axl, axh);
    unpack2(ay, ayl, ayh);
    unpack2(az, azl, azh);
    {
        const float dx = A.x + nxi, dy = A.z + nyi, dz = B.x + nzi;
        const float r2 = fmaf(dz, dz, fmaf(dy, dy, fmaf(dx, dx, e2)));
        const float inv = rsq(r2), s = (B.z * inv) * (inv * inv);
        axl = fmaf(s, dx, axl); ayl = fmaf(s, dy, ayl); azl = fmaf(s, dz, azl);
    }
    {
        const float dx = A.y + nxi, dy = A.w + nyi, dz = B.y + nzi;
        const float r2 = fmaf(dz, dz, fmaf(dy, dy, fmaf(dx, dx, e2)));
        const float inv = rsq(r2), s = (B.w * inv) * (inv * inv);
        axh = fmaf(s, dx, axh); ayh = fmaf(s, dy, ayh); azh = fmaf(s, dz, azh);
    }
    ax = pack2(axl, axh); ay = pack2(ayl, ayh); az = pack2(azl, azh);
}
template <> struct PairBytes<1> { static constexpr int v = 32; };

// fp64 reference for a sample of bodies: one block per sampled body
__global__ void ref_kernel(const float4* pos, const int* idx, int n, double eps2, double* out)
{
    const int i = idx[blockIdx.x];
    const double xi = pos[i].x, yi = pos[i].y, zi = pos[i].z;
    double ax = 0, ay = 0, az = 0;
    for (int j = threadIdx.x; j < n; j += blockDim.x) {
        const double dx = pos[j].x - xi, dy = pos[j].y - yi, dz = pos[j].z - zi;
        const double r2 = dx * dx + dy * dy + dz * dz + eps2;
        const double s = j == i ? 0.0 : double(pos[j].w) / (r2 * sqrt(r2));
        ax += s * dx; ay += s * dy; az += s * dz;
    }
    __shared__ double red[3][256];
    red[0][threadIdx.x] = ax; red[1][threadIdx.x] = ay; red[2][threadIdx.x] = az;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (threadIdx.x < o) for (int k = 0; k < 3; ++k) red[k][threadIdx.x] += red[k][threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) for (int k = 0; k < 3; ++k) out[3 * blockIdx.x + k] = red[k][0];
}

struct Lab {
    int n, reps, sms;
    double min_units;
    float eps2;
    float4* d_pos;
    float* d_j12;
    float* d_j11;
    float4* d_acc;
    std::vector<int> sample;
    std::vector<double> ref;
    double ref_max;
};

template <int R, int UNROLL, int WARPS, int STAGES, int TILE_PAIRS, int MINB, int FORM> void run(Lab& L, const char* note = "")
{
    constexpr int PB = PairBytes<FORM>::v;
    const size_t smem = size_t(STAGES) * TILE_PAIRS * PB + 2 * STAGES * 8;
    auto kern = lab_kernel<R, UNROLL, WARPS, STAGES, TILE_PAIRS, MINB, FORM>;
    CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    cudaFuncAttributes fa;
    CHECK(cudaFuncGetAttributes(&fa, kern));
    int occ = 0;
    CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, WARPS * 32 + 32, smem));
    const int itile = WARPS * 32 * R;
    if (L.n % itile || (L.n / 2) % TILE_PAIRS) { printf("skip (n not divisible)\n"); return; }
    const int tiles_i = L.n / itile, slots = occ * L.sms, maxc = (L.n / 2) / TILE_PAIRS;
    int chunks = 1;
    while (chunks < maxc && double(tiles_i) * chunks / slots < L.min_units) chunks *= 2;
    LabParams p{FORM == 11 ? L.d_j11 : L.d_j12, reinterpret_cast<const float*>(L.d_pos), L.d_acc, L.n, L.eps2, chunks};
    const int grid = tiles_i * chunks;
    cudaEvent_t e0, e1;
    CHECK(cudaEventCreate(&e0));
    CHECK(cudaEventCreate(&e1));
    kern<<<grid, WARPS * 32 + 32, smem>>>(p);
    CHECK(cudaGetLastError());
    CHECK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int r = 0; r < L.reps; ++r) {
        CHECK(cudaMemset(L.d_acc, 0, sizeof(float4) * L.n));
        CHECK(cudaEventRecord(e0));
        kern<<<grid, WARPS * 32 + 32, smem>>>(p);
        CHECK(cudaEventRecord(e1));
        CHECK(cudaEventSynchronize(e1));
        float ms;
        CHECK(cudaEventElapsedTime(&ms, e0, e1));
        best = ms < best ? ms : best;
    }
    std::vector<float4> acc(L.n);
    CHECK(cudaMemcpy(acc.data(), L.d_acc, sizeof(float4) * L.n, cudaMemcpyDeviceToHost));
    double err = 0;
    for (size_t k = 0; k < L.sample.size(); ++k) {
        const float4 a = acc[L.sample[k]];
        const double dx = a.x - L.ref[3 * k], dy = a.y - L.ref[3 * k + 1], dz = a.z - L.ref[3 * k + 2];
        err = fmax(err, sqrt(dx * dx + dy * dy + dz * dz));
    }
    const double gint = double(L.n) * L.n / (best * 1e-3) * 1e-9;
    printf("FORM=%2d R=%d U=%d W=%2d ST=%d TP=%4d MINB=%d | regs %3d smem %6zu occ %d warps/SM %2d grid %6d waves %6.2f | %8.3f ms %7.1f Gint/s  %5.2f int/clk/SM | err %.2e %s\n",
           FORM, R, UNROLL, WARPS, STAGES, TILE_PAIRS, MINB, fa.numRegs, smem, occ, occ * (WARPS + 1), grid,
           double(grid) / (occ * L.sms), best, gint, gint * 1e9 / (L.sms * 1.965e9), err / L.ref_max, note);
    fflush(stdout);
}

// the PRODUCT kernel (force_f32.cuh) on the same data, force-only epilogue, same unit balance
template <int R, int U> void run_product(Lab& L, int want_per_slot)
{
    using namespace nb;
    auto kern = force_f32_kernel<R, false, U>;
    const size_t smem = force_f32_smem_bytes();
    CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    cudaFuncAttributes fa;
    CHECK(cudaFuncGetAttributes(&fa, kern));
    int occ = 0;
    CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, kF32Threads + 32, smem));
    const int itile = kF32Threads * R, tiles_i = L.n / itile, slots = occ * L.sms;
    int chunks = (want_per_slot * slots + tiles_i - 1) / tiles_i;
    if (chunks < 1) chunks = 1;
    size_t len = (size_t(L.n) + chunks - 1) / chunks;
    len = (len + kJUnit - 1) / kJUnit * kJUnit;
    chunks = int((L.n + len - 1) / len);
    float* partial;
    unsigned* counters;
    float *vel;
    CHECK(cudaMalloc(&partial, size_t(chunks) * L.n * 16));
    CHECK(cudaMalloc(&counters, sizeof(unsigned) * tiles_i));
    CHECK(cudaMemset(counters, 0, sizeof(unsigned) * tiles_i));
    CHECK(cudaMalloc(&vel, size_t(L.n) * 16));
    ForceParams<float> p{};
    p.pos_cur = L.d_j12; p.out.n = 0; p.vel = vel; p.acc = reinterpret_cast<float*>(L.d_acc); p.partial = partial; p.temp = nullptr;
    p.tile_counter = counters; p.n_j = L.n; p.i_begin = 0; p.n_local = L.n; p.chunk_len = len; p.j_chunks = chunks; p.mode = kForceOnly;
    p.G = 1.f; p.eps2 = L.eps2; p.half_dt = 0.f; p.dt = 0.f; p.heat_decay = 0.f;
    const int grid = tiles_i * chunks;
    cudaEvent_t e0, e1;
    CHECK(cudaEventCreate(&e0));
    CHECK(cudaEventCreate(&e1));
    kern<<<grid, kF32Threads + 32, smem>>>(p);
    CHECK(cudaGetLastError());
    CHECK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int r = 0; r < L.reps; ++r) {
        CHECK(cudaEventRecord(e0));
        kern<<<grid, kF32Threads + 32, smem>>>(p);
        CHECK(cudaEventRecord(e1));
        CHECK(cudaEventSynchronize(e1));
        float ms;
        CHECK(cudaEventElapsedTime(&ms, e0, e1));
        best = ms < best ? ms : best;
    }
    std::vector<float4> acc(L.n);
    CHECK(cudaMemcpy(acc.data(), L.d_acc, sizeof(float4) * L.n, cudaMemcpyDeviceToHost));
    double err = 0;
    for (size_t k = 0; k < L.sample.size(); ++k) {
        const float4 a = acc[L.sample[k]];
        const double dx = a.x - L.ref[3 * k], dy = a.y - L.ref[3 * k + 1], dz = a.z - L.ref[3 * k + 2];
        err = fmax(err, sqrt(dx * dx + dy * dy + dz * dz));
    }
    const double gint = double(L.n) * L.n / (best * 1e-3) * 1e-9;
    printf("PRODUCT force_f32_kernel<R=%d,U=%d> want/slot %3d | regs %3d smem %6zu occ %d chunks %4d grid %6d waves %6.2f | %8.3f ms %7.1f Gint/s  %5.2f int/clk/SM | err %.2e\n",
           R, U, want_per_slot, fa.numRegs, smem, occ, chunks, grid, double(grid) / slots, best, gint, gint * 1e9 / (L.sms * 1.965e9), err / L.ref_max);
    fflush(stdout);
    cudaFree(partial); cudaFree(counters); cudaFree(vel);
}

static double urand(unsigned long long& s)
{
    s += 0x9E3779B97F4A7C15ull;
    unsigned long long z = s;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z ^= z >> 31;
    return double(z >> 11) * (1.0 / 9007199254740992.0);
}

int main(int argc, char** argv)
{
    Lab L;
    L.n = argc > 1 ? atoi(argv[1]) : 262144;
    L.reps = argc > 2 ? atoi(argv[2]) : 3;
    L.min_units = argc > 3 ? atof(argv[3]) : 40.0;
    cudaDeviceProp prop;
    CHECK(cudaGetDeviceProperties(&prop, 0));
    L.sms = prop.multiProcessorCount;
    L.eps2 = 1.0f;
    const int n = L.n;
    printf("force_lab: %s, %d SMs, N=%d, eps2=%g (uniform cube +-50, masses U[0.5,1.5]/N)\n", prop.name, L.sms, n, L.eps2);

    std::vector<float4> pos(n);
    unsigned long long seed = 0x5EED0004ull;
    for (int i = 0; i < n; ++i) {
        pos[i].x = float(100.0 * urand(seed) - 50.0);
        pos[i].y = float(100.0 * urand(seed) - 50.0);
        pos[i].z = float(100.0 * urand(seed) - 50.0);
        pos[i].w = float((0.5 + urand(seed)) / n);
    }
    std::vector<float> j12(size_t(n) * 4), j11(size_t(n) * 6);
    for (int p = 0; p < n / 2; ++p) {
        const float4 a = pos[2 * p], b = pos[2 * p + 1];
        float* r = &j12[size_t(p) * 8];
        r[0] = a.x; r[1] = b.x; r[2] = a.y; r[3] = b.y; r[4] = a.z; r[5] = b.z; r[6] = a.w; r[7] = b.w;
        const float ca = 1.0f / sqrtf(a.w), cb = 1.0f / sqrtf(b.w);
        float* s = &j11[size_t(p) * 12];
        s[0] = a.x * ca; s[1] = b.x * cb; s[2] = a.y * ca; s[3] = b.y * cb; s[4] = a.z * ca; s[5] = b.z * cb; s[6] = ca; s[7] = cb;
        s[8] = L.eps2 * ca * ca; s[9] = L.eps2 * cb * cb; s[10] = 0.f; s[11] = 0.f;
    }
    CHECK(cudaMalloc(&L.d_pos, sizeof(float4) * n));
    CHECK(cudaMalloc(&L.d_j12, sizeof(float) * j12.size()));
    CHECK(cudaMalloc(&L.d_j11, sizeof(float) * j11.size()));
    CHECK(cudaMalloc(&L.d_acc, sizeof(float4) * n));
    CHECK(cudaMemcpy(L.d_pos, pos.data(), sizeof(float4) * n, cudaMemcpyHostToDevice));
    CHECK(cudaMemcpy(L.d_j12, j12.data(), sizeof(float) * j12.size(), cudaMemcpyHostToDevice));
    CHECK(cudaMemcpy(L.d_j11, j11.data(), sizeof(float) * j11.size(), cudaMemcpyHostToDevice));

    // fp64 reference on a sample
    const int ns = 512;
    for (int k = 0; k < ns; ++k) L.sample.push_back(int((long long)k * (n - 1) / (ns - 1)));
    int* d_idx;
    double* d_ref;
    CHECK(cudaMalloc(&d_idx, sizeof(int) * ns));
    CHECK(cudaMalloc(&d_ref, sizeof(double) * 3 * ns));
    CHECK(cudaMemcpy(d_idx, L.sample.data(), sizeof(int) * ns, cudaMemcpyHostToDevice));
    ref_kernel<<<ns, 256>>>(L.d_pos, d_idx, n, double(L.eps2), d_ref);
    CHECK(cudaGetLastError());
    L.ref.resize(3 * ns);
    CHECK(cudaMemcpy(L.ref.data(), d_ref, sizeof(double) * 3 * ns, cudaMemcpyDeviceToHost));
    L.ref_max = 0;
    for (int k = 0; k < ns; ++k)
        L.ref_max = fmax(L.ref_max, sqrt(L.ref[3 * k] * L.ref[3 * k] + L.ref[3 * k + 1] * L.ref[3 * k + 1] + L.ref[3 * k + 2] * L.ref[3 * k + 2]));
    printf("reference: %d sampled bodies, max |acc| = %.6e; err column = max |acc - ref| / max |ref|\n", ns, L.ref_max);

    run_product<4, 2>(L, 48); run_product<4, 4>(L, 48); run_product<4, 6>(L, 48); run_product<4, 8>(L, 48); run_product<4, 16>(L, 48);
    run_product<2, 4>(L, 48); run_product<2, 8>(L, 48); run_product<2, 16>(L, 48);
    //   R  U  W  ST  TP   MINB FORM
    run<4, 4, 4, 2, 256, 1, 12>(L, "<- lab best");
    run<4, 4, 4, 4, 256, 1, 12>(L);
    run<4, 4, 4, 4, 256, 3, 12>(L);
    run<4, 2, 4, 2, 256, 1, 11>(L);
    run<4, 4, 4, 2, 256, 1, 11>(L);
    return 0;
}
