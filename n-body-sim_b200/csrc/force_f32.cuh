// fp32 all-pairs force kernel for sm_100a, fused with the kick/drift integrator.
// Replaces PhysicsSystem.computeGravity (/root/reference/index.html:264-291) and
// the two O(N) loops of integrate (idx:250-253, 258-261).
//
// Work unit = one CTA = (i-tile of kThreads*R bodies) x (j-chunk).  i-bodies live
// in registers; the j-chunk streams through a kStages-deep shared-memory ring
// filled by 1-D bulk copies (TMA, UBLKCP) issued by a dedicated producer warp and
// tracked with full/empty mbarriers.  Because pos_mass is pair-interleaved in
// HBM, one broadcast LDS.128 pair delivers two j-bodies as aligned 64-bit
// register pairs, and each interaction PAIR costs
//     3 FADD2 + 3 FFMA2 + 2 MUFU.RSQ + 3 FMUL2 + 3 FFMA2
// = 12 packed FP32 instructions + 2 MUFU for two interactions: the 12
// FP32-lane-ops per interaction of the roofline in 7 issue slots
// instead of 13, which leaves issue slots for LDS/loop overhead and makes the
// FP32 pipe, not instruction issue, the bound.  G is applied once per body.
#pragma once
#include "common.cuh"
#include "epilogue.cuh"
#include "tma.cuh"

namespace nb {

using f32x2 = unsigned long long;  // two packed floats in a 64-bit register pair

__device__ __forceinline__ f32x2 pack2(float lo, float hi)
{
    f32x2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack2(f32x2 v, float& lo, float& hi)
{
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c)
{
    f32x2 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b)
{
    f32x2 d;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b)
{
    f32x2 d;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ float rsqrt_approx(float x)
{
    float r;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));  // one MUFU.RSQ
    return r;
}

constexpr size_t force_f32_smem_bytes() { return size_t(kStages) * kTileBytes + 2 * kStages * 8; }

// ZERO_SOFT: softening == 0.  The reference never evaluates the self pair
// (idx:273, j starts at i+1); an all-pairs sweep does, and with eps = 0 it
// would be 0 * inf.  Clamping r^2 (FMNMX, ALU pipe, not the FMA pipe) makes the
// self term an exact 0 again.  Distinct coincident bodies then also give 0
// where the reference gives NaN.
template <int R, bool ZERO_SOFT>
__global__ void __launch_bounds__(kThreads + 32) force_f32_kernel(const ForceParams<float> p)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float4* tiles = reinterpret_cast<float4*>(smem_raw);
    uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + size_t(kStages) * kTileBytes);
    uint64_t* empty = full + kStages;
    __shared__ int s_last;

    const int tid = threadIdx.x;
    const int warp = tid >> 5;
    const int lane = tid & 31;
    const int chunk = blockIdx.x % p.j_chunks;
    const size_t tile_i = blockIdx.x / p.j_chunks;

    const size_t j0 = size_t(chunk) * p.chunk_len;
    const size_t j1 = (j0 + p.chunk_len < p.n_pad) ? j0 + p.chunk_len : p.n_pad;
    const size_t chunk_bytes = (j1 - j0) * 16;
    const int ntiles = static_cast<int>((chunk_bytes + kTileBytes - 1) / kTileBytes);
    const unsigned char* src = reinterpret_cast<const unsigned char*>(p.pos_cur) + j0 * 16;

    if (tid == 0) {
        for (int s = 0; s < kStages; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], kThreads / 32);
        }
        fence_mbar_init();
    }
    __syncthreads();

    if (warp == kThreads / 32) {  // ---- producer warp: one lane drives the TMA ring ----
        if (lane == 0) {
            for (int t = 0; t < ntiles; ++t) {
                const int slot = t % kStages;
                const int k = t / kStages;
                if (k > 0) mbar_wait(&empty[slot], (k - 1) & 1);
                const size_t off = size_t(t) * kTileBytes;
                const uint32_t bytes =
                    static_cast<uint32_t>(chunk_bytes - off < kTileBytes ? chunk_bytes - off : kTileBytes);
                mbar_expect_tx(&full[slot], bytes);
                bulk_g2s(smem_raw + size_t(slot) * kTileBytes, src + off, bytes, &full[slot]);
            }
        }
        return;
    }

    // ---- compute warps ----
    const size_t il0 = tile_i * size_t(kThreads * R) + tid;  // local index of this thread's body r=0
    f32x2 nx[R], ny[R], nz[R], ax[R], ay[R], az[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const size_t g = p.i_begin + il0 + size_t(r) * kThreads;
        const float x = p.pos_cur[pm_index<float>(g, 0)];
        const float y = p.pos_cur[pm_index<float>(g, 1)];
        const float z = p.pos_cur[pm_index<float>(g, 2)];
        nx[r] = pack2(-x, -x);
        ny[r] = pack2(-y, -y);
        nz[r] = pack2(-z, -z);
        ax[r] = ay[r] = az[r] = 0ull;
    }
    const f32x2 eps2 = pack2(p.eps2, p.eps2);

    for (int t = 0; t < ntiles; ++t) {
        const int slot = t % kStages;
        mbar_wait(&full[slot], (t / kStages) & 1);
        const float4* tp = tiles + size_t(slot) * (kTileBytes / 16);
        const size_t off = size_t(t) * kTileBytes;
        const int npairs =
            static_cast<int>((chunk_bytes - off < kTileBytes ? chunk_bytes - off : kTileBytes) / 32);
#pragma unroll 4
        for (int q = 0; q < npairs; ++q) {
            const float4 A = tp[2 * q];      // x0 x1 y0 y1
            const float4 B = tp[2 * q + 1];  // z0 z1 m0 m1
            const f32x2 jx = pack2(A.x, A.y), jy = pack2(A.z, A.w);
            const f32x2 jz = pack2(B.x, B.y), jm = pack2(B.z, B.w);
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const f32x2 dx = add2(jx, nx[r]);
                const f32x2 dy = add2(jy, ny[r]);
                const f32x2 dz = add2(jz, nz[r]);
                f32x2 r2 = fma2(dx, dx, eps2);
                r2 = fma2(dy, dy, r2);
                r2 = fma2(dz, dz, r2);
                float r2a, r2b;
                unpack2(r2, r2a, r2b);
                if (ZERO_SOFT) {
                    r2a = fmaxf(r2a, 1e-20f);
                    r2b = fmaxf(r2b, 1e-20f);
                }
                const f32x2 inv = pack2(rsqrt_approx(r2a), rsqrt_approx(r2b));
                const f32x2 inv2 = mul2(inv, inv);
                const f32x2 mi = mul2(jm, inv);
                const f32x2 s = mul2(mi, inv2);  // m_j / (r^2+eps^2)^1.5
                ax[r] = fma2(s, dx, ax[r]);
                ay[r] = fma2(s, dy, ay[r]);
                az[r] = fma2(s, dz, az[r]);
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[slot]);
    }

    float sx[R], sy[R], sz[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        float lo, hi;
        unpack2(ax[r], lo, hi);
        sx[r] = lo + hi;
        unpack2(ay[r], lo, hi);
        sy[r] = lo + hi;
        unpack2(az[r], lo, hi);
        sz[r] = lo + hi;
    }

    if (p.j_chunks == 1) {
#pragma unroll
        for (int r = 0; r < R; ++r) finish_body<float>(p, il0 + size_t(r) * kThreads, sx[r], sy[r], sz[r], p.G);
        return;
    }

    float4* part = reinterpret_cast<float4*>(p.partial);
#pragma unroll
    for (int r = 0; r < R; ++r)
        part[size_t(chunk) * p.n_local + il0 + size_t(r) * kThreads] = make_float4(sx[r], sy[r], sz[r], 0.f);

    if (!tile_arrive_last(&p.tile_counter[tile_i], p.j_chunks, &s_last)) return;
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const size_t il = il0 + size_t(r) * kThreads;
        float tx = 0.f, ty = 0.f, tz = 0.f;
        for (int c = 0; c < p.j_chunks; ++c) {
            const float4 v = __ldcg(&part[size_t(c) * p.n_local + il]);
            tx += v.x;
            ty += v.y;
            tz += v.z;
        }
        finish_body<float>(p, il, tx, ty, tz, p.G);
    }
    if (tid == 0) p.tile_counter[tile_i] = 0;  // ready for the next launch
}

}  // namespace nb
