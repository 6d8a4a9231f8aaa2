// fp32 all-pairs force kernel (one-sided sweep) for sm_100a, fused with the kick/drift integrator.
// Since round 2 the large systems run on the pairs-once kernel (force_f32_sym.cuh); this sweep serves
// softening == 0, contexts with radii (nearest-neighbour tracking for the contact stage), mid-size
// systems, multi-GPU contexts without peer access -- and the diagonal blocks of the pairs-once kernel.
// Replaces PhysicsSystem.computeGravity (/root/reference/index.html:264-291) and
// the two O(N) loops of integrate (idx:250-253, 258-261).
//
// Work unit = one CTA = (i-tile of kF32Threads*R bodies) x (j-chunk).  i-bodies live
// in registers; the j-chunk streams through a kF32Stages-deep shared-memory ring of
// FULL 8 KiB tiles (512 bodies) filled by 1-D bulk copies (TMA, UBLKCP) issued by a
// dedicated producer warp and tracked with full/empty mbarriers.  pos_mass is tile-SoA in
// HBM with couples of bodies adjacent (common.cuh), so one broadcast LDS.128 per plane
// delivers four j-bodies as aligned 64-bit register pairs, and each interaction PAIR costs
//     3 FADD2 + 3 FFMA2 + 2 MUFU.RSQ + 3 FMUL2 + 3 FFMA2
// = 12 packed FP32 instructions + 2 MUFU for two interactions: the 12 FP32-lane-ops
// per interaction of the roofline in 7 issue slots instead of 13.
// G is applied once per body.
//
// Shape (4 compute warps, R = 4, unroll 16, 128 registers, 3 CTAs/SM) is the best of
// the force_lab sweeps (profiles/r1_force_lab_sweep*.txt): 9.2 interactions/clk/SM.
// The plateau is insensitive to ring depth, unroll and occupancy, and an 11-op
// formulation (mass-scaled coordinates) does not move it: the packed FP32
// instructions sustain ~112 lanes/clk/SM (pipe_bench), i.e. 12 x 2.29 = 27.5 cycles
// per interaction pair per SM sub-partition, and the kernel runs at 27.8.
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

// Shapes re-measured after the layout change of round 2 (profiles/r2b_variants.jsonl, N=1M, G interactions/s): general masses
// unroll 12 at 3 CTAs/SM 2,619 (unroll 8: 2,552, unroll 16 at 2 CTAs: 2,356, unbounded registers: 2,471); equal masses unroll 8
// at 2 CTAs/SM 2,866 (unroll 16: 2,745, unbounded: 2,633).
#ifndef NB_F32_UNROLL
#define NB_F32_UNROLL 12
#endif
#ifndef NB_F32_LOOP_FORM
#define NB_F32_LOOP_FORM 0  // how consume_tile states its shared-memory loads (0: LDS.128 per plane and quad, re-stated per i-body; 1: hoisted; 2: LDS.64 per couple)
#endif
#ifndef NB_F32_MIN_CTAS
#define NB_F32_MIN_CTAS 3   // __launch_bounds__ second argument of force_f32_kernel: general masses ...
#endif
#ifndef NB_F32_MIN_CTAS_UNIFORM
#define NB_F32_MIN_CTAS_UNIFORM 2  // ... and the equal-mass instances
#endif
constexpr int kF32Unroll = NB_F32_UNROLL;  // j-pairs per inner-loop iteration
// TP = j-PAIRS per ring stage: kF32TilePairs (512 bodies, 8 KiB) for large systems, kF32SmallTilePairs (128 bodies,
// 2 KiB) for systems of up to 4,096 bodies, where a CTA's 512 x 512 interactions are a ~20 us dependent chain whatever
// the SM count and four times as many, four times shorter CTAs fill the chip instead: N=1,024 24.5 -> 13.1 us/step,
// N=4,096 24.6 -> 20.4 (nbody_capi.cu picks by body count; from 8,192 bodies up the large tile wins).
constexpr int kF32SmallTilePairs = 64;
template <int TP = kF32TilePairs> constexpr size_t force_f32_smem_bytes() { return size_t(kF32Stages) * TP * 32 + 2 * kF32Stages * 8; }

// ZERO_SOFT: softening == 0.  The reference never evaluates the self pair
// (idx:273, j starts at i+1); an all-pairs sweep does, and with eps = 0 it
// would be 0 * inf.  Clamping r^2 (FMNMX, ALU pipe, not the FMA pipe) makes the
// self term an exact 0 again.  Distinct coincident bodies then also give 0
// where the reference gives NaN.
//
// NEAR: also keep, per i-body, max over j != i of 1/sqrt(d^2+eps^2) -- the nearest
// neighbour -- for the collision broad phase (contacts.cuh; resolveCollisions'
// overlap search, idx:301-313, rides on the sweep that already visits every pair).
// One FMNMX per j-pair on the ALU pipe.  DIAG marks the one j-tile that contains
// the CTA's own i-bodies, where the self pair has to be masked out of the maximum.
//
// UNIFORM: every body has the same mass (BASELINE configs 4 and 5 do).  The per-pair m_j
// multiply drops out -- s = inv^3, 11 packed FP32 instructions per j-pair instead of 12 --
// and the common mass joins G in the epilogue scale.  Pad bodies then sit at 1e18 instead
// of carrying zero mass: inv^3 underflows to an exact 0 there.  Measured at N=1M: 2,829 G
// interactions/s against 2,651 for the general kernel (+6.7 %); unroll 16 (166 registers,
// 2 CTAs/SM) beats unroll 8 (2,809), 32 (2,741) and 4 (2,724); profiles/r1d_uniform_variants.txt.
template <int R, bool ZERO_SOFT, int U, bool NEAR, bool DIAG, bool UNIFORM, int TP>
__device__ __forceinline__ void consume_tile(const float* __restrict__ tp, const f32x2 (&nx)[R], const f32x2 (&ny)[R],
                                             const f32x2 (&nz)[R], f32x2 (&ax)[R], f32x2 (&ay)[R], f32x2 (&az)[R],
                                             float (&nr)[R], const f32x2 eps2, const int self0)
{
    // the stage holds 2 TP bodies as four planes x | y | z | m of 2 TP floats (common.cuh: tile-SoA); one broadcast
    // LDS.128 per plane delivers two couples = four j-bodies
    constexpr int kPlane = 2 * TP;
    int sslot[R];
    if (NEAR && DIAG) {
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int sb = self0 + r * kF32Threads;  // this body's index inside the stage (outside [0, 2 TP): not in it)
            sslot[r] = (sb >= 0 && sb < kPlane) ? int(tile_slot(unsigned(sb))) : -1;
        }
    }
    constexpr int kU4 = U / 2 > 0 ? U / 2 : 1;  // U j-pairs per iteration = U / 2 quads
#pragma unroll kU4
    for (int q4 = 0; q4 < TP / 2; ++q4) {
#if NB_F32_LOOP_FORM == 1
        const float4 X = reinterpret_cast<const float4*>(tp)[q4];
        const float4 Y = reinterpret_cast<const float4*>(tp + kPlane)[q4];
        const float4 Z = reinterpret_cast<const float4*>(tp + 2 * kPlane)[q4];
        float4 M = make_float4(0.f, 0.f, 0.f, 0.f);
        if (!UNIFORM) M = reinterpret_cast<const float4*>(tp + 3 * kPlane)[q4];
#endif
#pragma unroll
        for (int h = 0; h < 2; ++h) {
#if NB_F32_LOOP_FORM == 2
            const float2 X2 = reinterpret_cast<const float2*>(tp)[2 * q4 + h];
            const float2 Y2 = reinterpret_cast<const float2*>(tp + kPlane)[2 * q4 + h];
            const float2 Z2 = reinterpret_cast<const float2*>(tp + 2 * kPlane)[2 * q4 + h];
            float2 M2 = make_float2(0.f, 0.f);
            if (!UNIFORM) M2 = reinterpret_cast<const float2*>(tp + 3 * kPlane)[2 * q4 + h];
#endif
#pragma unroll
            for (int r = 0; r < R; ++r) {
#if NB_F32_LOOP_FORM == 0
                // (the loads are re-stated per i-body and merged by the compiler)
                const float4 X = reinterpret_cast<const float4*>(tp)[q4];
                const float4 Y = reinterpret_cast<const float4*>(tp + kPlane)[q4];
                const float4 Z = reinterpret_cast<const float4*>(tp + 2 * kPlane)[q4];
#endif
#if NB_F32_LOOP_FORM == 2
                const f32x2 jx = pack2(X2.x, X2.y), jy = pack2(Y2.x, Y2.y), jz = pack2(Z2.x, Z2.y);
#else
                const f32x2 jx = h ? pack2(X.z, X.w) : pack2(X.x, X.y);
                const f32x2 jy = h ? pack2(Y.z, Y.w) : pack2(Y.x, Y.y);
                const f32x2 jz = h ? pack2(Z.z, Z.w) : pack2(Z.x, Z.y);
#endif
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
                const float ia = rsqrt_approx(r2a), ib = rsqrt_approx(r2b);
                if (NEAR) {
                    float ma = ia, mb = ib;
                    if (DIAG) {
                        const int s0 = 4 * q4 + 2 * h;  // slot of this couple's first body
                        ma = (s0 == sslot[r]) ? 0.f : ia;
                        mb = (s0 + 1 == sslot[r]) ? 0.f : ib;
                    }
                    nr[r] = fmaxf(nr[r], fmaxf(ma, mb));
                }
                const f32x2 inv = pack2(ia, ib);
                const f32x2 inv2 = mul2(inv, inv);
                f32x2 mi = inv;
                if (!UNIFORM) {
#if NB_F32_LOOP_FORM == 0
                    const float4 M = reinterpret_cast<const float4*>(tp + 3 * kPlane)[q4];
#endif
#if NB_F32_LOOP_FORM == 2
                    mi = mul2(pack2(M2.x, M2.y), inv);
#else
                    mi = mul2(h ? pack2(M.z, M.w) : pack2(M.x, M.y), inv);
#endif
                }
                const f32x2 s = mul2(mi, inv2);  // m_j / (r^2+eps^2)^1.5   (UNIFORM: 1 / ...)
                ax[r] = fma2(s, dx, ax[r]);
                ay[r] = fma2(s, dy, ay[r]);
                az[r] = fma2(s, dz, az[r]);
            }
        }
    }
}

// UNIFORM with exceptions (config 3: two mass-500 centres among 131,070 equal disc bodies):
// the sweep treats every j as the common mass m0; here body i gets the difference,
//     sum = m0 * (unit-mass sum) + sum_k (m_k - m0) * d_ik / (d_ik^2 + eps^2)^1.5,
// from the current positions of the few exceptions -- O(N * outliers) scalar work in the
// CTA that finishes the tile, against the O(N^2) sweep's saved multiply.
template <bool ZERO_SOFT>
__device__ __forceinline__ void outlier_correction(const ForceParams<float>& p, float x, float y, float z, float& sx, float& sy,
                                                   float& sz)
{
    float cx = 0.f, cy = 0.f, cz = 0.f;
    for (int k = 0; k < p.n_outliers; ++k) {
        const MassOutlier o = p.outliers[k];
        const float dx = p.pos_cur[pm_index<float>(o.index, 0)] - x;
        const float dy = p.pos_cur[pm_index<float>(o.index, 1)] - y;
        const float dz = p.pos_cur[pm_index<float>(o.index, 2)] - z;
        float r2 = fmaf(dz, dz, fmaf(dy, dy, fmaf(dx, dx, p.eps2)));
        if (ZERO_SOFT) r2 = fmaxf(r2, 1e-20f);
        const float inv = rsqrt_approx(r2);
        const float w = inv * (inv * inv);  // the self pair (an exception's own row) is w * 0 = 0 whatever dm is
        cx = fmaf(o.dm, w * dx, cx);
        cy = fmaf(o.dm, w * dy, cy);
        cz = fmaf(o.dm, w * dz, cz);
    }
    sx = fmaf(sx, p.bulk_mass, cx);
    sy = fmaf(sy, p.bulk_mass, cy);
    sz = fmaf(sz, p.bulk_mass, cz);
}

template <int R, bool ZERO_SOFT, int U, bool NEAR, bool UNIFORM = false, int TP = kF32TilePairs>
__device__ __forceinline__ void force_f32_body(const ForceParams<float>& p)
{
    constexpr int kWarps = kF32Threads / 32;
    constexpr int kF32TileBytes = TP * 32;
    constexpr int kTileBodies = 2 * TP;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + size_t(kF32Stages) * kF32TileBytes);
    uint64_t* empty = full + kF32Stages;
    __shared__ int s_last;

    const int tid = threadIdx.x;
    const int warp = tid >> 5;
    const int lane = tid & 31;
    const int chunk = blockIdx.x % p.j_chunks;
    const size_t tile_i = blockIdx.x / p.j_chunks;

    // chunk_len and n_j are multiples of the j unit = one tile: every tile is full
    const size_t j0 = size_t(chunk) * p.chunk_len;
    const size_t j1 = (j0 + p.chunk_len < p.n_j) ? j0 + p.chunk_len : p.n_j;
    const int ntiles = static_cast<int>((j1 - j0) / kTileBodies);
    // the equal-mass sweep never reads the mass plane: three planes per stage
    constexpr int kPlanes = UNIFORM ? 3 : 4;

    if (tid == 0) {
        for (int s = 0; s < kF32Stages; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], kWarps);
        }
        fence_mbar_init();
    }
    __syncthreads();

    if (warp == kWarps) {  // ---- producer warp: one lane drives the TMA ring ----
        if (lane == 0) {
            for (int t = 0; t < ntiles; ++t) {
                const int slot = t % kF32Stages;
                const int k = t / kF32Stages;
                if (k > 0) mbar_wait(&empty[slot], (k - 1) & 1);
                const size_t jb = j0 + size_t(t) * kTileBodies;  // first body of the stage
                NB_CHECK(jb + kTileBodies <= p.n_j && slot >= 0 && slot < kF32Stages);
                const float* src = p.pos_cur + (jb >> 9) * size_t(4 * kPosTile) + (jb & 511);
                unsigned char* dst = smem_raw + size_t(slot) * kF32TileBytes;
                mbar_expect_tx(&full[slot], kPlanes * kTileBodies * 4);
                if (kTileBodies == kPosTile) {  // a whole tile: its planes are contiguous
                    bulk_g2s(dst, src, kPlanes * kTileBodies * 4, &full[slot]);
                } else {
                    for (int c = 0; c < kPlanes; ++c) bulk_g2s(dst + c * kTileBodies * 4, src + c * kPosTile, kTileBodies * 4, &full[slot]);
                }
            }
        }
        return;
    }

    // ---- compute warps ----
    const size_t il0 = tile_i * size_t(kF32Threads * R) + tid;  // local index of this thread's body r=0
    f32x2 nx[R], ny[R], nz[R], ax[R], ay[R], az[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const size_t g = p.i_begin + il0 + size_t(r) * kF32Threads;
        const float x = p.pos_cur[pm_index<float>(g, 0)];
        const float y = p.pos_cur[pm_index<float>(g, 1)];
        const float z = p.pos_cur[pm_index<float>(g, 2)];
        nx[r] = pack2(-x, -x);
        ny[r] = pack2(-y, -y);
        nz[r] = pack2(-z, -z);
        ax[r] = ay[r] = az[r] = 0ull;
    }
    const f32x2 eps2 = pack2(p.eps2, p.eps2);

    float nr[R];
#pragma unroll
    for (int r = 0; r < R; ++r) nr[r] = 0.f;
    // NEAR: the j-tiles that hold this CTA's own i-bodies (one 512-body tile, or four 128-body ones) mask the self pair
    const long long i0 = static_cast<long long>(p.i_begin + tile_i * size_t(kF32Threads * R));

    for (int t = 0; t < ntiles; ++t) {
        const int slot = t % kF32Stages;
        mbar_wait(&full[slot], (t / kF32Stages) & 1);
        const float* tp = reinterpret_cast<const float*>(smem_raw + size_t(slot) * kF32TileBytes);
        const long long off = i0 - (static_cast<long long>(j0) + static_cast<long long>(t) * kTileBodies);  // first own body, relative to the tile
        if (NEAR && off > -static_cast<long long>(kF32Threads * R) && off < kTileBodies)
            consume_tile<R, ZERO_SOFT, U, NEAR, true, UNIFORM, TP>(tp, nx, ny, nz, ax, ay, az, nr, eps2, static_cast<int>(off) + tid);
        else
            consume_tile<R, ZERO_SOFT, U, NEAR, false, UNIFORM, TP>(tp, nx, ny, nz, ax, ay, az, nr, eps2, 0);
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
        for (int r = 0; r < R; ++r) {
            if (UNIFORM && p.n_outliers > 0) {
                float x, y, z, hi;
                unpack2(nx[r], x, hi);
                unpack2(ny[r], y, hi);
                unpack2(nz[r], z, hi);
                outlier_correction<ZERO_SOFT>(p, -x, -y, -z, sx[r], sy[r], sz[r]);
            }
            finish_body<float>(p, il0 + size_t(r) * kF32Threads, sx[r], sy[r], sz[r], p.G);
            if (NEAR) p.near[il0 + size_t(r) * kF32Threads] = nr[r];
        }
        return;
    }

    float4* part = reinterpret_cast<float4*>(p.partial);
    NB_CHECK(chunk < p.j_chunks && il0 + size_t(R - 1) * kF32Threads < p.n_local && tile_i < p.n_local / (kF32Threads * R));
#pragma unroll
    for (int r = 0; r < R; ++r)
        part[size_t(chunk) * p.n_local + il0 + size_t(r) * kF32Threads] = make_float4(sx[r], sy[r], sz[r], NEAR ? nr[r] : 0.f);

    if (!tile_arrive_last(&p.tile_counter[tile_i], p.j_chunks, &s_last, kF32Threads)) return;
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const size_t il = il0 + size_t(r) * kF32Threads;
        float tx = 0.f, ty = 0.f, tz = 0.f, tn = 0.f;
        for (int c = 0; c < p.j_chunks; ++c) {
            const float4 v = __ldcg(&part[size_t(c) * p.n_local + il]);
            tx += v.x;
            ty += v.y;
            tz += v.z;
            if (NEAR) tn = fmaxf(tn, v.w);
        }
        if (UNIFORM && p.n_outliers > 0) {
            float x, y, z, hi;
            unpack2(nx[r], x, hi);
            unpack2(ny[r], y, hi);
            unpack2(nz[r], z, hi);
            outlier_correction<ZERO_SOFT>(p, -x, -y, -z, tx, ty, tz);
        }
        finish_body<float>(p, il, tx, ty, tz, p.G);
        if (NEAR) p.near[il] = tn;
    }
    if (tid == 0) p.tile_counter[tile_i] = 0;  // ready for the next launch
}

template <int R, bool ZERO_SOFT, int U = kF32Unroll, bool UNIFORM = false, int TP = kF32TilePairs>
__global__ void __launch_bounds__(kF32Threads + 32, UNIFORM ? NB_F32_MIN_CTAS_UNIFORM : NB_F32_MIN_CTAS) force_f32_kernel(const ForceParams<float> p)
{
    force_f32_body<R, ZERO_SOFT, U, false, UNIFORM, TP>(p);
}

// The NEAR variant keeps 4 more live values per thread and a second (masked) copy of the
// inner loop; it is built for 2 CTAs per SM (the force_lab sweeps found the plateau flat
// from 9 to 40 warps per SM).
#ifndef NB_NEAR_MIN_CTAS
#define NB_NEAR_MIN_CTAS 2
#endif
#ifndef NB_NEAR_UNROLL
#define NB_NEAR_UNROLL 4  // measured (profiles/r1c_near_variants.jsonl): U=4 @2 CTAs/SM +6.0 % over the plain sweep; U=8 +15 %, U=4 @3 CTAs +11 %
#endif
template <int R, bool ZERO_SOFT, bool UNIFORM = false, int U = NB_NEAR_UNROLL, int TP = kF32TilePairs>
__global__ void __launch_bounds__(kF32Threads + 32, NB_NEAR_MIN_CTAS) force_f32_near_kernel(const ForceParams<float> p)
{
    force_f32_body<R, ZERO_SOFT, U, true, UNIFORM, TP>(p);
}

}  // namespace nb
