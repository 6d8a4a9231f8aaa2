// fp32 pairs-once force pass with WARP-sized work units: the unit of force_f32_sym.cuh cut down to one warp.
//
// The CTA-wide kernel (force_f32_sym.cuh) meets an i-tile of 512 or 1,024 bodies with 512-body position tiles: a unit is
// 30-60 us of one CTA.  That is the right grain at a million bodies and far too coarse at ten thousand, where the whole step
// should take ~15 us: N = 8,192 has 136 such units for 444 resident CTAs.  Here every WARP is its own worker: it pulls
//     (an i-block of 32 R bodies = R/2 groups of 64)  x  (a run of 64-body groups above it)
// from the same kind of host-built table, keeps its i-bodies in registers, fetches the groups straight from L2 (one coalesced
// LDG.64 per plane and lane, prefetched one group ahead), parks each in a private 1 KiB shared-memory slot and rotates it
// through the lanes exactly as sym_stage does: LDS.64 per plane at the rotated index, R pair evaluations against the couple,
// the couple's running sums handed to the neighbour lane.  No TMA ring, no producer warp, no CTA-wide barrier: warps never
// wait for one another, and a unit is ~2 us (R = 4).  The i-block against its own groups (the "diagonal" unit) is evaluated
// one-sided: every ordered pair, i side only.
//
// Accumulation, bounds, poisoning, the finish kernel and the multi-GPU scheme are force_f32_sym.cuh's: partial sums are
// converted to 64-bit fixed point and added with RED.ADD.64, so results do not depend on which warp ran which unit.
#pragma once
#include "force_f32_sym.cuh"

namespace nb {

constexpr int kSymwWarps = 4;
constexpr int kSymwThreads = 32 * kSymwWarps;
constexpr int kSymwGroup = 64;                       // bodies per group: 32 couples (b, b + 32), adjacent in every plane
// One parked group: two arrays of 64 entries of 16 bytes, (x0, x1, y0, y1) and (z0, z1, m0, m1) of couple c at entries c AND
// c + 32 -- every couple is stored twice, so that lane l finds the couple of rotation step q, (l + q) & 31, at entry l + q
// without the wrap: the address is one register plus an immediate (no integer work in the loop) and a step is two LDS.128.
constexpr int kSymwSlotFloats = 2 * 64 * 4;
#ifndef NB_SYMW_UNROLL
#define NB_SYMW_UNROLL 4   // rotation steps per loop iteration (1 / 2 / 4 / 8 / 32 at N=8,192, R=4: 1,906 / 1,934 / 2,167 / 2,164 / 1,963 G int/s)
#endif

// float index of group g's first couple in the tile-SoA position array (common.cuh): tile, plane 0, group inside the tile
__device__ __forceinline__ size_t symw_group_base(unsigned g) { return size_t(g >> 3) * (4 * kPosTile) + size_t(g & 7u) * kSymwGroup; }

struct SymwCouple {
    float2 x, y, z, m;
};

// COHERENT: the positions change DURING the kernel (sym_coop.cuh: other SMs drift them between grid barriers): L2 loads
// (ld.global.cg) instead of the non-coherent path.
template <bool UNIFORM, bool COHERENT> __device__ __forceinline__ SymwCouple symw_fetch(const float* __restrict__ pos, unsigned g, int lane)
{
    const float* b = pos + symw_group_base(g) + 2 * lane;
    SymwCouple c;
    if (COHERENT) {
        c.x = __ldcg(reinterpret_cast<const float2*>(b));
        c.y = __ldcg(reinterpret_cast<const float2*>(b + kPosTile));
        c.z = __ldcg(reinterpret_cast<const float2*>(b + 2 * kPosTile));
        c.m = UNIFORM ? make_float2(0.f, 0.f) : __ldcg(reinterpret_cast<const float2*>(b + 3 * kPosTile));
    } else {
        c.x = __ldg(reinterpret_cast<const float2*>(b));
        c.y = __ldg(reinterpret_cast<const float2*>(b + kPosTile));
        c.z = __ldg(reinterpret_cast<const float2*>(b + 2 * kPosTile));
        c.m = UNIFORM ? make_float2(0.f, 0.f) : __ldg(reinterpret_cast<const float2*>(b + 3 * kPosTile));
    }
    return c;
}

template <bool UNIFORM> __device__ __forceinline__ void symw_park(float* slot, const SymwCouple& c, int lane)
{
    float4* xy = reinterpret_cast<float4*>(slot);
    float4* zm = xy + 64;
    const float4 a = make_float4(c.x.x, c.x.y, c.y.x, c.y.y), b = make_float4(c.z.x, c.z.y, c.m.x, c.m.y);
    xy[lane] = a;
    xy[lane + 32] = a;
    zm[lane] = b;
    zm[lane + 32] = b;
}

// One parked group against this warp's 32 R i-bodies.  DIAG: the group belongs to the i-block itself -- i side only.
template <int R, bool UNIFORM, bool DIAG>
__device__ __forceinline__ void symw_round(const float* __restrict__ slot, const int lane, const f32x2 (&nx)[R], const f32x2 (&ny)[R],
                                           const f32x2 (&nz)[R], const f32x2 (&mi)[R], f32x2 (&ax)[R], f32x2 (&ay)[R], f32x2 (&az)[R],
                                           const f32x2 eps2, f32x2& gx, f32x2& gy, f32x2& gz)
{
    gx = gy = gz = 0ull;
    constexpr int kUnroll = NB_SYMW_UNROLL;
    const float4* xy = reinterpret_cast<const float4*>(slot) + lane;
#pragma unroll kUnroll
    for (int q = 0; q < 32; ++q) {
        const float4 XY = xy[q];
        const float4 ZM = xy[q + 64];
        const f32x2 jx = pack2(XY.x, XY.y), jy = pack2(XY.z, XY.w), jz = pack2(ZM.x, ZM.y);
        const f32x2 jm = UNIFORM ? 0ull : pack2(ZM.z, ZM.w);
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
            const f32x2 inv = pack2(rsqrt_approx(r2a), rsqrt_approx(r2b));
            const f32x2 w = mul2(inv, mul2(inv, inv));  // (distSq + epsSq)^-1.5, idx:280
            const f32x2 wi = UNIFORM ? w : mul2(w, jm);  // f * m_j, idx:282
            ax[r] = fma2(wi, dx, ax[r]);
            ay[r] = fma2(wi, dy, ay[r]);
            az[r] = fma2(wi, dz, az[r]);
            if (!DIAG) {
                const f32x2 wj = UNIFORM ? w : mul2(w, mi[r]);  // f * m_i, idx:286
                gx = fma2(wj, dx, gx);
                gy = fma2(wj, dy, gy);
                gz = fma2(wj, dz, gz);
            }
        }
        if (!DIAG) {  // at step q + 1 this lane meets the couple that lane + 1 met at step q: its sums come along
            gx = __shfl_sync(0xffffffffu, gx, (lane + 1) & 31);
            gy = __shfl_sync(0xffffffffu, gy, (lane + 1) & 31);
            gz = __shfl_sync(0xffffffffu, gz, (lane + 1) & 31);
        }
    }
}

template <int R, bool UNIFORM, bool COHERENT = false> __device__ __forceinline__ void force_f32_symw_body(const Sym32Params& p)
{
    static_assert(R == 2 || R == 4 || R == 8, "an i-block is R/2 whole groups (and the padded body count a whole number of i-blocks: powers of two)");
    constexpr int kH = R / 2;  // groups per i-block
    __shared__ __align__(16) float park[kSymwWarps][2][kSymwSlotFloats];  // 16 KiB
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    float* slot0 = park[warp][0];
    float* slot1 = park[warp][1];
    const f32x2 eps2 = pack2(p.eps2, p.eps2);

    unsigned u = 0;
    if (lane == 0) u = p.u_begin + atomicAdd(p.sched, 1u);
    u = __shfl_sync(0xffffffffu, u, 0);
    while (u < p.u_end) {
        unsigned u_next = 0;
        if (lane == 0) u_next = p.u_begin + atomicAdd(p.sched, 1u);  // (its latency hides behind this unit)
        const uint2 raw = __ldg(reinterpret_cast<const uint2*>(p.units) + u);
        const unsigned ib = raw.x, g0 = raw.y & 0xffffffu, count = (raw.y >> 24) & 0x7fu;
        const bool diag = (raw.y & kSymUnitDiag) != 0u;
        NB_CHECK(count >= 1 && size_t(g0 + count) * kSymwGroup <= p.n_pad && size_t(ib + 1) * kH * kSymwGroup <= p.n_pad);
        SymwCouple nxt = symw_fetch<UNIFORM, COHERENT>(p.pos, g0, lane);

        f32x2 nx[R], ny[R], nz[R], mi[R], ax[R], ay[R], az[R];
#pragma unroll
        for (int h = 0; h < kH; ++h) {
            const SymwCouple c = symw_fetch<UNIFORM, COHERENT>(p.pos, ib * kH + h, lane);
            nx[2 * h] = pack2(-c.x.x, -c.x.x);
            nx[2 * h + 1] = pack2(-c.x.y, -c.x.y);
            ny[2 * h] = pack2(-c.y.x, -c.y.x);
            ny[2 * h + 1] = pack2(-c.y.y, -c.y.y);
            nz[2 * h] = pack2(-c.z.x, -c.z.x);
            nz[2 * h + 1] = pack2(-c.z.y, -c.z.y);
            mi[2 * h] = UNIFORM ? 0ull : pack2(c.m.x, c.m.x);
            mi[2 * h + 1] = UNIFORM ? 0ull : pack2(c.m.y, c.m.y);
        }
#pragma unroll
        for (int r = 0; r < R; ++r) ax[r] = ay[r] = az[r] = 0ull;

#pragma unroll 1
        for (unsigned k = 0; k < count; ++k) {
            float* slot = (k & 1u) ? slot1 : slot0;
            __syncwarp();  // (the lanes are done with this slot's previous group: two rounds ago)
            symw_park<UNIFORM>(slot, nxt, lane);
            if (k + 1 < count) nxt = symw_fetch<UNIFORM, COHERENT>(p.pos, g0 + k + 1, lane);
            __syncwarp();
            f32x2 gx, gy, gz;
            if (diag) {
                symw_round<R, UNIFORM, true>(slot, lane, nx, ny, nz, mi, ax, ay, az, eps2, gx, gy, gz);
            } else {
                symw_round<R, UNIFORM, false>(slot, lane, nx, ny, nz, mi, ax, ay, az, eps2, gx, gy, gz);
                // 32 rotations: lane l holds the sums of couple l = bodies (l, l + 32) of the group again.  acc_j -= (f * m_i) * vec, idx:286-288
                unsigned long long* dst = p.accfix + size_t(g0 + k) * kSymwGroup + 2 * lane;
                float lo, hi;
                unpack2(gx, lo, hi);
                sym_red(dst, -lo, p);
                sym_red(dst + 1, -hi, p);
                unpack2(gy, lo, hi);
                sym_red(dst + p.n_pad, -lo, p);
                sym_red(dst + p.n_pad + 1, -hi, p);
                unpack2(gz, lo, hi);
                sym_red(dst + 2 * p.n_pad, -lo, p);
                sym_red(dst + 2 * p.n_pad + 1, -hi, p);
            }
        }
        // acc_i += (f * m_j) * vec, idx:282-284: the unit's sums of this lane's R i-bodies (couple `lane` of each of the block's groups)
#pragma unroll
        for (int h = 0; h < kH; ++h) {
            unsigned long long* dst = p.accfix + size_t(ib * kH + h) * kSymwGroup + 2 * lane;
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                float lo, hi;
                unpack2(ax[2 * h + e], lo, hi);
                sym_red(dst + e, lo + hi, p);
                unpack2(ay[2 * h + e], lo, hi);
                sym_red(dst + p.n_pad + e, lo + hi, p);
                unpack2(az[2 * h + e], lo, hi);
                sym_red(dst + 2 * p.n_pad + e, lo + hi, p);
            }
        }
        u = __shfl_sync(0xffffffffu, u_next, 0);
    }
}

template <int R, bool UNIFORM>
__global__ void __launch_bounds__(kSymwThreads, R == 8 ? 2 : 3) force_f32_symw_kernel(const Sym32Params p)
{
    force_f32_symw_body<R, UNIFORM>(p);
}

}  // namespace nb
