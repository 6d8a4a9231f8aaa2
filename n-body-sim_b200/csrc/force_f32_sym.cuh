// fp32 force pass that evaluates every pair ONCE -- the production kernel for large systems.
//
// The reference walks i < j and applies each pair to both bodies (/root/reference/index.html:272-288:
// acc_i += (f*m_j)*vec, acc_j -= (f*m_i)*vec).  The one-sided sweep (force_f32.cuh) is bound by FP32
// issue slots -- 12 packed instructions + 2 MUFU per couple of interactions, every packed instruction
// holding the dispatch port for two cycles -- and 8 of those 12 (the three differences,
// distSq, inv^2, inv^3) are the same for both bodies of a pair.  Evaluated once, a couple of PAIRS costs
//     3 FADD2 + 3 FFMA2 + 2 MUFU.RSQ + 2 FMUL2 (+ 2 FMUL2 with general masses) + 3 FFMA2 (i) + 3 FFMA2 (j)
// = 14 (16) packed + 2 MUFU for FOUR one-sided interactions: 7.5 (8.5) issue slots per interaction
// instead of 12 (13).
//
// Work unit = (i-tile of 128 R bodies held in registers) x (a run of 512-body position tiles above it),
// or the i-tile against its own bodies (a "diagonal" unit: the one-sided loop of force_f32.cuh, i-side
// only).  Units come from a table the host builds per body count (a pure function of n), and CTAs are
// PERSISTENT: one CTA per resident slot pulls unit after unit from an atomic counter, so there is no
// wave tail and no per-chunk partial-sum array.  A producer warp streams the units' position tiles
// through a ring of 8 KiB shared-memory stages with 1-D bulk copies (TMA) and full/empty mbarriers,
// running ahead across unit boundaries; each stage carries the unit it belongs to.
//
// j side.  A tile is eight groups of 64 bodies = 32 couples (b, b + 32) adjacent in memory (common.cuh).
// A warp takes a group at a time: at step q lane l fetches couple (l + q) & 31 with one conflict-free
// LDS.64 per plane, meets its R i-bodies with it, and hands the couple's running sums to lane l - 1
// (three 64-bit shuffles): every lane works on a different couple, and after 32 steps the sums are home.
// No atomics and no shared-memory read-modify-write inside the loop.
//
// Accumulation.  Every partial sum -- an i-body's sum over one unit, a j-body's sum over one warp's
// 32 R i-bodies -- is an fp32 number that depends on the unit table only.  It is converted to 64-bit
// fixed point (scale 2^k, k from an a-priori bound on |sum|: sum|m| * max_r r/(r^2+eps^2)^1.5) and added
// to the body's accumulator with RED.ADD.64.  Integer addition is exact and associative, so the result
// does not depend on which CTA, which GPU or in which order the partial sums arrive: runs are bit-identical
// whatever the scheduling and whatever the number of GPUs, with no fixed-order reduction pass.
// f32_sym_finish_kernel then converts each body's sums back (one rounding), adds the peers' accumulators
// on several GPUs (each GPU works through its share of the unit table for ALL bodies), and runs the
// integrator epilogue (epilogue.cuh) -- kick-2, thermal, and the next step's kick-1 + drift with the
// position exchange.  A non-finite partial sum (a NaN or infinite coordinate) poisons the step: every
// acceleration becomes NaN, as the reference's does one step later (every pair with that body is NaN).
// Needs softening > 0 (the bound); softening == 0 stays on the one-sided sweep.
#pragma once
#include "common.cuh"
#include "epilogue.cuh"
#include "force_f32.cuh"
#include "tma.cuh"

namespace nb {

#ifndef NB_SYM_STAGES
#define NB_SYM_STAGES 4
#endif
// rotation steps per inner-loop iteration, by kernel variant.  Measured at N=1M (profiles/r2b_variants.jsonl, r2c_variants.jsonl;
// G interactions/s): equal masses, 8 bodies per thread: 4,240 / 4,061 / 4,109 / 4,183 / 4,187 / 3,403 with 1 / 2 / 4 / 8 / 16 / 32;
// general masses, 8 per thread: 3,541 / 3,608 with 1 / 16; 4 per thread (3 CTAs per SM): 3,169 / 3,374 / 3,469 / 3,489 / 3,526 with
// 1 / 2 / 4 / 8 / 16 (general) and 3,859 / 3,920 with 8 / 32 (equal masses).
#ifndef NB_SYM32_SPLIT_G
#define NB_SYM32_SPLIT_G 0  // lab switch: two half-length j-side accumulation chains per step (measured: see DESIGN.md 4.1)
#endif
#ifndef NB_SYM32_REPARK
#define NB_SYM32_REPARK 0  // lab switch: each warp copies the 64-body group it is about to rotate into a private slot where every couple
                           // is stored twice (force_f32_symw.cuh's layout: no index arithmetic in the loop, two LDS.128 per step).
                           // 143 instead of 148 instructions per step and still slower HERE: N=1M equal masses 4,009 / 4,083 / 4,138 /
                           // 4,148 G int/s with unroll 1 / 2 / 4 / 8 against 4,228 (general masses 3,597 -> 3,666): profiles/r2t_repark.jsonl
#endif
#ifndef NB_SYM32_UNROLL_EQUAL
#define NB_SYM32_UNROLL_EQUAL 1
#endif
#ifndef NB_SYM32_UNROLL_GENERAL
#define NB_SYM32_UNROLL_GENERAL 16
#endif
constexpr int kSym32Stages = NB_SYM_STAGES;    // ring depth, 8 KiB (one position tile) per stage
constexpr int kSym32StageBytes = 4 * kPosTile * 4;
constexpr size_t force_f32_sym_smem_bytes() { return size_t(kSym32Stages) * kSym32StageBytes + kSym32Stages * (2 * 8 + 8); }

// (i-tile, run of position tiles): stage0 in bits 0..23, count in 24..30, bit 31 = diagonal unit
struct SymUnit {
    unsigned tile;
    unsigned run;
};
constexpr unsigned kSymUnitDiag = 0x80000000u;
constexpr int kSymUnitMaxRun = 127;

enum : unsigned { kSymFirst = 1u, kSymLast = 2u, kSymDiag = 4u, kSymStop = 8u };

struct Sym32Params {
    const float* pos;            // tile-SoA positions of all bodies (common.cuh)
    const SymUnit* units;        // the unit table (the same on every GPU)
    unsigned u_begin, u_end;     // this GPU's share of it
    unsigned* sched;             // next unit of the share, relative to u_begin; zero at launch
    unsigned long long* accfix;  // [3][n_pad] fixed-point sums, bodies in position-slot order (fix_slot), then ONE more word:
                                 // the poison mark, set when a partial sum is not finite (or beyond the bound); zero at launch
    size_t n_pad;
    float eps2;
    float scale;                 // 2^k
    float bound;                 // |partial sum| <= bound, or the step is poisoned
};

// accumulator index of body g: its tile, then its slot inside a position plane
__host__ __device__ __forceinline__ size_t fix_slot(size_t g) { return (g & ~size_t(511)) + tile_slot(unsigned(g) & 511u); }

__device__ __forceinline__ void sym_red(unsigned long long* dst, float v, const Sym32Params& p)
{
    if (!(fabsf(v) <= p.bound)) atomicOr(p.accfix + 3 * p.n_pad, 1ull);  // NaN, infinity, or a broken bound: never silently wrong
    atomicAdd(dst, static_cast<unsigned long long>(__float2ll_rn(v * p.scale)));  // RED.E.ADD.64
}

// One 512-body stage against this warp's 32 R i-bodies, every pair once.
template <int R, bool UNIFORM>
__device__ __forceinline__ void sym_stage(const float* __restrict__ tp, const int lane, const int warp, const f32x2 (&nx)[R],
                                          const f32x2 (&ny)[R], const f32x2 (&nz)[R], const f32x2 (&mi)[R], f32x2 (&ax)[R], f32x2 (&ay)[R],
                                          f32x2 (&az)[R], const f32x2 eps2, const unsigned stage, const Sym32Params& p, float4* __restrict__ slot)
{
#pragma unroll 1
    for (int gi = 0; gi < kPosTile / 64; ++gi) {
        const int grp = (gi + 2 * warp) & 7;  // the four warps start on different groups
        const float* px = tp + grp * 64;
        f32x2 gx = 0ull, gy = 0ull, gz = 0ull;  // sum over this warp's i-bodies of w * m_i * d for the couple this lane holds
#if NB_SYM32_SPLIT_G
        f32x2 hx = 0ull, hy = 0ull, hz = 0ull;  // (lab: the odd i-bodies' half of the chain)
#endif
        constexpr int kUnroll = UNIFORM ? (R == 8 ? NB_SYM32_UNROLL_EQUAL : 32) : NB_SYM32_UNROLL_GENERAL;
#if NB_SYM32_REPARK
        {
            const float2 X = *reinterpret_cast<const float2*>(px + 2 * lane);
            const float2 Y = *reinterpret_cast<const float2*>(px + kPosTile + 2 * lane);
            const float2 Z = *reinterpret_cast<const float2*>(px + 2 * kPosTile + 2 * lane);
            float2 M = make_float2(0.f, 0.f);
            if (!UNIFORM) M = *reinterpret_cast<const float2*>(px + 3 * kPosTile + 2 * lane);
            const float4 a = make_float4(X.x, X.y, Y.x, Y.y), b = make_float4(Z.x, Z.y, M.x, M.y);
            __syncwarp();  // (every lane is done with the previous group)
            slot[lane] = a;
            slot[lane + 32] = a;
            slot[lane + 64] = b;
            slot[lane + 96] = b;
            __syncwarp();
        }
        const float4* xy = slot + lane;
#endif
#pragma unroll kUnroll
        for (int q = 0; q < 32; ++q) {
#if NB_SYM32_REPARK
            const float4 XY = xy[q];
            const float4 ZM = xy[q + 64];
            const f32x2 jx = pack2(XY.x, XY.y), jy = pack2(XY.z, XY.w), jz = pack2(ZM.x, ZM.y);
            const f32x2 jm = UNIFORM ? 0ull : pack2(ZM.z, ZM.w);
#else
            const int o = 2 * ((lane + q) & 31);
            const float2 X = *reinterpret_cast<const float2*>(px + o);
            const float2 Y = *reinterpret_cast<const float2*>(px + kPosTile + o);
            const float2 Z = *reinterpret_cast<const float2*>(px + 2 * kPosTile + o);
            const f32x2 jx = pack2(X.x, X.y), jy = pack2(Y.x, Y.y), jz = pack2(Z.x, Z.y);
            f32x2 jm = 0ull;
            if (!UNIFORM) {
                const float2 M = *reinterpret_cast<const float2*>(px + 3 * kPosTile + o);
                jm = pack2(M.x, M.y);
            }
#endif
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
                const f32x2 wi = UNIFORM ? w : mul2(w, jm);     // f * m_j, idx:282
                const f32x2 wj = UNIFORM ? w : mul2(w, mi[r]);  // f * m_i, idx:286
                ax[r] = fma2(wi, dx, ax[r]);
                ay[r] = fma2(wi, dy, ay[r]);
                az[r] = fma2(wi, dz, az[r]);
#if NB_SYM32_SPLIT_G
                if (r & 1) {
                    hx = fma2(wj, dx, hx);
                    hy = fma2(wj, dy, hy);
                    hz = fma2(wj, dz, hz);
                } else
#endif
                {
                    gx = fma2(wj, dx, gx);
                    gy = fma2(wj, dy, gy);
                    gz = fma2(wj, dz, gz);
                }
            }
#if NB_SYM32_SPLIT_G
            gx = add2(gx, hx);
            gy = add2(gy, hy);
            gz = add2(gz, hz);
            hx = hy = hz = 0ull;
#endif
            // at step q + 1 this lane meets the couple that lane + 1 met at step q: its sums come along
            gx = __shfl_sync(0xffffffffu, gx, (lane + 1) & 31);
            gy = __shfl_sync(0xffffffffu, gy, (lane + 1) & 31);
            gz = __shfl_sync(0xffffffffu, gz, (lane + 1) & 31);
        }
        // 32 rotations: lane l holds the sums of couple l = bodies (l, l + 32) of the group again.
        // acc_j -= (f * m_i) * vec, idx:286-288
        NB_CHECK(size_t(stage) * kPosTile + grp * 64 + 2 * lane + 1 < p.n_pad);
        unsigned long long* dst = p.accfix + size_t(stage) * kPosTile + grp * 64 + 2 * lane;
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

template <int R, bool UNIFORM>
__device__ __forceinline__ void force_f32_sym_body(const Sym32Params& p)
{
    constexpr int kWarps = kF32Threads / 32;
    constexpr int kITile = kF32Threads * R;
    constexpr unsigned kBytes = (UNIFORM ? 3 : 4) * kPosTile * 4;  // the equal-mass kernel never reads the mass plane
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ __align__(16) float4 repark[kF32Threads / 32][NB_SYM32_REPARK ? 128 : 1];  // 2 KiB per compute warp
    uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + size_t(kSym32Stages) * kSym32StageBytes);
    uint64_t* empty = full + kSym32Stages;
    uint2* info = reinterpret_cast<uint2*>(empty + kSym32Stages);  // per stage: (i-tile, flags | position tile << 4)

    const int tid = threadIdx.x;
    const int warp = tid >> 5;
    const int lane = tid & 31;

    if (tid == 0) {
        for (int s = 0; s < kSym32Stages; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], kWarps);
        }
        fence_mbar_init();
    }
    __syncthreads();

    if (warp == kWarps) {  // ---- producer warp: one lane pulls units and drives the TMA ring ----
        if (lane != 0) return;
        unsigned t = 0;
        for (;;) {
            const unsigned u = p.u_begin + atomicAdd(p.sched, 1u);
            if (u >= p.u_end) break;
            const uint2 raw = __ldg(reinterpret_cast<const uint2*>(p.units) + u);
            const SymUnit un{raw.x, raw.y};
            const unsigned stage0 = un.run & 0xffffffu, count = (un.run >> 24) & 0x7fu;
            NB_CHECK(count >= 1 && size_t(stage0 + count) * kPosTile <= p.n_pad && size_t(un.tile + 1) * kITile <= p.n_pad);
            const unsigned diag = (un.run & kSymUnitDiag) ? kSymDiag : 0u;
            for (unsigned k = 0; k < count; ++k, ++t) {
                const unsigned slot = t % kSym32Stages;
                if (t >= unsigned(kSym32Stages)) mbar_wait(&empty[slot], ((t / kSym32Stages) - 1) & 1);
                info[slot] = make_uint2(un.tile, (k == 0 ? kSymFirst : 0u) | (k + 1 == count ? kSymLast : 0u) | diag | ((stage0 + k) << 4));
                mbar_expect_tx(&full[slot], kBytes);
                bulk_g2s(smem_raw + size_t(slot) * kSym32StageBytes, p.pos + size_t(stage0 + k) * (4 * kPosTile), kBytes, &full[slot]);
            }
        }
        const unsigned slot = t % kSym32Stages;
        if (t >= unsigned(kSym32Stages)) mbar_wait(&empty[slot], ((t / kSym32Stages) - 1) & 1);
        info[slot] = make_uint2(0u, kSymStop);
        mbar_arrive(&full[slot]);  // (release: the consumers' wait sees the stop mark)
        return;
    }

    // ---- compute warps ----
    f32x2 nx[R], ny[R], nz[R], mi[R], ax[R], ay[R], az[R];
    float nr[R];  // (consume_tile's nearest-neighbour slot: unused here)
    const f32x2 eps2 = pack2(p.eps2, p.eps2);
    size_t g0 = 0;  // this thread's body r = 0 of the current unit

    for (unsigned t = 0;; ++t) {
        const unsigned slot = t % kSym32Stages;
        mbar_wait(&full[slot], (t / kSym32Stages) & 1);
        const uint2 inf = info[slot];
        if (inf.y & kSymStop) break;
        NB_CHECK(size_t(inf.y >> 4) * kPosTile < p.n_pad);
        const float* tp = reinterpret_cast<const float*>(smem_raw + size_t(slot) * kSym32StageBytes);
        if (inf.y & kSymFirst) {
            g0 = size_t(inf.x) * kITile + tid;
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const size_t g = g0 + size_t(r) * kF32Threads;
                const float x = p.pos[pm_index<float>(g, 0)];
                const float y = p.pos[pm_index<float>(g, 1)];
                const float z = p.pos[pm_index<float>(g, 2)];
                nx[r] = pack2(-x, -x);
                ny[r] = pack2(-y, -y);
                nz[r] = pack2(-z, -z);
                if (!UNIFORM) {
                    const float m = p.pos[pm_index<float>(g, 3)];
                    mi[r] = pack2(m, m);
                } else {
                    mi[r] = 0ull;
                }
                ax[r] = ay[r] = az[r] = 0ull;
            }
        }
        if (inf.y & kSymDiag)  // the i-tile's own bodies: every ordered pair, i side only (the self pair is 0 * w: softening > 0)
            consume_tile<R, false, 4, false, false, UNIFORM, kPosTile / 2>(tp, nx, ny, nz, ax, ay, az, nr, eps2, 0);
        else
            sym_stage<R, UNIFORM>(tp, lane, warp, nx, ny, nz, mi, ax, ay, az, eps2, inf.y >> 4, p, repark[warp]);
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[slot]);
        if (inf.y & kSymLast) {
#pragma unroll
            for (int r = 0; r < R; ++r) {
                NB_CHECK(fix_slot(g0 + size_t(r) * kF32Threads) < p.n_pad);
                unsigned long long* dst = p.accfix + fix_slot(g0 + size_t(r) * kF32Threads);
                float lo, hi;
                unpack2(ax[r], lo, hi);
                sym_red(dst, lo + hi, p);
                unpack2(ay[r], lo, hi);
                sym_red(dst + p.n_pad, lo + hi, p);
                unpack2(az[r], lo, hi);
                sym_red(dst + 2 * p.n_pad, lo + hi, p);
            }
        }
    }
}

#ifndef NB_SYM32_CTAS_R4
#define NB_SYM32_CTAS_R4 3
#endif
#ifndef NB_SYM32_CTAS_R8
#define NB_SYM32_CTAS_R8 2
#endif
template <int R, bool UNIFORM>
__global__ void __launch_bounds__(kF32Threads + 32, R == 8 ? NB_SYM32_CTAS_R8 : NB_SYM32_CTAS_R4) force_f32_sym_kernel(const Sym32Params p)
{
    force_f32_sym_body<R, UNIFORM>(p);
}

// Where the sums of the bodies this GPU owns have been accumulated: its own array and, on several GPUs,
// every peer's (peer-mapped or IPC-opened pointers).
struct Sym32Sources {
    unsigned long long* acc[kMaxPeers];  // (each followed by its poison word at [3 n_pad])
    int n;
    double inv_scale;  // 2^-k
    // One GPU: no memsets between the kernels.  The LAST block of the finish kernel to retire (counted in `done`, which it
    // leaves at zero) rewinds the unit counter and clears the poison word: every thread of the pass has read it by then.
    unsigned* sched_rewind;
    unsigned* done;
};

// One thread per owned body: add the accumulators (exact), clear them for the next pass, convert, and run the
// integrator epilogue.  Bodies are walked in slot order so that the 64-bit loads are coalesced.
constexpr int kSym32FinishThreads = 256;
template <bool UNIFORM>
__global__ void __launch_bounds__(kSym32FinishThreads) f32_sym_finish_kernel(const ForceParams<float> p, const Sym32Sources src, const size_t n_pad)
{
    const size_t sl = size_t(blockIdx.x) * kSym32FinishThreads + threadIdx.x;  // local slot index
    const unsigned bad0 = src.sched_rewind ? unsigned(src.acc[0][3 * n_pad] != 0ull) : 0u;
    if (src.sched_rewind) {  // (one GPU) every thread of the block has the poison word; the last block to say so clears it
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence();
            if (atomicAdd(src.done, 1u) == gridDim.x - 1) {
                *src.done = 0u;
                *src.sched_rewind = 0u;
                src.acc[0][3 * n_pad] = 0ull;
            }
        }
    }
    if (sl >= p.n_local) return;
    const size_t il = (sl & ~size_t(511)) + tile_body(unsigned(sl) & 511u);  // (i_begin is a multiple of the tile)
    const size_t gs = p.i_begin + sl;
    NB_CHECK(gs < n_pad && il < p.n_local);
    long long s[3] = {0, 0, 0};
    unsigned bad = bad0;
    for (int k = 0; k < src.n; ++k) {
        unsigned long long* a = src.acc[k] + gs;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            s[c] += static_cast<long long>(a[c * n_pad]);
            a[c * n_pad] = 0ull;
        }
        if (!src.sched_rewind) bad |= unsigned(src.acc[k][3 * n_pad] != 0ull);
    }
    float sx = float(double(s[0]) * src.inv_scale);
    float sy = float(double(s[1]) * src.inv_scale);
    float sz = float(double(s[2]) * src.inv_scale);
    if (bad) sx = sy = sz = __int_as_float(0x7fc00000);
    if (UNIFORM && p.n_outliers > 0) {
        const size_t g = p.i_begin + il;
        outlier_correction<false>(p, p.pos_cur[pm_index<float>(g, 0)], p.pos_cur[pm_index<float>(g, 1)], p.pos_cur[pm_index<float>(g, 2)], sx, sy, sz);
    }
    finish_body<float>(p, il, sx, sy, sz, p.G);
}

}  // namespace nb
