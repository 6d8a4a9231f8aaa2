// fp64 force pass that evaluates every pair ONCE (one GPU, MATH 2 arithmetic).
//
// The reference itself walks i < j and applies each pair to both bodies
// (/root/reference/index.html:272-288: acc_i += (f*m_j)*vec, acc_j -= (f*m_i)*vec).  On the
// fp32 path that sharing buys nothing; in fp64 the kernel is bound by the
// FP64 pipe (force_f64.cuh: 15-16 DP instructions per one-sided interaction), and the
// expensive part of a pair -- the three subtractions, distSq and s^-1.5, 12 of those
// instructions -- is the same for both bodies.  Evaluated once, a pair costs 18 DP instructions
// (20 with general masses) for TWO interactions: 9-10 per interaction instead of 15-16.
//
// Decomposition.  Bodies are cut into i-blocks of 512 (R = 4 per thread, 128 threads) and j-blocks
// of 128.  CTA (I, c) takes i-block I against the c-th run of `k` j-blocks at or above I's
// own; j-blocks below I belong to the rows of lower i-blocks (they reach I through the
// j-side sums).  A j-block inside I's own range is walked the ordinary one-sided way (every
// ordered pair of the 512 x 128 square, i-side only).  For a j-block above I the four warps
// take its four 32-body segments in rotation; inside a segment lane l meets body (l + q) mod
// 32 at step q, so every lane of a warp works on a different j, and the j-side sum of a body
// travels with it from lane to lane (three 64-bit shuffles per step of 4 pairs) -- no atomics,
// no shared-memory read-modify-write inside the loop.  After 32 steps the sums are back in
// their home lanes and are added to the block's shared-memory totals, which the warp owns
// exclusively during that phase.
//
// Results leave the kernel as partial sums in a fixed layout -- part_i[c][body] from CTA
// (I, c), part_j[I][body] for every body above i-block I -- and f64_sym_finish_kernel adds
// them per body in one fixed order (lower i-blocks first, then the body's own row) and runs
// the integrator epilogue.  Nothing depends on CTA scheduling: two runs agree bit for bit.
// Every term is the value force_f64_kernel<2> computes for that ordered pair (d_ji = -d_ij
// and the negation are exact); only the order of the additions differs.
#pragma once
#include "common.cuh"
#include "epilogue.cuh"
#include "force_f64.cuh"

namespace nb {

#ifndef NB_SYM_UNROLL
#define NB_SYM_UNROLL 2
#endif
constexpr int kSymUnroll = NB_SYM_UNROLL;            // steps per inner-loop iteration
constexpr int kSymThreads = 128;
constexpr int kSymMaxR = 4;                       // i-bodies per thread: 4 (2 and 1 for small systems: more CTAs)
constexpr int kSymIBlock = kSymThreads * kSymMaxR;  // 512: the padding granularity whatever R is
constexpr int kSymJBlock = kSymThreads;           // 128: one 32-body segment per warp and phase
constexpr int kSymFinishWarps = 8;

struct SymParams {
    const double* pos;   // double4 (x, y, z, m) per body, n_pad bodies; pads are far away and massless
    double* part_i;      // [chunks_max][3][n_pad]
    double* part_j;      // [n_iblocks][3][n_pad]; row I is written for bodies above i-block I only
    size_t n_pad;        // multiple of kSymIBlock
    int r;               // i-bodies per thread: an i-block is r j-blocks long
    int n_iblocks;       // n_pad / (kSymThreads * r)
    int k;               // j-blocks per CTA
    double eps2;
    // FIXED: the sums leave as exact fixed-point contributions instead (round 2; several GPUs, any n).  Every partial sum v is
    // split into hi = rint(v 2^k1) and lo = rint((v - hi 2^-k1) 2^(k1+50)) -- both exact -- and added to the body's two 64-bit
    // accumulators with RED.ADD.64: integer sums do not depend on the order or the GPU the contributions come from.
    int i_block_begin;           // first i-block (row) of this device's share; gridDim.y rows are launched
    unsigned long long* acc_hi;  // [3][n_pad] + the poison word at [3 n_pad]
    unsigned long long* acc_lo;  // [3][n_pad]
    double scale_hi, inv_scale_hi, scale_lo;  // 2^k1, 2^-k1, 2^(k1+50); |sum| <= bound maps to 2^61
    double bound;
};

__device__ __forceinline__ void sym64_red(const SymParams& p, size_t idx, double v)
{
    if (!(fabs(v) <= p.bound)) atomicOr(p.acc_hi + 3 * p.n_pad, 1ull);  // NaN, infinity or a broken bound poisons the step
    const long long h = __double2ll_rn(v * p.scale_hi);
    const double r = v - double(h) * p.inv_scale_hi;  // exact: |r| <= 2^-k1-1
    const long long l = __double2ll_rn(r * p.scale_lo);
    atomicAdd(p.acc_hi + idx, static_cast<unsigned long long>(h));
    atomicAdd(p.acc_lo + idx, static_cast<unsigned long long>(l));
}

__host__ __device__ inline int sym_chunks_of_row(int n_iblocks, int r, int k, int I)
{
    return ((n_iblocks - I) * r + k - 1) / k;
}

// (distSq + epsSq)^-1.5 of one pair, MATH 2's arithmetic (force_f64.cuh)
__device__ __forceinline__ double sym_pair_weight(double dx, double dy, double dz, double eps2)
{
    return inv_cube_f64(__fma_rn(dz, dz, __fma_rn(dy, dy, __fma_rn(dx, dx, eps2))));
}

template <bool ZERO_SOFT, bool UNIFORM, int kSymR, bool FIXED = false>
__global__ void __launch_bounds__(kSymThreads) force_f64_sym_kernel(const SymParams p)
{
    constexpr int kSymJPerI = kSymR;
    constexpr int kIBlock = kSymThreads * kSymR;
    __shared__ double sx[kSymJBlock], sy[kSymJBlock], sz[kSymJBlock], sm[kSymJBlock];
    __shared__ double fx[kSymJBlock], fy[kSymJBlock], fz[kSymJBlock];

    const int I = int(blockIdx.y) + (FIXED ? p.i_block_begin : 0), c = blockIdx.x;
    if (c >= sym_chunks_of_row(p.n_iblocks, kSymR, p.k, I)) return;
    const int jb0 = I * kSymJPerI + c * p.k;
    const int jb_end = p.n_iblocks * kSymJPerI;
    const int jb1 = jb0 + p.k < jb_end ? jb0 + p.k : jb_end;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const double4* pos4 = reinterpret_cast<const double4*>(p.pos);
    const size_t i0 = size_t(I) * kIBlock + tid;  // body r of this thread: i0 + r * kSymThreads

    double xi[kSymR], yi[kSymR], zi[kSymR], mi[kSymR], ax[kSymR], ay[kSymR], az[kSymR];
#pragma unroll
    for (int r = 0; r < kSymR; ++r) {
        const double4 b = pos4[i0 + size_t(r) * kSymThreads];
        xi[r] = b.x;
        yi[r] = b.y;
        zi[r] = b.z;
        mi[r] = b.w;
        ax[r] = ay[r] = az[r] = 0.0;
    }

    for (int jb = jb0; jb < jb1; ++jb) {
        __syncthreads();  // the previous block's positions and totals are no longer read
        {
            const double4 b = pos4[size_t(jb) * kSymJBlock + tid];
            sx[tid] = b.x;
            sy[tid] = b.y;
            sz[tid] = b.z;
            sm[tid] = b.w;
            fx[tid] = fy[tid] = fz[tid] = 0.0;
        }
        __syncthreads();

        if (jb / kSymJPerI == I) {
            // inside the i-block's own range: every ordered pair, i-side only (force_f64_kernel's loop)
            const int self0 = tid - (jb - I * kSymJPerI) * kSymJBlock;  // slot of body r in this block: self0 + r * kSymThreads
#pragma unroll kSymUnroll
            for (int q = 0; q < kSymJBlock; ++q) {
                const double xj = sx[q], yj = sy[q], zj = sz[q];
                const double mj = UNIFORM ? 0.0 : sm[q];
#pragma unroll
                for (int r = 0; r < kSymR; ++r) {
                    const double dx = __dsub_rn(xj, xi[r]), dy = __dsub_rn(yj, yi[r]), dz = __dsub_rn(zj, zi[r]);
                    double w = sym_pair_weight(dx, dy, dz, p.eps2);
                    if (!UNIFORM) w = __dmul_rn(w, mj);
                    if (ZERO_SOFT && q == self0 + r * kSymThreads) w = 0.0;  // the reference never evaluates the self pair (idx:273)
                    ax[r] = __fma_rn(w, dx, ax[r]);
                    ay[r] = __fma_rn(w, dy, ay[r]);
                    az[r] = __fma_rn(w, dz, az[r]);
                }
            }
            continue;
        }

        for (int ph = 0; ph < 4; ++ph) {
            const int seg = ((warp + ph) & 3) * 32;
            double gx = 0.0, gy = 0.0, gz = 0.0;  // sum over this warp's i-bodies of w * m_i * d for the j this lane holds
#pragma unroll kSymUnroll
            for (int q = 0; q < 32; ++q) {
                const int j = seg + ((lane + q) & 31);
                const double xj = sx[j], yj = sy[j], zj = sz[j];
                const double mj = UNIFORM ? 0.0 : sm[j];
#pragma unroll
                for (int r = 0; r < kSymR; ++r) {
                    const double dx = __dsub_rn(xj, xi[r]), dy = __dsub_rn(yj, yi[r]), dz = __dsub_rn(zj, zi[r]);
                    const double w = sym_pair_weight(dx, dy, dz, p.eps2);
                    const double wi = UNIFORM ? w : __dmul_rn(w, mj);
                    const double wj = UNIFORM ? w : __dmul_rn(w, mi[r]);
                    ax[r] = __fma_rn(wi, dx, ax[r]);
                    ay[r] = __fma_rn(wi, dy, ay[r]);
                    az[r] = __fma_rn(wi, dz, az[r]);
                    gx = __fma_rn(wj, dx, gx);
                    gy = __fma_rn(wj, dy, gy);
                    gz = __fma_rn(wj, dz, gz);
                }
                // at step q + 1 this lane meets the body that lane + 1 met at step q: its sum comes along
                gx = __shfl_sync(0xffffffffu, gx, (lane + 1) & 31);
                gy = __shfl_sync(0xffffffffu, gy, (lane + 1) & 31);
                gz = __shfl_sync(0xffffffffu, gz, (lane + 1) & 31);
            }
            // 32 rotations: lane l holds the sum of body seg + l again; the segment is this warp's alone in this phase
            fx[seg + lane] = __dadd_rn(fx[seg + lane], gx);
            fy[seg + lane] = __dadd_rn(fy[seg + lane], gy);
            fz[seg + lane] = __dadd_rn(fz[seg + lane], gz);
            __syncthreads();
        }
        // acc_j -= (f*m_i)*vec, idx:286-288
        if (FIXED) {
            const size_t at = size_t(jb) * kSymJBlock + tid;
            sym64_red(p, at, -fx[tid]);
            sym64_red(p, p.n_pad + at, -fy[tid]);
            sym64_red(p, 2 * p.n_pad + at, -fz[tid]);
        } else {
            double* row = p.part_j + size_t(I) * 3 * p.n_pad + size_t(jb) * kSymJBlock + tid;
            row[0] = -fx[tid];
            row[p.n_pad] = -fy[tid];
            row[2 * p.n_pad] = -fz[tid];
        }
    }

    if (FIXED) {
#pragma unroll
        for (int r = 0; r < kSymR; ++r) {
            const size_t at = i0 + size_t(r) * kSymThreads;
            sym64_red(p, at, ax[r]);
            sym64_red(p, p.n_pad + at, ay[r]);
            sym64_red(p, 2 * p.n_pad + at, az[r]);
        }
        return;
    }
    double* row = p.part_i + size_t(c) * 3 * p.n_pad + i0;
#pragma unroll
    for (int r = 0; r < kSymR; ++r) {
        row[size_t(r) * kSymThreads] = ax[r];
        row[p.n_pad + size_t(r) * kSymThreads] = ay[r];
        row[2 * p.n_pad + size_t(r) * kSymThreads] = az[r];
    }
}

// FIXED: one thread per owned body adds the accumulators of every rank (exact), clears them, converts
// (hi 2^-k1 + lo 2^-(k1+50): one rounding each, a fixed formula of integers) and runs the integrator epilogue.
struct Sym64Sources {
    unsigned long long* hi[kMaxPeers];
    unsigned long long* lo[kMaxPeers];
    int n;
    double inv_scale_hi, inv_scale_lo;
};
__global__ void __launch_bounds__(256) f64_sym_fixed_finish_kernel(const ForceParams<double> p, const Sym64Sources src, const size_t n_pad, double scale)
{
    const size_t il = size_t(blockIdx.x) * 256 + threadIdx.x;
    if (il >= p.n_local) return;
    const size_t g = p.i_begin + il;
    long long h[3] = {0, 0, 0}, l[3] = {0, 0, 0};
    bool bad = false;
    for (int k = 0; k < src.n; ++k) {
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            h[c] += static_cast<long long>(src.hi[k][c * n_pad + g]);
            l[c] += static_cast<long long>(src.lo[k][c * n_pad + g]);
            src.hi[k][c * n_pad + g] = 0ull;
            src.lo[k][c * n_pad + g] = 0ull;
        }
        bad = bad || src.hi[k][3 * n_pad] != 0ull;
    }
    double v[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) v[c] = __fma_rn(double(l[c]), src.inv_scale_lo, __dmul_rn(double(h[c]), src.inv_scale_hi));
    if (bad) v[0] = v[1] = v[2] = __longlong_as_double(0x7ff8000000000000LL);
    finish_body<double>(p, il, v[0], v[1], v[2], scale);
}

// One CTA finishes 32 bodies: warp w adds the slots w, w + 8, ... of each body (rows of lower
// i-blocks first, then the chunks of the body's own row), the eight sums are added in warp order,
// and the integrator epilogue follows (epilogue.cuh).  A fixed order per body.
__global__ void __launch_bounds__(kSymFinishWarps * 32) f64_sym_finish_kernel(const ForceParams<double> p, const SymParams s, double scale)
{
    __shared__ double red[kSymFinishWarps][3][32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const size_t b = size_t(blockIdx.x) * 32 + lane;
    const int I = int(b / (size_t(kSymThreads) * s.r));  // the same for all 32 bodies
    const int nj = I, ni = sym_chunks_of_row(s.n_iblocks, s.r, s.k, I);
    double tx = 0.0, ty = 0.0, tz = 0.0;
#pragma unroll 4
    for (int k = w; k < nj + ni; k += kSymFinishWarps) {
        const double* row = (k < nj ? s.part_j + size_t(k) * 3 * s.n_pad : s.part_i + size_t(k - nj) * 3 * s.n_pad) + b;
        tx = __dadd_rn(tx, row[0]);
        ty = __dadd_rn(ty, row[s.n_pad]);
        tz = __dadd_rn(tz, row[2 * s.n_pad]);
    }
    red[w][0][lane] = tx;
    red[w][1][lane] = ty;
    red[w][2][lane] = tz;
    __syncthreads();
    if (w != 0) return;
#pragma unroll
    for (int k = 1; k < kSymFinishWarps; ++k) {
        tx = __dadd_rn(tx, red[k][0][lane]);
        ty = __dadd_rn(ty, red[k][1][lane]);
        tz = __dadd_rn(tz, red[k][2][lane]);
    }
    finish_body<double>(p, b, tx, ty, tz, scale);
}

}  // namespace nb
