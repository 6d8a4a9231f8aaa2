// fp64 all-pairs force kernel (reference-matching mode), same work decomposition
// and TMA ring as force_f32.cuh, scalar double arithmetic, one i-body per thread.
// The reference's statement (/root/reference/index.html:277-288) is
//   vec = pos_j - pos_i; distSq = (vx*vx + vy*vy) + vz*vz;
//   f = G / pow(distSq + epsSq, 1.5);  acc_i += (f*m_j)*vec
// and the gate is 1e-12 norm-relative per step (north_star), four orders above double
// rounding, so the kernel is free to pick the cheapest sequence that stays at the
// 1e-16 level per term.  MATH selects it:
//   0  IEEE, the reference's association, no contraction: pow(s,1.5) = s*sqrt(s) with
//      IEEE sqrt and division (<= ~1.5 ulp per term).  33 DP-pipe instructions + 2 MUFU
//      per interaction, two slow-path branches.  NBODY_F64_MATH=0 selects it.
//   1  1/pow(s,1.5) = y^3 with y = rsqrt(s): MUFU.RSQ64H seed (2^-26) and one cubically
//      convergent correction (5 DP ops, y to ~1 ulp); G applied once per body.  Everything
//      else as in 0.  23 DP instructions + 1 MUFU, no branches.
//   2  (default) the multiply-adds of distSq and of the accumulation contracted (DFMA: one
//      rounding instead of two) and s^-1.5 taken directly from the seed (inv_cube_f64: 6 DP ops
//      instead of 5 + 2, and closer to the exact value).  16 DP instructions + 1 MUFU.
// Per-step error against the oracle, all three: ~3e-16 (tests/test_gpu_parity.py prints it).
// For body i with j < i the reference accumulates -(f*m_j)*(pos_i - pos_j),
// which is bit-identical to +(f*m_j)*(pos_j - pos_i) (negation is exact), so a
// thread that walks j upward reproduces the reference's summation order when
// the j range is not split (j_chunks == 1).
#pragma once
#include "common.cuh"
#include "epilogue.cuh"
#include "tma.cuh"

namespace nb {

constexpr size_t force_f64_smem_bytes() { return size_t(kStages) * kTileBytes + 2 * kStages * 8; }

// rsqrt(s) to ~1 ulp: e = 1 - s*y0^2 is ~2^-25 after the MUFU seed, and
// y0 + y0*e*(1/2 + 3/8 e) leaves an O(e^3) term far below the final rounding.
__device__ __forceinline__ double rsqrt_f64(double s)
{
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(s));  // MUFU.RSQ64H
    const double e = __fma_rn(s, -__dmul_rn(y, y), 1.0);
    const double c = __fma_rn(e, 0.375, 0.5);
    return __fma_rn(c, __dmul_rn(y, e), y);
}

// s^-1.5 to ~2 ulp in 6 DP instructions: with c = y0^2 and e = 1 - s*c,
//   s^-1.5 = y0^3 * (1 - e)^-1.5 = y0^3 * (1 + e*(3/2 + 15/8 e) + O(e^3)).
// Measured on the host against quad precision with a seed truncated to the high word (2^-21, worse
// than MUFU.RSQ64H's): max 2.1 x 2^-53 relative, against 4.9 x 2^-53 for rsqrt_f64(s)^3 (7 instructions).
__device__ __forceinline__ double inv_cube_f64(double s)
{
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(s));  // MUFU.RSQ64H
    const double c = __dmul_rn(y, y);
    const double e = __fma_rn(s, -c, 1.0);
    const double c3 = __dmul_rn(c, y);
    const double p = __fma_rn(e, 1.875, 1.5);
    return __fma_rn(c3, __dmul_rn(p, e), c3);
}

#ifndef NB_F64_UNROLL
#define NB_F64_UNROLL 4  // N=65,536: unroll 1 / 2 / 4 = 830 / 882 / 907 G interactions/s (profiles/r1e_f64_variants.jsonl)
#endif
constexpr int kF64Unroll = NB_F64_UNROLL;
#ifndef NB_F64_R
#define NB_F64_R 1
#endif
constexpr int kF64BodiesPerThread = NB_F64_R;  // default i-bodies per thread (NBODY_F64_R overrides at create)

// ZERO_SOFT: softening == 0, where the self pair would be 0 * inf and is selected away (the
// reference never evaluates it, idx:273).  With softening it contributes an exact zero by
// itself and the select is not compiled in.
// UNIFORM (MATH 2 only): every mass is the same value (config 2's Plummer sphere: 1/N), the
// per-pair multiply by m_j drops out -- 15 DP instructions -- and the common mass joins G; pads sit
// at x = 1e150, where y^3 underflows to an exact 0.  R: i-bodies per thread (more independent chains).
template <int MATH, bool ZERO_SOFT, bool UNIFORM = false, int R = 1>
__global__ void __launch_bounds__(kThreads + 32) force_f64_kernel(const ForceParams<double> p)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    double4* tiles = reinterpret_cast<double4*>(smem_raw);
    uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + size_t(kStages) * kTileBytes);
    uint64_t* empty = full + kStages;
    __shared__ int s_last;

    const int tid = threadIdx.x;
    const int warp = tid >> 5;
    const int lane = tid & 31;
    const int chunk = blockIdx.x % p.j_chunks;
    const size_t tile_i = blockIdx.x / p.j_chunks;

    const size_t j0 = size_t(chunk) * p.chunk_len;
    const size_t j1 = (j0 + p.chunk_len < p.n_j) ? j0 + p.chunk_len : p.n_j;
    const size_t chunk_bytes = (j1 - j0) * 32;
    const int ntiles = static_cast<int>((chunk_bytes + kTileBytes - 1) / kTileBytes);
    const unsigned char* src = reinterpret_cast<const unsigned char*>(p.pos_cur) + j0 * 32;

    if (tid == 0) {
        for (int s = 0; s < kStages; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], kThreads / 32);
        }
        fence_mbar_init();
    }
    __syncthreads();

    if (warp == kThreads / 32) {  // producer warp
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

    const size_t il0 = tile_i * size_t(kThreads * R) + tid;  // body r of this thread: il0 + r * kThreads
    double4 me[R];
    double ax[R], ay[R], az[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        me[r] = reinterpret_cast<const double4*>(p.pos_cur)[p.i_begin + il0 + size_t(r) * kThreads];
        ax[r] = ay[r] = az[r] = 0.0;
    }

    for (int t = 0; t < ntiles; ++t) {
        const int slot = t % kStages;
        mbar_wait(&full[slot], (t / kStages) & 1);
        const double4* tp = tiles + size_t(slot) * (kTileBytes / 32);
        const size_t off = size_t(t) * kTileBytes;
        const int nj = static_cast<int>((chunk_bytes - off < kTileBytes ? chunk_bytes - off : kTileBytes) / 32);
        int self_q[R];  // this body's slot in the tile, if it is there
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const long long d = static_cast<long long>(p.i_begin + il0 + size_t(r) * kThreads) - static_cast<long long>(j0 + off / 32);
            self_q[r] = (d >= 0 && d < nj) ? static_cast<int>(d) : -1;
        }
#pragma unroll kF64Unroll
        for (int q = 0; q < nj; ++q) {
            const double4 b = tp[q];
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const double vx = __dsub_rn(b.x, me[r].x);
                const double vy = __dsub_rn(b.y, me[r].y);
                const double vz = __dsub_rn(b.z, me[r].z);
                double s;
                if (MATH == 2) {
                    s = __fma_rn(vz, vz, __fma_rn(vy, vy, __fma_rn(vx, vx, p.eps2)));
                } else {
                    const double distSq = __dadd_rn(__dadd_rn(__dmul_rn(vx, vx), __dmul_rn(vy, vy)), __dmul_rn(vz, vz));
                    s = __dadd_rn(distSq, p.eps2);
                }
                double fm;
                if (MATH == 0) {
                    fm = __dmul_rn(__ddiv_rn(p.G, __dmul_rn(s, __dsqrt_rn(s))), b.w);
                } else {
                    if (MATH == 2) {
                        fm = inv_cube_f64(s);  // G (and the common mass) join in the epilogue
                    } else {
                        const double y = rsqrt_f64(s);
                        fm = __dmul_rn(__dmul_rn(y, y), y);
                    }
                    if (!UNIFORM) fm = __dmul_rn(fm, b.w);
                }
                if (ZERO_SOFT && q == self_q[r]) fm = 0.0;  // the reference never evaluates the self pair (idx:273)
                if (MATH == 2) {
                    ax[r] = __fma_rn(fm, vx, ax[r]);
                    ay[r] = __fma_rn(fm, vy, ay[r]);
                    az[r] = __fma_rn(fm, vz, az[r]);
                } else {
                    ax[r] = __dadd_rn(ax[r], __dmul_rn(fm, vx));
                    ay[r] = __dadd_rn(ay[r], __dmul_rn(fm, vy));
                    az[r] = __dadd_rn(az[r], __dmul_rn(fm, vz));
                }
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[slot]);
    }

    const double scale = MATH == 0 ? 1.0 : p.G;  // UNIFORM: the host folded the common mass into p.G
    if (p.j_chunks == 1) {
#pragma unroll
        for (int r = 0; r < R; ++r) finish_body<double>(p, il0 + size_t(r) * kThreads, ax[r], ay[r], az[r], scale);
        return;
    }
    double4* part = reinterpret_cast<double4*>(p.partial);
    NB_CHECK(chunk < p.j_chunks && il0 + size_t(R - 1) * kThreads < p.n_local);
#pragma unroll
    for (int r = 0; r < R; ++r) part[size_t(chunk) * p.n_local + il0 + size_t(r) * kThreads] = make_double4(ax[r], ay[r], az[r], 0.0);
    if (!tile_arrive_last(&p.tile_counter[tile_i], p.j_chunks, &s_last, kThreads)) return;
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const size_t il = il0 + size_t(r) * kThreads;
        double tx = 0.0, ty = 0.0, tz = 0.0;
        for (int c = 0; c < p.j_chunks; ++c) {
            const double2* q = reinterpret_cast<const double2*>(&part[size_t(c) * p.n_local + il]);
            const double2 xy = __ldcg(q);
            const double2 zw = __ldcg(q + 1);
            tx = __dadd_rn(tx, xy.x);
            ty = __dadd_rn(ty, xy.y);
            tz = __dadd_rn(tz, zw.x);
        }
        finish_body<double>(p, il, tx, ty, tz, scale);
    }
    if (tid == 0) p.tile_counter[tile_i] = 0;
}

}  // namespace nb
