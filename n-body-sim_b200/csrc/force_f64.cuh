// fp64 all-pairs force kernel (reference-matching mode), same work decomposition
// and TMA ring as force_f32.cuh but scalar IEEE double arithmetic in the
// reference's own association (/root/reference/index.html:277-288):
//   vec = pos_j - pos_i; distSq = (vx*vx + vy*vy) + vz*vz;
//   f = G / pow(distSq + epsSq, 1.5);  acc_i += (f*m_j)*vec
// with no multiply-add contraction (JS has none).  pow(s,1.5) is evaluated as
// s*sqrt(s) with IEEE sqrt and division: <= ~1.5 ulp from the correctly
// rounded value, i.e. ~2e-16 relative per term against a 1e-12 gate.
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

    const size_t il = tile_i * size_t(kThreads) + tid;
    const size_t gi = p.i_begin + il;
    const double4 me = reinterpret_cast<const double4*>(p.pos_cur)[gi];
    double ax = 0.0, ay = 0.0, az = 0.0;

    for (int t = 0; t < ntiles; ++t) {
        const int slot = t % kStages;
        mbar_wait(&full[slot], (t / kStages) & 1);
        const double4* tp = tiles + size_t(slot) * (kTileBytes / 32);
        const size_t off = size_t(t) * kTileBytes;
        const int nj = static_cast<int>((chunk_bytes - off < kTileBytes ? chunk_bytes - off : kTileBytes) / 32);
        const size_t jg0 = j0 + off / 32;
#pragma unroll 2
        for (int q = 0; q < nj; ++q) {
            const double4 b = tp[q];
            const double vx = __dsub_rn(b.x, me.x);
            const double vy = __dsub_rn(b.y, me.y);
            const double vz = __dsub_rn(b.z, me.z);
            const double distSq =
                __dadd_rn(__dadd_rn(__dmul_rn(vx, vx), __dmul_rn(vy, vy)), __dmul_rn(vz, vz));
            const double s = __dadd_rn(distSq, p.eps2);
            double fm = __dmul_rn(__ddiv_rn(p.G, __dmul_rn(s, __dsqrt_rn(s))), b.w);
            if (jg0 + q == gi) fm = 0.0;  // the reference never evaluates the self pair (idx:273)
            ax = __dadd_rn(ax, __dmul_rn(fm, vx));
            ay = __dadd_rn(ay, __dmul_rn(fm, vy));
            az = __dadd_rn(az, __dmul_rn(fm, vz));
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[slot]);
    }

    if (p.j_chunks == 1) {
        finish_body<double>(p, il, ax, ay, az, 1.0);
        return;
    }
    double4* part = reinterpret_cast<double4*>(p.partial);
    part[size_t(chunk) * p.n_local + il] = make_double4(ax, ay, az, 0.0);
    if (!tile_arrive_last(&p.tile_counter[tile_i], p.j_chunks, &s_last, kThreads)) return;
    double tx = 0.0, ty = 0.0, tz = 0.0;
    for (int c = 0; c < p.j_chunks; ++c) {
        const double2* q = reinterpret_cast<const double2*>(&part[size_t(c) * p.n_local + il]);
        const double2 xy = __ldcg(q);
        const double2 zw = __ldcg(q + 1);
        tx = __dadd_rn(tx, xy.x);
        ty = __dadd_rn(ty, xy.y);
        tz = __dadd_rn(tz, zw.x);
    }
    finish_body<double>(p, il, tx, ty, tz, 1.0);
    if (tid == 0) p.tile_counter[tile_i] = 0;
}

}  // namespace nb
