// Broad phase of PhysicsSystem.resolveCollisions (/root/reference/index.html:301-313):
// find every pair i<j whose spheres overlap, distSq < (r_i + r_j)^2, on the device.
// SURVEY.md 8(f) row 1.  The order-dependent response (idx:314-389) stays on the host.
//
// Two stages, so that the O(N^2) overlap search costs (almost) nothing on top of the
// force sweep that already visits every pair:
//   1. force_f32_kernel<.., NEAR=true> keeps, per i-body, the largest 1/sqrt(d^2+eps^2)
//      over all j != i (one FMNMX on the ALU pipe per j-pair, off the FMA pipe that
//      bounds the kernel).  flag_near_kernel turns it into a candidate list: body i is
//      a candidate iff its nearest neighbour could touch it even if that neighbour had
//      the largest radius in the scene (conservative; slack covers fp32 and MUFU error).
//   2. contact_pairs_kernel tests the candidates (or, without stage 1, every body)
//      against all j > i.  The test that decides is evaluated in fp64 in the reference's
//      association -- (vx*vx + vy*vy) + vz*vz < minDist*minDist -- on the state as stored,
//      so the pair SET is exactly the reference's for those positions; in fp32 contexts
//      an fp32 pre-test with slack keeps the fp64 work to near-contacts only.
// Pairs are appended with an atomic counter; the host sorts them into the reference's
// visiting order (ascending i, then ascending j).
#pragma once
#include "common.cuh"

namespace nb {

constexpr int kContactThreads = 128;

// near[il]   : max over j != i of rsqrt(d^2 + eps2) from the force kernel (local index)
// radius[g]  : Body.radius (idx:221), global index, replicated on every device
// cand[]     : out, global indices of candidate bodies (unordered)
// Bodies with radius > r_split ("big": a few dominant bodies, e.g. a galaxy's central mass)
// are left out here and out of `reach`; the host scans them against everybody instead, so
// one huge radius does not make every body a candidate.
__global__ void flag_near_kernel(const float* near, const float* radius, size_t i_begin, size_t n_local_real,
                                 float reach /* largest non-big radius + margin */, float r_split, float eps2,
                                 unsigned* cand, unsigned* n_cand)
{
    const size_t il = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (il >= n_local_real) return;
    const float nr = near[il];
    const float ri = radius[i_begin + il];
    if (ri > r_split) return;
    const float D = ri + reach;
    // touching possible  <=>  d^2 < D^2  <=>  near^2 * (D^2 + eps2) > 1   (1e-5 slack >> fp32 + MUFU error)
    if (nr * nr * (D * D + eps2) * 1.00001f >= 1.0f) {
        const unsigned k = atomicAdd(n_cand, 1u);
        cand[k] = static_cast<unsigned>(i_begin + il);
    }
}

__device__ __forceinline__ bool overlap_exact(double xi, double yi, double zi, double ri, double xj, double yj,
                                              double zj, double rj, double margin)
{
    // vec.subVectors(B.pos, A.pos); distSq = vec.lengthSq(); minDist = A.radius + B.radius   (idx:309-311)
    const double vx = __dsub_rn(xj, xi), vy = __dsub_rn(yj, yi), vz = __dsub_rn(zj, zi);
    const double distSq = __dadd_rn(__dadd_rn(__dmul_rn(vx, vx), __dmul_rn(vy, vy)), __dmul_rn(vz, vz));
    const double minDist = __dadd_rn(__dadd_rn(ri, rj), margin);
    return distSq < __dmul_rn(minDist, minDist);  // idx:313
}

// grid = (candidate blocks, j-chunks).  cand == nullptr: candidates are i_begin .. i_begin+n_cand-1.
// BOTH: test every j != i and emit (min, max) -- used for the big bodies, whose partners
// with a smaller index need not be candidates themselves (the host removes duplicates).
// Each thread owns kContactPerThread candidates in registers and the j-tile is one 16/32-byte
// (x, y, z, r) record per body, so a j costs one shared-memory load per kContactPerThread tests.
constexpr int kContactPerThread = 4;

template <typename Real, bool BOTH>
__global__ void __launch_bounds__(kContactThreads)
    contact_pairs_kernel(const Real* pos, const Real* radius, const unsigned* cand, unsigned n_cand, size_t i_begin,
                         size_t n, size_t chunk_len, double margin, uint2* pairs, unsigned long long cap,
                         unsigned long long* count)
{
    using V4 = typename Vec4<Real>::type;
    constexpr int C = kContactPerThread;
    __shared__ V4 tile[kContactThreads];
    const int tid = threadIdx.x;
    size_t ci[C];
    Real xi[C], yi[C], zi[C], ri[C];
    size_t i_min = ~size_t(0);
#pragma unroll
    for (int c = 0; c < C; ++c) {
        const unsigned k = (blockIdx.x * C + c) * kContactThreads + tid;
        if (k < n_cand) {
            ci[c] = cand ? size_t(cand[k]) : i_begin + k;
            xi[c] = pos[pm_index<Real>(ci[c], 0)];
            yi[c] = pos[pm_index<Real>(ci[c], 1)];
            zi[c] = pos[pm_index<Real>(ci[c], 2)];
            ri[c] = radius[ci[c]];
            i_min = ci[c] < i_min ? ci[c] : i_min;
        } else {  // no candidate: parked so far away that every pre-test fails (d2 = inf)
            ci[c] = ~size_t(0);
            xi[c] = Real(1e30);
            yi[c] = zi[c] = ri[c] = Real(0);
        }
    }
    // fp32 pre-test: d2 <= ((r_i + margin) * k + r_j * k)^2 with k^2 = 1 + 1e-5 (slack for fp32 rounding), 9 instructions
    const float kSlack = 1.000005f;
    float rik[C];
#pragma unroll
    for (int c = 0; c < C; ++c) rik[c] = (float(ri[c]) + float(margin)) * kSlack;
    const size_t j0 = size_t(blockIdx.y) * chunk_len;
    const size_t j1 = (j0 + chunk_len < n) ? j0 + chunk_len : n;
    for (size_t jb = j0; jb < j1; jb += kContactThreads) {
        __syncthreads();
        if (jb + tid < j1) {
            const size_t j = jb + tid;
            V4 v;
            v.x = pos[pm_index<Real>(j, 0)];
            v.y = pos[pm_index<Real>(j, 1)];
            v.z = pos[pm_index<Real>(j, 2)];
            v.w = radius[j];
            tile[tid] = v;
        }
        __syncthreads();
        if (!BOTH && (i_min == ~size_t(0) || jb + kContactThreads <= i_min + 1)) continue;  // the tile holds no j > i for this thread
        const int cnt = static_cast<int>(j1 - jb < size_t(kContactThreads) ? j1 - jb : size_t(kContactThreads));
#pragma unroll 4
        for (int q = 0; q < cnt; ++q) {
            const V4 b = tile[q];
#pragma unroll
            for (int c = 0; c < C; ++c) {
                bool hit;
                if (sizeof(Real) == 4) {
                    const float dx = float(b.x) - float(xi[c]), dy = float(b.y) - float(yi[c]), dz = float(b.z) - float(zi[c]);
                    const float d2 = fmaf(dz, dz, fmaf(dy, dy, dx * dx));
                    const float md = fmaf(float(b.w), kSlack, rik[c]);
                    hit = d2 <= md * md;
                } else {
                    hit = ci[c] != ~size_t(0);
                }
                if (!hit) continue;
                const size_t i = ci[c], j = jb + q;  // rare from here on
                if (BOTH ? j == i : j <= i) continue;
                // A = the lower index, B = the higher, as the reference's i<j loop has them (idx:306-309)
                const bool swap = BOTH && j < i;
                const double ax = swap ? double(b.x) : double(xi[c]), ay = swap ? double(b.y) : double(yi[c]);
                const double az = swap ? double(b.z) : double(zi[c]), ar = swap ? double(b.w) : double(ri[c]);
                const double bx = swap ? double(xi[c]) : double(b.x), by = swap ? double(yi[c]) : double(b.y);
                const double bz = swap ? double(zi[c]) : double(b.z), br = swap ? double(ri[c]) : double(b.w);
                if (overlap_exact(ax, ay, az, ar, bx, by, bz, br, margin)) {
                    const unsigned long long at = atomicAdd(count, 1ull);
                    if (at < cap)
                        pairs[at] = swap ? make_uint2(static_cast<unsigned>(j), static_cast<unsigned>(i))
                                         : make_uint2(static_cast<unsigned>(i), static_cast<unsigned>(j));
                }
            }
        }
    }
}

// ---- candidate x candidate -----------------------------------------------------------------
// When two ordinary (not "big") bodies overlap, EACH of them has the other within its reach, so
// both are candidates: the partners of a candidate need only be looked for among the candidates.
// With F candidates of N bodies that is F^2 tests instead of F*N (N=1M, F=64k: 16x fewer).  The
// candidates of all devices, concatenated, are first gathered into dense (x, y, z, r) records.
template <typename Real>
__global__ void gather_candidates_kernel(const Real* pos, const Real* radius, const unsigned* cand, unsigned n_cand,
                                         typename Vec4<Real>::type* rec)
{
    const unsigned k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n_cand) return;
    const size_t i = cand[k];
    typename Vec4<Real>::type v;
    v.x = pos[pm_index<Real>(i, 0)];
    v.y = pos[pm_index<Real>(i, 1)];
    v.z = pos[pm_index<Real>(i, 2)];
    v.w = radius[i];
    rec[k] = v;
}

// grid = (blocks of this device's own candidates [i_off, i_off + n_i), chunks of the whole list [0, n_all)).
// The list is in no particular order, so every (i, j) is visited from both sides and kept where
// index(i) < index(j).
template <typename Real>
__global__ void __launch_bounds__(kContactThreads)
    contact_pairs_compact_kernel(const typename Vec4<Real>::type* rec, const unsigned* cand, unsigned i_off, unsigned n_i,
                                 unsigned n_all, unsigned chunk_len, double margin, uint2* pairs, unsigned long long cap,
                                 unsigned long long* count)
{
    using V4 = typename Vec4<Real>::type;
    constexpr int C = kContactPerThread;
    __shared__ V4 tile[kContactThreads];
    __shared__ unsigned tile_idx[kContactThreads];
    const int tid = threadIdx.x;
    unsigned ci[C];
    V4 me[C];
    float rik[C];
    const float kSlack = 1.000005f;  // see contact_pairs_kernel
#pragma unroll
    for (int c = 0; c < C; ++c) {
        const unsigned k = (blockIdx.x * C + c) * kContactThreads + tid;
        if (k < n_i) {
            ci[c] = cand[i_off + k];
            me[c] = rec[i_off + k];
        } else {  // parked so far away that every pre-test fails
            ci[c] = 0xffffffffu;
            me[c].x = Real(1e30);
            me[c].y = me[c].z = me[c].w = Real(0);
        }
        rik[c] = (float(me[c].w) + float(margin)) * kSlack;
    }
    const unsigned j0 = blockIdx.y * chunk_len;
    const unsigned j1 = (j0 + chunk_len < n_all) ? j0 + chunk_len : n_all;
    for (unsigned jb = j0; jb < j1; jb += kContactThreads) {
        __syncthreads();
        if (jb + tid < j1) {
            tile[tid] = rec[jb + tid];
            tile_idx[tid] = cand[jb + tid];
        }
        __syncthreads();
        const int cnt = static_cast<int>(j1 - jb < unsigned(kContactThreads) ? j1 - jb : unsigned(kContactThreads));
#pragma unroll 4
        for (int q = 0; q < cnt; ++q) {
            const V4 b = tile[q];
#pragma unroll
            for (int c = 0; c < C; ++c) {
                const float dx = float(b.x) - float(me[c].x), dy = float(b.y) - float(me[c].y), dz = float(b.z) - float(me[c].z);
                const float d2 = fmaf(dz, dz, fmaf(dy, dy, dx * dx));
                const float md = fmaf(float(b.w), kSlack, rik[c]);
                if (!(d2 <= md * md)) continue;
                const unsigned i = ci[c], j = tile_idx[q];  // rare from here on
                if (j <= i) continue;                        // (also drops the parked slots: i = 0xffffffff)
                if (overlap_exact(double(me[c].x), double(me[c].y), double(me[c].z), double(me[c].w), double(b.x), double(b.y),
                                  double(b.z), double(b.w), margin)) {
                    const unsigned long long at = atomicAdd(count, 1ull);
                    if (at < cap) pairs[at] = make_uint2(i, j);
                }
            }
        }
    }
}

}  // namespace nb
