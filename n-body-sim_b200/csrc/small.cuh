// The reference's own regime: N <= 300 bodies (CONFIG.maxBodies, /root/reference/index.html:57;
// the default scene has 151), where a step is ~23,000 interactions and one kernel launch per
// step costs more than the arithmetic.  nbody_step(ctx, dt, G, eps, nsteps) for a small system
// on one GPU is ONE launch of one thread-block cluster that runs all nsteps steps
// (animate()'s sub-step loop, idx:877-879, and integrate(), idx:247-262, gravity only):
//
//   * kSmallCluster CTAs; CTA c owns bodies [c*per, (c+1)*per) and keeps its own copy of every
//     body's (x, y, z, m) in shared memory, double-buffered;
//   * per step: the owner does kick-1 + drift (idx:250-253) and stores the new position into the
//     next buffer of EVERY CTA of the cluster through distributed shared memory -- the position
//     exchange of the multi-GPU path in miniature -- then one cluster barrier;
//   * every CTA sums the force on its own bodies: threads are (body, j-slice) pairs, a thread
//     walks its slice of j upward, the slices are added in slice order (a fixed order: results do
//     not depend on scheduling), then kick-2 and the thermal decay (idx:258-261);
//   * state goes back to global memory once, after the last step.
//
// Arithmetic per pair is the large kernels' (fp64: force_f64.cuh MATH 2; fp32: rsqrt.approx and
// FMA, general masses); the integrator is the shared rounded-multiply-then-add of epilogue.cuh.
#pragma once
#include <cooperative_groups.h>

#include "common.cuh"
#include "epilogue.cuh"
#include "force_f32.cuh"
#include "force_f64.cuh"

namespace nb {

namespace cg = cooperative_groups;

constexpr int kSmallCluster = 8;    // portable cluster size
constexpr int kSmallThreads = 1024;  // 8 warps per scheduler: the per-pair dependency chain (~13 ops) hides behind other warps
constexpr int kSmallMaxN = 512;     // what the kernel's shared memory holds
constexpr int kSmallDefaultMaxN64 = 224, kSmallDefaultMaxN32 = 384;  // above this the all-SM one-launch kernel (medium.cuh) takes over

template <typename Real> constexpr size_t small_smem_bytes() { return (2 * size_t(kSmallMaxN) + kSmallThreads) * 4 * sizeof(Real); }

template <typename Real> struct SmallParams {
    Real* pos;    // device pos_mass (pm_index layout), the CURRENT buffer: read, then overwritten
    Real* vel;    // Real4 per body
    Real* acc;    // Real4 per body: acc of the previous step on entry (first kick-1), of the last step on exit
    Real* temp;   // may be null
    Real* rec_out;  // may be null; else the final state also goes out as one record [pos_mass 4n | vel 4n | acc 4n | temp n]
                    // (integrate.cuh) -- nbody_step_host points it at the mapped pinned record, so no unpack kernel follows
    int n;
    int nsteps;
    Real G, eps2, r2_floor, half_dt, dt, heat_decay;  // r2_floor: 1e-20 when eps == 0 (fp32 self pair), else 0
};

__device__ __forceinline__ void small_pair(const float4& me, const float4& b, bool, float eps2, float r2_floor, float& ax, float& ay,
                                           float& az)
{
    const float dx = b.x - me.x, dy = b.y - me.y, dz = b.z - me.z;
    float r2 = fmaf(dz, dz, fmaf(dy, dy, fmaf(dx, dx, eps2)));
    r2 = fmaxf(r2, r2_floor);  // the self pair with eps = 0: 0 * finite instead of 0 * inf (force_f32.cuh, ZERO_SOFT)
    const float inv = rsqrt_approx(r2);
    const float s = (b.w * inv) * (inv * inv);
    ax = fmaf(s, dx, ax);
    ay = fmaf(s, dy, ay);
    az = fmaf(s, dz, az);
}

__device__ __forceinline__ void small_pair(const double4& me, const double4& b, bool self, double eps2, double, double& ax, double& ay,
                                           double& az)
{
    const double vx = __dsub_rn(b.x, me.x), vy = __dsub_rn(b.y, me.y), vz = __dsub_rn(b.z, me.z);
    const double s = __fma_rn(vz, vz, __fma_rn(vy, vy, __fma_rn(vx, vx, eps2)));
    double fm = __dmul_rn(inv_cube_f64(s), b.w);
    if (self) fm = 0.0;  // the reference never evaluates the self pair (idx:273); with eps = 0 it would be 0 * inf
    ax = __fma_rn(fm, vx, ax);
    ay = __fma_rn(fm, vy, ay);
    az = __fma_rn(fm, vz, az);
}

// Four pairs at once, stage by stage, so that the four dependent chains (subtract, distSq, rsqrt,
// cube, accumulate) overlap; `self` is the index (0..3) of the self pair among them, or -1.
__device__ __forceinline__ void small_pair4(const float4& me, const float4& b0, const float4& b1, const float4& b2, const float4& b3, int,
                                            float eps2, float r2_floor, float& ax, float& ay, float& az)
{
    const float4 b[4] = {b0, b1, b2, b3};
    float dx[4], dy[4], dz[4], s[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        dx[k] = b[k].x - me.x;
        dy[k] = b[k].y - me.y;
        dz[k] = b[k].z - me.z;
        const float r2 = fmaxf(fmaf(dz[k], dz[k], fmaf(dy[k], dy[k], fmaf(dx[k], dx[k], eps2))), r2_floor);
        const float inv = rsqrt_approx(r2);
        s[k] = (b[k].w * inv) * (inv * inv);
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        ax = fmaf(s[k], dx[k], ax);
        ay = fmaf(s[k], dy[k], ay);
        az = fmaf(s[k], dz[k], az);
    }
}

__device__ __forceinline__ void small_pair4(const double4& me, const double4& b0, const double4& b1, const double4& b2, const double4& b3,
                                            int self, double eps2, double, double& ax, double& ay, double& az)
{
    const double4 b[4] = {b0, b1, b2, b3};
    double vx[4], vy[4], vz[4], fm[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        vx[k] = __dsub_rn(b[k].x, me.x);
        vy[k] = __dsub_rn(b[k].y, me.y);
        vz[k] = __dsub_rn(b[k].z, me.z);
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const double s = __fma_rn(vz[k], vz[k], __fma_rn(vy[k], vy[k], __fma_rn(vx[k], vx[k], eps2)));
        fm[k] = __dmul_rn(inv_cube_f64(s), b[k].w);
        if (k == self) fm[k] = 0.0;
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        ax = __fma_rn(fm[k], vx[k], ax);
        ay = __fma_rn(fm[k], vy[k], ay);
        az = __fma_rn(fm[k], vz[k], az);
    }
}

// Cluster-wide barrier with release/acquire at cluster scope (what the distributed-shared-memory
// stores need).  cooperative_groups' cluster.sync() puts a MEMBAR.ALL.GPU in front of it, which at
// ~1 us per step was most of the step.
__device__ __forceinline__ void cluster_barrier()
{
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

template <typename Real>
__global__ void __cluster_dims__(kSmallCluster, 1, 1) __launch_bounds__(kSmallThreads) small_step_kernel(const SmallParams<Real> p)
{
    using V4 = typename Vec4<Real>::type;
    extern __shared__ __align__(16) unsigned char small_smem[];
    V4(*body)[kSmallMaxN] = reinterpret_cast<V4(*)[kSmallMaxN]>(small_smem);  // [2][kSmallMaxN]: (x, y, z, m) of every body, this step's and the next
    V4* part = reinterpret_cast<V4*>(small_smem) + 2 * kSmallMaxN;            // [kSmallThreads]: per-thread partial sums of one step
    cg::cluster_group cluster = cg::this_cluster();
    const int rank = int(cluster.block_rank());
    const int tid = threadIdx.x;
    const int n = p.n;

    // (body, j-slice) thread grid: `per` bodies (the CTA's share), as many slices as fit the CTA
    const int per = (n + kSmallCluster - 1) / kSmallCluster;
    const int fit = kSmallThreads / per;     // >= 16: per <= kSmallMaxN / kSmallCluster = 64
    const int slices = fit < 32 ? fit : 32;  // (a handful of bodies: more slices would only lengthen the owner's sum)
    const int slice = tid / per, il = tid - slice * per;
    const int gi = rank * per + il;
    const bool mine = slice < slices && gi < n;  // this thread works on a body of this CTA
    const bool owner = mine && slice == 0;       // ... and carries its velocity, acceleration and temperature

    for (int j = tid; j < n; j += kSmallThreads) {
        V4 b;
        b.x = p.pos[pm_index<Real>(size_t(j), 0)];
        b.y = p.pos[pm_index<Real>(size_t(j), 1)];
        b.z = p.pos[pm_index<Real>(size_t(j), 2)];
        b.w = p.pos[pm_index<Real>(size_t(j), 3)];
        body[0][j] = b;
    }
    V4 v = {}, a = {};
    Real t = Real(0);
    const bool thermal_on = p.temp != nullptr && p.heat_decay > Real(0);
    if (owner) {
        v = reinterpret_cast<const V4*>(p.vel)[gi];
        a = reinterpret_cast<const V4*>(p.acc)[gi];
        if (p.temp != nullptr) t = p.temp[gi];
    }
    cluster_barrier();  // every CTA is resident and has its copy before anyone stores into a peer

    int cur = 0;
    for (int s = 0; s < p.nsteps; ++s) {
        const int nxt = cur ^ 1;
        if (owner) {
            V4 b = body[cur][gi];
            v.x = add_rn(v.x, mul_rn(a.x, p.half_dt));  // kick-1, idx:251
            v.y = add_rn(v.y, mul_rn(a.y, p.half_dt));
            v.z = add_rn(v.z, mul_rn(a.z, p.half_dt));
            b.x = add_rn(b.x, mul_rn(v.x, p.dt));  // drift, idx:252
            b.y = add_rn(b.y, mul_rn(v.y, p.dt));
            b.z = add_rn(b.z, mul_rn(v.z, p.dt));
#pragma unroll
            for (int r = 0; r < kSmallCluster; ++r) *cluster.map_shared_rank(&body[nxt][gi], r) = b;
        }
        // Peers read body[cur] of step s-1 only before this barrier and body[nxt] only after it, and the
        // stores of step s+1 go to the other buffer: one barrier per step is enough.
        cluster_barrier();

        Real sx = Real(0), sy = Real(0), sz = Real(0);
        if (mine) {
            const V4 me = body[nxt][gi];
            const V4* bj = body[nxt];
            int j = slice;
            for (; j + 3 * slices < n; j += 4 * slices) {  // four independent chains in flight, added in ascending j
                const int d = gi - j;                       // the self pair is the (d / slices)-th of the four, if any
                const int self = d == 0 ? 0 : d == slices ? 1 : d == 2 * slices ? 2 : d == 3 * slices ? 3 : -1;
                small_pair4(me, bj[j], bj[j + slices], bj[j + 2 * slices], bj[j + 3 * slices], self, p.eps2, p.r2_floor, sx, sy, sz);
            }
            for (; j < n; j += slices) small_pair(me, bj[j], j == gi, p.eps2, p.r2_floor, sx, sy, sz);
        }
        V4 mine_sum;
        mine_sum.x = sx;
        mine_sum.y = sy;
        mine_sum.z = sz;
        mine_sum.w = Real(0);
        part[tid] = mine_sum;
        __syncthreads();
        if (owner) {
            Real tx = Real(0), ty = Real(0), tz = Real(0);
            for (int q = 0; q < slices; ++q) {
                const V4 w = part[q * per + il];
                tx = add_rn(tx, w.x);
                ty = add_rn(ty, w.y);
                tz = add_rn(tz, w.z);
            }
            a.x = mul_rn(tx, p.G);
            a.y = mul_rn(ty, p.G);
            a.z = mul_rn(tz, p.G);
            v.x = add_rn(v.x, mul_rn(a.x, p.half_dt));  // kick-2, idx:259
            v.y = add_rn(v.y, mul_rn(a.y, p.half_dt));
            v.z = add_rn(v.z, mul_rn(a.z, p.half_dt));
            if (thermal_on) t = thermal(t, p.heat_decay);  // idx:260
        }
        cur = nxt;  // (the next write to `part` comes after the next cluster barrier)
    }

    if (owner) {
        const V4 b = body[cur][gi];
        p.pos[pm_index<Real>(size_t(gi), 0)] = b.x;
        p.pos[pm_index<Real>(size_t(gi), 1)] = b.y;
        p.pos[pm_index<Real>(size_t(gi), 2)] = b.z;
        a.w = Real(0);
        reinterpret_cast<V4*>(p.vel)[gi] = v;
        reinterpret_cast<V4*>(p.acc)[gi] = a;
        if (thermal_on) p.temp[gi] = t;
        if (p.rec_out != nullptr) {
            V4* rec = reinterpret_cast<V4*>(p.rec_out);
            rec[gi] = b;
            rec[n + gi] = v;
            rec[2 * n + gi] = a;
            if (p.temp != nullptr) p.rec_out[12 * size_t(n) + gi] = t;
        }
    }
}

}  // namespace nb
