// Between the reference's own sizes and the large-N kernels: a few hundred to medium_max_n bodies on one GPU.
// A launch per step costs 12-30 us there (force_f32.cuh small tiles; force_f64.cuh), several times
// the arithmetic.  medium_step_kernel is small.cuh's one-launch form spread over the whole chip:
// nbody_step(ctx, dt, G, eps, nsteps) is ONE cooperative launch of one CTA per SM that runs all
// nsteps steps of integrate() (/root/reference/index.html:247-262, gravity only; animate()'s
// sub-step loop, idx:877-879):
//
//   * CTA c owns bodies [c*per, (c+1)*per); per step its owner threads do kick-1 + drift
//     (idx:250-253) and store the new positions into the context's OTHER position buffer in
//     global memory (the same double buffer the launch-per-step path uses);
//   * one grid-wide barrier (an arrival counter in global memory, release/acquire at GPU scope);
//   * every CTA copies all positions (L2 hits) into shared memory and sums the force on its own
//     bodies: threads are (body, j-slice) pairs, a thread walks its slice of j upward, the slices
//     are added in slice order (fixed: no dependence on scheduling), then kick-2 and the thermal
//     decay (idx:258-261) in the owner's registers;
//   * velocities, accelerations and temperatures return to global memory once, after the last step.
//
// One barrier per step is enough: step s+1 writes the buffer step s read BEFORE its barrier, and
// nobody passes barrier s+1 until everyone has finished copying the buffer of step s.
// Pair arithmetic and integrator are small.cuh's (the large kernels': fp64 MATH 2, fp32 rsqrt.approx + FMA).
#pragma once
#include "common.cuh"
#include "epilogue.cuh"
#include "small.cuh"

namespace nb {

constexpr int kMediumThreads = 1024;
constexpr int kMediumMaxSlices = 32;
// Largest system: where the launch-per-step kernels catch up.  Measured us/step, this kernel / launch per step
// (profiles/r1f_small_n_sweep_final.jsonl): fp32 N=1,024 3.4 / 12.6, 2,048 6.0 / 14.3, 4,096 12.7 / 19.4, 6,144 25.8 / 24.7;
// fp64 N=1,024 5.9 / 12.7, 2,048 10.7 / 16.4, 4,096 26.6 / 30.7 (and 160 KiB of shared memory there).
template <typename Real> __host__ __device__ constexpr int medium_max_n() { return sizeof(Real) == 4 ? 5120 : 4096; }

template <typename Real> constexpr size_t medium_smem_bytes() { return (size_t(medium_max_n<Real>()) + kMediumThreads) * 4 * sizeof(Real); }

template <typename Real> struct MediumParams {
    Real* pos[2];       // the context's position buffers (pm_index layout, masses in both); pos[0] is current on entry
    Real* vel;          // Real4 per body
    Real* acc;          // Real4 per body: acc of the previous step on entry (first kick-1), of the last step on exit
    Real* temp;         // may be null
    unsigned* arrivals; // grid barrier counter, zero on entry
    int n;
    int nsteps;
    Real G, eps2, r2_floor, half_dt, dt, heat_decay;
};

// All CTAs of a cooperative launch (co-resident by construction).  `goal` = arrivals expected so far.
__device__ __forceinline__ void medium_grid_barrier(unsigned* arrivals, unsigned goal)
{
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();  // the CTA's position stores (ordered before this by the barrier above) become visible GPU-wide
        atomicAdd(arrivals, 1u);
        unsigned seen;
        do {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(arrivals) : "memory");
        } while (seen < goal);
    }
    __syncthreads();
}

template <typename Real>
__global__ void __launch_bounds__(kMediumThreads) medium_step_kernel(const MediumParams<Real> p)
{
    using V4 = typename Vec4<Real>::type;
    extern __shared__ __align__(16) unsigned char medium_smem[];
    V4* body = reinterpret_cast<V4*>(medium_smem);  // [medium_max_n]: (x, y, z, m) of every body at the step's force evaluation
    V4* part = body + medium_max_n<Real>();          // [kMediumThreads]: per-thread partial sums of one step
    const int rank = int(blockIdx.x), ctas = int(gridDim.x);
    const int tid = threadIdx.x;
    const int n = p.n;

    // (body, j-slice) thread grid: `per` bodies (the CTA's share), as many slices as fit the CTA
    const int per = (n + ctas - 1) / ctas;       // <= kMediumThreads: the host launches at least n / kMediumThreads CTAs
    const int fit = kMediumThreads / per;
    const int slices = fit < kMediumMaxSlices ? fit : kMediumMaxSlices;
    const int slice = tid / per, il = tid - slice * per;
    const int gi = rank * per + il;
    const bool mine = slice < slices && gi < n;
    const bool owner = mine && slice == 0;

    auto load_all = [&](const Real* src) {
        for (int j = tid; j < n; j += kMediumThreads) {
            V4 b;
            b.x = __ldcg(&src[pm_index<Real>(size_t(j), 0)]);  // L2: another SM wrote it
            b.y = __ldcg(&src[pm_index<Real>(size_t(j), 1)]);
            b.z = __ldcg(&src[pm_index<Real>(size_t(j), 2)]);
            b.w = __ldcg(&src[pm_index<Real>(size_t(j), 3)]);
            body[j] = b;
        }
    };

    load_all(p.pos[0]);
    V4 v = {}, a = {};
    Real t = Real(0);
    const bool thermal_on = p.temp != nullptr && p.heat_decay > Real(0);
    if (owner) {
        v = reinterpret_cast<const V4*>(p.vel)[gi];
        a = reinterpret_cast<const V4*>(p.acc)[gi];
        if (thermal_on) t = p.temp[gi];
    }
    __syncthreads();

    for (int s = 0; s < p.nsteps; ++s) {
        Real* nxt = (s & 1) ? p.pos[0] : p.pos[1];
        if (owner) {
            V4 b = body[gi];
            v.x = add_rn(v.x, mul_rn(a.x, p.half_dt));  // kick-1, idx:251
            v.y = add_rn(v.y, mul_rn(a.y, p.half_dt));
            v.z = add_rn(v.z, mul_rn(a.z, p.half_dt));
            b.x = add_rn(b.x, mul_rn(v.x, p.dt));  // drift, idx:252
            b.y = add_rn(b.y, mul_rn(v.y, p.dt));
            b.z = add_rn(b.z, mul_rn(v.z, p.dt));
            nxt[pm_index<Real>(size_t(gi), 0)] = b.x;
            nxt[pm_index<Real>(size_t(gi), 1)] = b.y;
            nxt[pm_index<Real>(size_t(gi), 2)] = b.z;
        }
        medium_grid_barrier(p.arrivals, unsigned(s + 1) * unsigned(ctas));
        load_all(nxt);
        __syncthreads();

        Real sx = Real(0), sy = Real(0), sz = Real(0);
        if (mine) {
            const V4 me = body[gi];
            int j = slice;
            for (; j + 3 * slices < n; j += 4 * slices) {  // four independent chains in flight, added in ascending j
                const int d = gi - j;                       // the self pair is the (d / slices)-th of the four, if any
                const int self = d == 0 ? 0 : d == slices ? 1 : d == 2 * slices ? 2 : d == 3 * slices ? 3 : -1;
                small_pair4(me, body[j], body[j + slices], body[j + 2 * slices], body[j + 3 * slices], self, p.eps2, p.r2_floor, sx, sy, sz);
            }
            for (; j < n; j += slices) small_pair(me, body[j], j == gi, p.eps2, p.r2_floor, sx, sy, sz);
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
        // (the next write to `part` and to `body` comes after the next grid barrier's first __syncthreads)
    }

    if (owner) {
        a.w = Real(0);
        reinterpret_cast<V4*>(p.vel)[gi] = v;
        reinterpret_cast<V4*>(p.acc)[gi] = a;
        if (thermal_on) p.temp[gi] = t;
    }
}

}  // namespace nb
