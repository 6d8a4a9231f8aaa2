// Integrator arithmetic shared by the fused force epilogue and the O(N) kernels.
// Every update is a rounded multiply followed by a rounded add, never an FMA:
// that is THREE.Vector3.addScaledVector as the reference calls it
// (/root/reference/index.html:251,252,259 -- `this.x += v.x * s`), so the fp64
// integrator is bit-identical to the oracle's given the same acc.
#pragma once
#include "common.cuh"

namespace nb {

__device__ __forceinline__ float mul_rn(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float add_rn(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ double mul_rn(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double add_rn(double a, double b) { return __dadd_rn(a, b); }

// Body.updateThermal, idx:227-230.
template <typename Real> __device__ __forceinline__ Real thermal(Real t, Real heat_decay)
{
    t = mul_rn(t, heat_decay);
    return t < Real(0.1) ? Real(0) : t;
}

// Finish body `il` (local index) whose summed acceleration is (sx,sy,sz)*scale.
// mode: see EpilogueMode.  Writes acc, vel, temp and -- in mode 2 -- the next
// step's drifted position into p.out (the OTHER position buffer of every rank:
// other CTAs, here and on the peers, are still reading pos_cur).
template <typename Real>
__device__ __forceinline__ void finish_body(const ForceParams<Real>& p, size_t il, Real sx, Real sy,
                                            Real sz, Real scale)
{
    using V4 = typename Vec4<Real>::type;
    NB_CHECK(il < p.n_local);
    V4 a;
    a.x = mul_rn(sx, scale);
    a.y = mul_rn(sy, scale);
    a.z = mul_rn(sz, scale);
    a.w = Real(0);
    reinterpret_cast<V4*>(p.acc)[il] = a;
    if (p.mode == kForceOnly) return;

    V4 v = reinterpret_cast<V4*>(p.vel)[il];
    v.x = add_rn(v.x, mul_rn(a.x, p.half_dt));  // kick-2, idx:259
    v.y = add_rn(v.y, mul_rn(a.y, p.half_dt));
    v.z = add_rn(v.z, mul_rn(a.z, p.half_dt));
    if (p.temp != nullptr && p.heat_decay > Real(0)) p.temp[il] = thermal(p.temp[il], p.heat_decay);

    if (p.mode == kForceKick2Kick1) {
        // first loop of the NEXT integrate(dt) call, idx:250-253
        v.x = add_rn(v.x, mul_rn(a.x, p.half_dt));
        v.y = add_rn(v.y, mul_rn(a.y, p.half_dt));
        v.z = add_rn(v.z, mul_rn(a.z, p.half_dt));
        const size_t g = p.i_begin + il;
        const size_t ix = pm_index<Real>(g, 0), iy = pm_index<Real>(g, 1), iz = pm_index<Real>(g, 2);
        const Real nx = add_rn(p.pos_cur[ix], mul_rn(v.x, p.dt));
        const Real ny = add_rn(p.pos_cur[iy], mul_rn(v.y, p.dt));
        const Real nz = add_rn(p.pos_cur[iz], mul_rn(v.z, p.dt));
        // the position exchange, fused: one store per rank (own buffer first, then NVLink peers)
        for (int k = 0; k < p.out.n; ++k) {
            Real* o = p.out.dst[k];
            o[ix] = nx;
            o[iy] = ny;
            o[iz] = nz;
        }
    }
    reinterpret_cast<V4*>(p.vel)[il] = v;
}

// Cross-CTA completion of one i-tile: every CTA of the tile has stored its
// per-chunk partial sums; the CTA that arrives last adds them in ascending
// chunk order (a fixed order: results do not depend on CTA scheduling or on
// how many GPUs share the work) and runs the integrator.  Returns true in the
// last CTA only.  Called by the CTA's `nthreads` compute threads (named barrier 1).
__device__ __forceinline__ bool tile_arrive_last(unsigned* counter, int j_chunks, int* s_flag, int nthreads)
{
    __threadfence();
    asm volatile("bar.sync 1, %0;" ::"r"(nthreads) : "memory");
    if (threadIdx.x == 0) {
        const unsigned prev = atomicAdd(counter, 1u);
        NB_CHECK(prev < unsigned(j_chunks));
        *s_flag = (prev == static_cast<unsigned>(j_chunks - 1));
    }
    asm volatile("bar.sync 1, %0;" ::"r"(nthreads) : "memory");
    const bool last = (*s_flag != 0);
    if (last) __threadfence();
    return last;
}

}  // namespace nb
