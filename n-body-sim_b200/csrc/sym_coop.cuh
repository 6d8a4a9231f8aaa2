// All steps of an nbody_step call in ONE cooperative launch for mid-size fp32 systems (about 5k to 32k bodies):
// the warp-unit pairs-once force pass (force_f32_symw.cuh) between two grid-wide barriers, with the integrator
// (/root/reference/index.html:247-262, gravity only; animate()'s sub-step loop, idx:877-879) in the registers of the
// thread that owns each body.
//
// At these sizes a step is 10-90 us of arithmetic and the launch-per-step path spends three launches on it (kick-1 + drift,
// force, finish) with the machine draining and refilling around each.  Here the CTAs stay resident:
//
//   A  owner threads: kick-1 + drift (idx:250-253) from registers; the new positions go to the position array IN PLACE
//      (nobody reads it between barrier 2 of the previous step and barrier 1 of this one)
//      -- grid barrier 1 --
//   B  every warp pulls work units from the atomic counter and adds fixed-point partial sums to the accumulators
//      (force_f32_symw_body, positions read through L2: other SMs have just written them)
//      -- grid barrier 2 --
//   C  owner threads: read and clear their three accumulators, one rounding to fp32, times G, kick-2 and the thermal
//      decay (idx:258-261); thread 0 rewinds the unit counter
//
// Same unit table, same partial sums, same integer accumulation, same rounded operations in the integrator as the
// launch-per-step kernels: the result is bit-identical to theirs (tests/test_gpu_sym32.py), whatever the schedule.
// The poison word (a non-finite partial sum) is sticky for the rest of the launch -- a NaN acceleration makes every
// later step NaN anyway -- and the host clears it behind the launch.
#pragma once
#include "epilogue.cuh"
#include "force_f32_symw.cuh"
#include "medium.cuh"

namespace nb {

constexpr size_t kSymCoopMaxN = 32768;  // one owner thread per body slot: n_pad <= resident threads (296 CTAs x 128 at R = 8)

struct SymCoopParams {
    Sym32Params sym;     // pos: the current position buffer (updated in place); u_begin = 0; sched and accumulators zero on entry
    float* pos;          // ... the same buffer, writable
    float* vel;          // float4 per body
    float* acc;          // float4 per body: acc of the previous step on entry (first kick-1), of the last step on exit
    float* temp;         // may be null
    unsigned* arrivals;  // grid barrier counter, zero on entry
    unsigned n;
    int nsteps;
    float G, half_dt, dt, heat_decay;
    double inv_scale;    // 2^-k
};

template <int R, bool UNIFORM>
__global__ void __launch_bounds__(kSymwThreads, R == 8 ? 2 : 3) sym_coop_step_kernel(const SymCoopParams p)
{
    // owner of position slot sl (slot order: accumulators and position planes are coalesced; vel/acc/temp are touched twice per launch)
    const size_t sl = size_t(blockIdx.x) * kSymwThreads + threadIdx.x;
    const size_t il = (sl & ~size_t(511)) + tile_body(unsigned(sl) & 511u);
    const bool slot = sl < p.sym.n_pad;  // (pads collect sums too: their accumulators are cleared like everybody's)
    const bool owner = slot && il < p.n;
    const size_t ix = (sl >> 9) * (4 * kPosTile) + (sl & 511);  // plane x; y and z follow at + kPosTile, + 2 kPosTile
    float x = 0.f, y = 0.f, z = 0.f, t = 0.f;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f), a = v;
    const bool thermal_on = p.temp != nullptr && p.heat_decay > 0.f;
    if (owner) {
        x = p.pos[ix];
        y = p.pos[ix + kPosTile];
        z = p.pos[ix + 2 * kPosTile];
        v = reinterpret_cast<const float4*>(p.vel)[il];
        a = reinterpret_cast<const float4*>(p.acc)[il];
        if (thermal_on) t = p.temp[il];
    }
    unsigned long long* fix = p.sym.accfix + sl;
    const unsigned long long* poison = p.sym.accfix + 3 * p.sym.n_pad;
    unsigned goal = 0;
    for (int s = 0; s < p.nsteps; ++s) {
        if (owner) {
            v.x = add_rn(v.x, mul_rn(a.x, p.half_dt));  // kick-1, idx:251
            v.y = add_rn(v.y, mul_rn(a.y, p.half_dt));
            v.z = add_rn(v.z, mul_rn(a.z, p.half_dt));
            x = add_rn(x, mul_rn(v.x, p.dt));  // drift, idx:252
            y = add_rn(y, mul_rn(v.y, p.dt));
            z = add_rn(z, mul_rn(v.z, p.dt));
            p.pos[ix] = x;
            p.pos[ix + kPosTile] = y;
            p.pos[ix + 2 * kPosTile] = z;
        }
        goal += gridDim.x;
        medium_grid_barrier(p.arrivals, goal);
        force_f32_symw_body<R, UNIFORM, true>(p.sym);  // computeGravity, idx:264-291
        goal += gridDim.x;
        medium_grid_barrier(p.arrivals, goal);
        if (sl == 0) *p.sym.sched = 0u;  // (everybody is past the counter; the next readers come after barrier 1)
        float sum[3] = {0.f, 0.f, 0.f};
        if (slot) {
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                const long long q = static_cast<long long>(__ldcg(fix + c * p.sym.n_pad));
                fix[c * p.sym.n_pad] = 0ull;
                sum[c] = float(double(q) * p.inv_scale);
            }
        }
        if (owner) {
            if (__ldcg(poison) != 0ull) sum[0] = sum[1] = sum[2] = __int_as_float(0x7fc00000);
            a.x = mul_rn(sum[0], p.G);
            a.y = mul_rn(sum[1], p.G);
            a.z = mul_rn(sum[2], p.G);
            v.x = add_rn(v.x, mul_rn(a.x, p.half_dt));  // kick-2, idx:259
            v.y = add_rn(v.y, mul_rn(a.y, p.half_dt));
            v.z = add_rn(v.z, mul_rn(a.z, p.half_dt));
            if (thermal_on) t = thermal(t, p.heat_decay);  // idx:260
        }
    }
    if (owner) {
        a.w = 0.f;
        reinterpret_cast<float4*>(p.vel)[il] = v;
        reinterpret_cast<float4*>(p.acc)[il] = a;
        if (thermal_on) p.temp[il] = t;
    }
}

}  // namespace nb
