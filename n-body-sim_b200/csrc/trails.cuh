// Device-side position history ring and ribbon-trail vertex builder: the renderer feed
// of SURVEY.md 8(f) row 3.  In the reference every Body carries `history`, the last
// CONFIG.trailLength = 80 positions, newest first (/root/reference/index.html:58, 540,
// 564-565), and VisualSystem.update turns it into a camera-facing ribbon of
// trailLength x 2 vertices per body every frame (idx:567-592).  Keeping the ring on the
// device lets a host renderer pull trails (or finished ribbon vertices) with one copy
// instead of re-uploading 80 positions per body per frame.
//
// Ring layout: hist[slot][body] as Real4 (x, y, z, unused), body-fastest, so a push is one
// coalesced n x 16 B (fp32) store.  `head` is the slot of the newest sample; sample i
// (0 = newest) lives in slot (head + i) % depth.  Both kernels are HBM-bound:
//   push    : 16 B read + 16 B written per body (fp32; 32 + 32 fp64)
//   ribbons : depth x 16 B read + depth x 24 B written per body (fp32 state; output is
//             always float, like the reference's Float32Array, idx:524)
#pragma once
#include "common.cuh"
#include "epilogue.cuh"

namespace nb {

// body.history.unshift(body.pos.clone()) for every body (idx:564); fill_all: a body seen
// for the first time gets trailLength copies of its position (idx:540).  `source` (nullable,
// history_remap_kernel below): bodies with source[i] < 0 are new to the ring and are filled alone.
template <typename Real>
__global__ void history_push_kernel(const Real* pos, typename Vec4<Real>::type* hist, size_t n, size_t stride,
                                    int slot, int depth, int fill_all, const int* source)
{
    using V4 = typename Vec4<Real>::type;
    const size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    V4 v;
    v.x = pos[pm_index<Real>(i, 0)];
    v.y = pos[pm_index<Real>(i, 1)];
    v.z = pos[pm_index<Real>(i, 2)];
    v.w = Real(0);
    if (fill_all || (source != nullptr && source[i] < 0)) {
        for (int s = 0; s < depth; ++s) hist[size_t(s) * stride + i] = v;
    } else {
        hist[size_t(slot) * stride + i] = v;
    }
}

// The body list changed (merge, fracture, spawn; idx:383-389, 835) but Body objects carry their
// history with them: body i of the new list continues the ring of body source[i] of the old one.
// Slots keep their meaning (same head), so this is one gather per slot; rows of new bodies
// (source < 0) are left for the next push to fill.
template <typename Real>
__global__ void history_remap_kernel(const typename Vec4<Real>::type* from, typename Vec4<Real>::type* to, const int* source,
                                     size_t n, size_t stride, int depth)
{
    const size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int src = source[i];
    if (src < 0) return;
    for (int s = 0; s < depth; ++s) to[size_t(s) * stride + i] = from[size_t(s) * stride + size_t(src)];
}

// out[body][i][0..2] = history[i], newest first (what Body.history holds).
template <typename Real>
__global__ void history_gather_kernel(const typename Vec4<Real>::type* hist, Real* out, size_t n, size_t stride,
                                      int head, int depth)
{
    const size_t t = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (t >= n * size_t(depth)) return;
    const size_t b = t / depth;
    const int i = int(t % depth);
    const auto v = hist[size_t((head + i) % depth) * stride + b];
    out[3 * t + 0] = v.x;
    out[3 * t + 1] = v.y;
    out[3 * t + 2] = v.z;
}

__device__ __forceinline__ float sub_rn(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ double sub_rn(double a, double b) { return __dsub_rn(a, b); }
__device__ __forceinline__ float sqrt_rn(float a) { return __fsqrt_rn(a); }
__device__ __forceinline__ double sqrt_rn(double a) { return __dsqrt_rn(a); }
__device__ __forceinline__ float div_rn(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ double div_rn(double a, double b) { return __ddiv_rn(a, b); }
__device__ __forceinline__ float to_f32(float a) { return a; }
__device__ __forceinline__ float to_f32(double a) { return __double2float_rn(a); }

// The ribbon loop of VisualSystem.update (idx:567-592) for kTrailBodies bodies per CTA.
// The CTA stages its bodies' rings in shared memory with coalesced body-fastest reads,
// then every thread builds vertex pairs with point-fastest coalesced stores.  Arithmetic
// is rounded operation by operation in the reference's association (no FMA), so the
// fp64 context reproduces the JS doubles and their Float32Array rounding exactly.
constexpr int kTrailThreads = 256;
// bodies per CTA: one ring row of the CTA is 512 contiguous bytes in either precision
template <typename Real> __host__ __device__ constexpr int trail_bodies() { return sizeof(Real) == 4 ? 32 : 16; }

template <typename Real>
__global__ void __launch_bounds__(kTrailThreads)
    trail_ribbons_kernel(const typename Vec4<Real>::type* hist, const Real* radius, float* out, size_t n,
                         size_t stride, int head, int depth, int trail_len, Real cx, Real cy, Real cz)
{
    using V4 = typename Vec4<Real>::type;
    constexpr int kTrailBodies = trail_bodies<Real>();
    extern __shared__ __align__(16) unsigned char trail_smem[];
    V4* sh = reinterpret_cast<V4*>(trail_smem);  // [kTrailBodies][depth + 1]: body-major, one element of padding per row
    const int row = depth + 1;                   // (sample-major rows made every read of the vertex loop a 32-way bank conflict)
    const size_t b0 = size_t(blockIdx.x) * kTrailBodies;
    const int nb_here = int(n - b0 < size_t(kTrailBodies) ? n - b0 : size_t(kTrailBodies));
    // slot of sample i is (head + i) mod depth; both loops advance their indices incrementally
    // (an integer division per element costs as much as the vertex arithmetic)
    for (int t = threadIdx.x; t < depth * kTrailBodies; t += kTrailThreads) {
        const int i = t / kTrailBodies, lb = t % kTrailBodies;  // kTrailBodies is a power of two
        int slot = head + i;
        if (slot >= depth) slot -= depth;
        if (lb < nb_here) sh[lb * row + i] = hist[size_t(slot) * stride + b0 + lb];
    }
    __syncthreads();
    const int step_lb = kTrailThreads / depth, step_i = kTrailThreads % depth;
    int lb = threadIdx.x / depth, i = threadIdx.x % depth;
    const Real inv_span = div_rn(Real(1), Real(trail_len - 1));
    for (int t = threadIdx.x; t < nb_here * depth; t += kTrailThreads) {
        const V4 p = sh[lb * row + i];
        // widthRatio = 1 - i/(trailLength-1); currentWidth = radius*0.8*max(0, widthRatio)   (idx:569-570)
        // fp64 divides like the reference; fp32 (tolerance parity only) multiplies by the reciprocal
        const Real ratio = sizeof(Real) == 8 ? sub_rn(Real(1), div_rn(Real(i), Real(trail_len - 1))) : Real(1) - Real(i) * inv_span;
        const Real width = mul_rn(mul_rn(radius[b0 + lb], Real(0.8)), ratio > Real(0) ? ratio : Real(0));
        const Real tx = sub_rn(cx, p.x), ty = sub_rn(cy, p.y), tz = sub_rn(cz, p.z);  // vecToCam, idx:572
        V4 a, b;                                                                      // vecDir, idx:574-578
        if (i < depth - 1) {
            a = p;
            b = sh[lb * row + i + 1];
        } else if (i > 0) {
            a = sh[lb * row + i - 1];
            b = p;
        } else {
            a = p;
            b = p;
        }
        const Real dx = sub_rn(a.x, b.x), dy = sub_rn(a.y, b.y), dz = sub_rn(a.z, b.z);
        // crossVectors(vecToCam, vecDir).normalize().multiplyScalar(currentWidth), idx:580
        Real px = sub_rn(mul_rn(ty, dz), mul_rn(tz, dy));
        Real py = sub_rn(mul_rn(tz, dx), mul_rn(tx, dz));
        Real pz = sub_rn(mul_rn(tx, dy), mul_rn(ty, dx));
        const Real len2 = add_rn(add_rn(mul_rn(px, px), mul_rn(py, py)), mul_rn(pz, pz));
        Real s;
        if (sizeof(Real) == 8) {
            Real len = sqrt_rn(len2);
            if (len == Real(0)) len = Real(1);  // length() || 1
            s = div_rn(Real(1), len);
        } else {
            s = len2 > Real(0) ? Real(rsqrtf(float(len2))) : Real(0);  // one MUFU; a zero vector stays zero
        }
        px = mul_rn(mul_rn(px, s), width);
        py = mul_rn(mul_rn(py, s), width);
        pz = mul_rn(mul_rn(pz, s), width);
        // posArr[i*6 + 0..5], idx:582-588; 24 B per thread, 8-byte aligned: three 64-bit stores
        float2* o = reinterpret_cast<float2*>(out + (b0 * size_t(depth) + size_t(t)) * 6);
        o[0] = make_float2(to_f32(add_rn(p.x, px)), to_f32(add_rn(p.y, py)));
        o[1] = make_float2(to_f32(add_rn(p.z, pz)), to_f32(sub_rn(p.x, px)));
        o[2] = make_float2(to_f32(sub_rn(p.y, py)), to_f32(sub_rn(p.z, pz)));
        lb += step_lb;
        i += step_i;
        if (i >= depth) {
            i -= depth;
            ++lb;
        }
    }
}

}  // namespace nb
