// Headless body renderer: the reference's planet shaders
// (/root/reference/index.html:83-202) as CUDA kernels, no graphics API involved.
//
// The page rasterises an icosphere per body (idx:458, 545-546) and shades it with
// planetFragmentShader; here each body's sphere is ray-cast analytically, the vertex
// shader's varyings (vPosition, vNormal, vViewPosition, idx:89-94) are evaluated at the
// hit point, and the fragment shader (3-D simplex noise, colour mix, diffuse + fresnel
// rim + heat glow, idx:108-201) runs per pixel.  Output is gl_FragColor before three.js's
// tone mapping and bloom (idx:711-727, library code that is not restated).
//
// Two passes over a 64-bit depth/id buffer:
//   depth : one warp per body walks the body's screen bounding box; every hit does
//           atomicMin(zbuf[pixel], float_bits(t) << 32 | body) -- nearest body wins, the
//           lower index on a tie, independent of scheduling;
//   shade : one thread per pixel re-derives the winner's hit point and shades it.
// GLSL `float` is fp32.  This translation unit is compiled with -fmad=false and uses
// IEEE-rounded division and square root, so every operation is a single rounded fp32
// operation, the same as the CPU oracle's (tests compare the images bit for bit).
// Bound: FP32 issue (about 900 flop per shaded pixel against 24 B of traffic).
#include "render.cuh"

namespace nb {
namespace {

struct V3 {
    float x, y, z;
};
__device__ __forceinline__ V3 mk(float x, float y, float z) { return V3{x, y, z}; }
__device__ __forceinline__ float dot3(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ V3 normalize3(V3 a)
{
    const float l = sqrtf(dot3(a, a));
    return mk(a.x / l, a.y / l, a.z / l);
}
__device__ __forceinline__ float stepf(float edge, float x) { return x < edge ? 0.0f : 1.0f; }
__device__ __forceinline__ float clampf(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }
__device__ __forceinline__ float smoothstepf(float e0, float e1, float x)
{
    const float t = clampf((x - e0) / (e1 - e0), 0.0f, 1.0f);
    return t * t * (3.0f - 2.0f * t);
}
__device__ __forceinline__ float mod289(float x) { return x - floorf(x * (1.0f / 289.0f)) * 289.0f; }  // idx:109-110
__device__ __forceinline__ float permute(float x) { return mod289(((x * 34.0f) + 1.0f) * x); }          // idx:111

// float snoise(vec3 v), idx:114-162, the vec4 lanes unrolled by hand
__device__ float snoise(V3 v)
{
    const float Cx = 1.0f / 6.0f, Cy = 1.0f / 3.0f;
    const float s = dot3(v, mk(Cy, Cy, Cy));
    float ix = floorf(v.x + s), iy = floorf(v.y + s), iz = floorf(v.z + s);
    const float t = dot3(mk(ix, iy, iz), mk(Cx, Cx, Cx));
    const V3 x0 = mk(v.x - ix + t, v.y - iy + t, v.z - iz + t);
    const V3 g = mk(stepf(x0.y, x0.x), stepf(x0.z, x0.y), stepf(x0.x, x0.z));
    const V3 l = mk(1.0f - g.x, 1.0f - g.y, 1.0f - g.z);
    const V3 i1 = mk(fminf(g.x, l.z), fminf(g.y, l.x), fminf(g.z, l.y));
    const V3 i2 = mk(fmaxf(g.x, l.z), fmaxf(g.y, l.x), fmaxf(g.z, l.y));
    V3 xs[4];
    xs[0] = x0;
    xs[1] = mk(x0.x - i1.x + Cx, x0.y - i1.y + Cx, x0.z - i1.z + Cx);
    xs[2] = mk(x0.x - i2.x + Cy, x0.y - i2.y + Cy, x0.z - i2.z + Cy);
    xs[3] = mk(x0.x - 0.5f, x0.y - 0.5f, x0.z - 0.5f);
    ix = mod289(ix);
    iy = mod289(iy);
    iz = mod289(iz);
    const float ox[4] = {0.0f, i1.x, i2.x, 1.0f}, oy[4] = {0.0f, i1.y, i2.y, 1.0f}, oz[4] = {0.0f, i1.z, i2.z, 1.0f};
    const float n_ = 0.142857142857f;
    const float nsx = n_ * 2.0f - 0.0f, nsy = n_ * 0.5f - 1.0f, nsz = n_ * 1.0f - 0.0f;
    float gx[4], gy[4], h[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const float p = permute(permute(permute(iz + oz[k]) + iy + oy[k]) + ix + ox[k]);
        const float j = p - 49.0f * floorf(p * nsz * nsz);
        const float x_ = floorf(j * nsz);
        const float y_ = floorf(j - 7.0f * x_);
        gx[k] = x_ * nsx + nsy;
        gy[k] = y_ * nsx + nsy;
        h[k] = 1.0f - fabsf(gx[k]) - fabsf(gy[k]);
    }
    // b0 = (x.xy, y.xy), b1 = (x.zw, y.zw); a = b.xzyw + s.xzyw * sh.{xxyy|zzww}: gradient k gets (gx, gy) moved by sh[k]
    float acc = 0.0f;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const float sh = -stepf(h[k], 0.0f);
        const float ax = gx[k] + (floorf(gx[k]) * 2.0f + 1.0f) * sh;
        const float ay = gy[k] + (floorf(gy[k]) * 2.0f + 1.0f) * sh;
        V3 p = mk(ax, ay, h[k]);
        const float nrm = 1.79284291400159f - 0.85373472095314f * dot3(p, p);  // taylorInvSqrt, idx:112
        p = mk(p.x * nrm, p.y * nrm, p.z * nrm);
        float m = fmaxf(0.6f - dot3(xs[k], xs[k]), 0.0f);
        m = m * m;
        const float term = (m * m) * dot3(p, xs[k]);
        acc = k == 0 ? term : acc + term;
    }
    return 42.0f * acc;
}

__constant__ float kColor1[3][3] = {{0.4f, 0.3f, 0.2f}, {0.6f, 0.6f, 0.7f}, {0.8f, 0.9f, 1.0f}};  // CONFIG.materials, idx:71-73
__constant__ float kColor2[3][3] = {{0.1f, 0.1f, 0.1f}, {0.3f, 0.3f, 0.4f}, {0.1f, 0.3f, 0.6f}};

// planetFragmentShader main(), idx:164-201
__device__ float4 planet_fragment(V3 vPosition, V3 vNormal, V3 vViewPosition, int type, float temperature, float seed)
{
    const float n = snoise(mk(vPosition.x * 0.5f + seed, vPosition.y * 0.5f + seed, vPosition.z * 0.5f + seed));
    const float sd = seed * 2.0f;
    const float n2 = snoise(mk(vPosition.x * 2.0f + sd, vPosition.y * 2.0f + sd, vPosition.z * 2.0f + sd));
    const float detail = n * 0.6f + n2 * 0.4f;
    const float tmix = smoothstepf(-0.2f, 0.5f, detail);
    const V3 lightDir = normalize3(mk(50.0f - vPosition.x, 50.0f - vPosition.y, 50.0f - vPosition.z));  // sunPosition, idx:493
    const V3 viewDir = normalize3(vViewPosition);
    V3 normal = normalize3(vNormal);
    const float bump = detail * 0.1f;
    normal = normalize3(mk(normal.x + bump, normal.y + bump, normal.z + bump));
    const float diff = fmaxf(dot3(normal, lightDir), 0.0f);
    const float f1 = fmaxf(1.0f - dot3(viewDir, normal), 0.0f);
    const float fresnel = f1 * f1 * f1;  // pow(x, 3.0), idx:184
    const float heatIntensity = clampf(temperature / 50.0f, 0.0f, 1.0f);
    const float crack = smoothstepf(0.4f, 0.6f, fabsf(n2));
    const float glow = fmaxf(0.0f, temperature - 50.0f);
    const float heatC[3] = {1.0f, 0.3f, 0.1f}, glowC[3] = {1.0f, 0.5f, 0.2f};
    float out[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float fin = kColor2[type][k] * (1.0f - tmix) + kColor1[type][k] * tmix;  // mix(color2, color1, t)
        const float atmosphere = kColor1[type][k] * fresnel * 0.5f;
        const float heat = heatC[k] * (1.0f - crack) * heatIntensity * 5.0f;
        const float body = glowC[k] * glow * 0.005f;
        out[k] = fin * (diff + 0.05f) + atmosphere + heat + body;
    }
    return make_float4(out[0], out[1], out[2], 1.0f);
}

__device__ __forceinline__ V3 pixel_dir(const RenderCamera& c, int px, int py, int w, int h, float& dx, float& dy)
{
    const float ndx = (2.0f * (float(px) + 0.5f)) / float(w) - 1.0f;
    const float ndy = 1.0f - (2.0f * (float(py) + 0.5f)) / float(h);
    dx = ndx * c.tan_half * c.aspect;
    dy = ndy * c.tan_half;
    return mk(c.xa[0] * dx + c.ya[0] * dy - c.za[0], c.xa[1] * dx + c.ya[1] * dy - c.za[1], c.xa[2] * dx + c.ya[2] * dy - c.za[2]);
}

__device__ __forceinline__ float hit_sphere(const RenderCamera& c, V3 d, V3 ctr, float r)
{
    const V3 oc = mk(c.pos[0] - ctr.x, c.pos[1] - ctr.y, c.pos[2] - ctr.z);
    const float a = dot3(d, d), b = dot3(oc, d), cc = dot3(oc, oc) - r * r;
    const float disc = b * b - a * cc;
    if (!(disc > 0.0f)) return -1.0f;
    const float t = (-b - sqrtf(disc)) / a;
    return t > c.near ? t : -1.0f;
}

template <typename Real> __device__ __forceinline__ V3 body_centre(const Real* pos, size_t b)
{
    return mk(float(pos[pm_index<Real>(b, 0)]), float(pos[pm_index<Real>(b, 1)]), float(pos[pm_index<Real>(b, 2)]));
}

constexpr unsigned long long kEmpty = ~0ull;

// One warp per body: conservative screen bounding box, exact per-pixel hit test, atomicMin on (depth, id).
template <typename Real>
__global__ void render_depth_kernel(const Real* pos, const Real* radius, size_t n, RenderCamera cam, int w, int h,
                                    unsigned long long* zbuf)
{
    const size_t b = (size_t(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (b >= n) return;
    const V3 ctr = body_centre(pos, b);
    const float r = float(radius[b]);
    if (!(r > 0.0f)) return;
    const V3 rel = mk(ctr.x - cam.pos[0], ctr.y - cam.pos[1], ctr.z - cam.pos[2]);
    const float vx = dot3(mk(cam.xa[0], cam.xa[1], cam.xa[2]), rel), vy = dot3(mk(cam.ya[0], cam.ya[1], cam.ya[2]), rel);
    const float depth = -dot3(mk(cam.za[0], cam.za[1], cam.za[2]), rel);
    if (depth + r <= cam.near) return;  // wholly behind the near plane
    int x0 = 0, x1 = w - 1, y0 = 0, y1 = h - 1;
    const float dmin = depth - r * 1.001f, dmax = depth + r * 1.001f;
    if (dmin > 1e-3f * dmax) {  // otherwise the camera is (almost) inside: test the whole screen
        const float sx = 1.0f / (cam.tan_half * cam.aspect), sy = 1.0f / cam.tan_half;
        const float lox = fminf((vx - r) / dmin, (vx - r) / dmax) * sx, hix = fmaxf((vx + r) / dmin, (vx + r) / dmax) * sx;
        const float loy = fminf((vy - r) / dmin, (vy - r) / dmax) * sy, hiy = fmaxf((vy + r) / dmin, (vy + r) / dmax) * sy;
        // ndc -> pixel: px = (ndx + 1)/2 * w - 0.5 ; py = (1 - ndy)/2 * h - 0.5 ; one pixel of slack on each side
        const float fx0 = (lox + 1.0f) * 0.5f * float(w) - 1.5f, fx1 = (hix + 1.0f) * 0.5f * float(w) + 0.5f;
        const float fy0 = (1.0f - hiy) * 0.5f * float(h) - 1.5f, fy1 = (1.0f - loy) * 0.5f * float(h) + 0.5f;
        if (fx1 < 0.0f || fy1 < 0.0f || fx0 > float(w - 1) || fy0 > float(h - 1)) return;
        x0 = max(0, int(floorf(fmaxf(fx0, 0.0f))));
        y0 = max(0, int(floorf(fmaxf(fy0, 0.0f))));
        x1 = min(w - 1, int(ceilf(fminf(fx1, float(w - 1)))));
        y1 = min(h - 1, int(ceilf(fminf(fy1, float(h - 1)))));
    }
    const int bw = x1 - x0 + 1, cnt = bw * (y1 - y0 + 1);
    for (int k = lane; k < cnt; k += 32) {
        const int px = x0 + k % bw, py = y0 + k / bw;
        float dx, dy;
        const V3 d = pixel_dir(cam, px, py, w, h, dx, dy);
        const float t = hit_sphere(cam, d, ctr, r);
        if (t > 0.0f) atomicMin(&zbuf[size_t(py) * w + px], (unsigned long long)__float_as_uint(t) << 32 | (unsigned long long)b);
    }
}

template <typename Real>
__global__ void render_shade_kernel(const Real* pos, const Real* radius, const Real* temp, const int* type, const float* seed,
                                    RenderCamera cam, int w, int h, const unsigned long long* zbuf, float4* rgba, int* ids)
{
    const int px = blockIdx.x * blockDim.x + threadIdx.x, py = blockIdx.y * blockDim.y + threadIdx.y;
    if (px >= w || py >= h) return;
    const size_t pix = size_t(py) * w + px;
    const unsigned long long key = zbuf[pix];
    if (key == kEmpty) {
        rgba[pix] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (ids) ids[pix] = -1;
        return;
    }
    const size_t b = size_t(key & 0xffffffffull);
    const float t = __uint_as_float(unsigned(key >> 32));
    float dx, dy;
    const V3 d = pixel_dir(cam, px, py, w, h, dx, dy);
    const V3 ctr = body_centre(pos, b);
    const float r = float(radius[b]);
    const V3 P = mk(cam.pos[0] + t * d.x, cam.pos[1] + t * d.y, cam.pos[2] + t * d.z);                 // vPosition, idx:91
    const V3 nw = mk((P.x - ctr.x) / r, (P.y - ctr.y) / r, (P.z - ctr.z) / r);
    const V3 nv = mk(dot3(mk(cam.xa[0], cam.xa[1], cam.xa[2]), nw), dot3(mk(cam.ya[0], cam.ya[1], cam.ya[2]), nw),
                     dot3(mk(cam.za[0], cam.za[1], cam.za[2]), nw));                                  // normalMatrix * normal, idx:90
    const V3 vview = mk(-(t * dx), -(t * dy), t);                                                     // -mvPosition.xyz, idx:93
    rgba[pix] = planet_fragment(P, nv, vview, type ? type[b] : 0, temp ? float(temp[b]) : 0.0f, seed ? seed[b] : 0.0f);
    if (ids) ids[pix] = int(b);
}

}  // namespace

template <typename Real>
cudaError_t render_launch(const Real* pos, const Real* radius, const Real* temp, const int* type, const float* seed, size_t n,
                          const RenderCamera& cam, int w, int h, unsigned long long* zbuf, float4* rgba, int* ids,
                          cudaStream_t stream)
{
    cudaError_t e = cudaMemsetAsync(zbuf, 0xff, size_t(w) * h * sizeof(unsigned long long), stream);
    if (e != cudaSuccess) return e;
    const unsigned blocks = unsigned((n * 32 + 255) / 256);
    render_depth_kernel<Real><<<blocks, 256, 0, stream>>>(pos, radius, n, cam, w, h, zbuf);
    const dim3 blk(32, 8), grid((w + 31) / 32, (h + 7) / 8);
    render_shade_kernel<Real><<<grid, blk, 0, stream>>>(pos, radius, temp, type, seed, cam, w, h, zbuf, rgba, ids);
    return cudaGetLastError();
}

template cudaError_t render_launch<float>(const float*, const float*, const float*, const int*, const float*, size_t,
                                          const RenderCamera&, int, int, unsigned long long*, float4*, int*, cudaStream_t);
template cudaError_t render_launch<double>(const double*, const double*, const double*, const int*, const float*, size_t,
                                           const RenderCamera&, int, int, unsigned long long*, float4*, int*, cudaStream_t);

}  // namespace nb
