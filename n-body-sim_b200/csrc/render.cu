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
// Tile-binned, no depth buffer and no per-pixel atomics:
//   bin   : one warp per body walks the conservative rectangle of 16x16-pixel tiles the
//           body can touch -- a counting pass, a one-CTA prefix sum, a fill pass;
//   tile  : one CTA per tile streams the tile's bodies through shared memory; every
//           thread keeps the nearest hit of its pixel (the lower index on equal depth, so
//           the arbitrary list order does not matter) and shades it.
// (The first version did atomicMin(depth << 32 | id) per hit from one warp per body: 7x
// overdraw made that pass 6x slower than the shading itself, profiles/r1c_render_launches.csv.)
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

// Nearest intersection of the pixel ray with a sphere, as render_tile_kernel evaluates it:
//   oc = cam - centre; a = d.d; b = oc.d; cc = oc.oc - r*r; disc = b*b - a*cc;
//   hit iff disc > 0 and t = (-b - sqrt(disc)) / a > near.   t is the view-space depth (view z of d is -1).

template <typename Real> __device__ __forceinline__ V3 body_centre(const Real* pos, size_t b)
{
    return mk(float(pos[pm_index<Real>(b, 0)]), float(pos[pm_index<Real>(b, 1)]), float(pos[pm_index<Real>(b, 2)]));
}

constexpr int kTile = 16;  // pixels per tile edge; one CTA of kTile*kTile threads renders a tile

// Conservative rectangle of tiles that body (ctr, r) can touch; false if it is culled.
__device__ __forceinline__ bool body_tiles(const RenderCamera& cam, V3 ctr, float r, int w, int h, int& tx0, int& tx1, int& ty0,
                                           int& ty1)
{
    if (!(r > 0.0f)) return false;
    const V3 rel = mk(ctr.x - cam.pos[0], ctr.y - cam.pos[1], ctr.z - cam.pos[2]);
    const float vx = dot3(mk(cam.xa[0], cam.xa[1], cam.xa[2]), rel), vy = dot3(mk(cam.ya[0], cam.ya[1], cam.ya[2]), rel);
    const float depth = -dot3(mk(cam.za[0], cam.za[1], cam.za[2]), rel);
    if (depth + r <= cam.near) return false;  // wholly behind the near plane
    int x0 = 0, x1 = w - 1, y0 = 0, y1 = h - 1;
    const float dmin = depth - r * 1.001f, dmax = depth + r * 1.001f;
    if (dmin > 1e-3f * dmax) {  // otherwise the camera is (almost) inside: the whole screen
        const float sx = 1.0f / (cam.tan_half * cam.aspect), sy = 1.0f / cam.tan_half;
        const float lox = fminf((vx - r) / dmin, (vx - r) / dmax) * sx, hix = fmaxf((vx + r) / dmin, (vx + r) / dmax) * sx;
        const float loy = fminf((vy - r) / dmin, (vy - r) / dmax) * sy, hiy = fmaxf((vy + r) / dmin, (vy + r) / dmax) * sy;
        // ndc -> pixel: px = (ndx + 1)/2 * w - 0.5 ; py = (1 - ndy)/2 * h - 0.5 ; one pixel of slack on each side
        const float fx0 = (lox + 1.0f) * 0.5f * float(w) - 1.5f, fx1 = (hix + 1.0f) * 0.5f * float(w) + 0.5f;
        const float fy0 = (1.0f - hiy) * 0.5f * float(h) - 1.5f, fy1 = (1.0f - loy) * 0.5f * float(h) + 0.5f;
        if (fx1 < 0.0f || fy1 < 0.0f || fx0 > float(w - 1) || fy0 > float(h - 1)) return false;
        x0 = max(0, int(floorf(fmaxf(fx0, 0.0f))));
        y0 = max(0, int(floorf(fmaxf(fy0, 0.0f))));
        x1 = min(w - 1, int(ceilf(fminf(fx1, float(w - 1)))));
        y1 = min(h - 1, int(ceilf(fminf(fy1, float(h - 1)))));
    }
    tx0 = x0 / kTile;
    tx1 = x1 / kTile;
    ty0 = y0 / kTile;
    ty1 = y1 / kTile;
    return true;
}

// Binning, one warp per body.  FILL = false: count the bodies of every tile; FILL = true: write each body into
// its tiles' lists (the order inside a list is arbitrary; the render pass resolves depth ties by index).
template <typename Real, bool FILL>
__global__ void render_bin_kernel(const Real* pos, const Real* radius, size_t n, RenderCamera cam, int w, int h, int tiles_x,
                                  unsigned* counts, const unsigned* offsets, unsigned* list)
{
    const size_t b = (size_t(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (b >= n) return;
    int tx0, tx1, ty0, ty1;
    if (!body_tiles(cam, body_centre(pos, b), float(radius[b]), w, h, tx0, tx1, ty0, ty1)) return;
    const int bw = tx1 - tx0 + 1, cnt = bw * (ty1 - ty0 + 1);
    for (int k = lane; k < cnt; k += 32) {
        const unsigned tile = unsigned(ty0 + k / bw) * unsigned(tiles_x) + unsigned(tx0 + k % bw);
        if (FILL) list[offsets[tile] + atomicAdd(&counts[tile], 1u)] = unsigned(b);
        else atomicAdd(&counts[tile], 1u);
    }
}

// Exclusive prefix sum of counts[0..ntiles) into offsets[0..ntiles], one CTA; zeroes counts for the fill pass.
__global__ void render_scan_kernel(unsigned* counts, unsigned* offsets, int ntiles)
{
    __shared__ unsigned part[1024];
    const int tid = threadIdx.x, per = (ntiles + 1023) / 1024;
    const int lo = min(tid * per, ntiles), hi = min(lo + per, ntiles);
    unsigned s = 0;
    for (int k = lo; k < hi; ++k) s += counts[k];
    part[tid] = s;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {
        const unsigned v = tid >= o ? part[tid - o] : 0u;
        __syncthreads();
        part[tid] += v;
        __syncthreads();
    }
    unsigned run = part[tid] - s;
    for (int k = lo; k < hi; ++k) {
        const unsigned c = counts[k];
        offsets[k] = run;
        run += c;
        counts[k] = 0;
    }
    if (tid == 1023) offsets[ntiles] = part[1023];
}

// One CTA per 16x16-pixel tile.  The tile's bodies stream through shared memory in batches; every thread keeps the
// nearest hit of its pixel (lower index on equal depth, so the list order does not matter), then shades it.
// The hit arithmetic is hit_sphere()'s, operation for operation (oc and oc.oc - r*r are per-body values).
template <typename Real>
__global__ void __launch_bounds__(kTile* kTile)
    render_tile_kernel(const Real* pos, const Real* radius, const Real* temp, const int* type, const float* seed, RenderCamera cam,
                       int w, int h, int tiles_x, const unsigned* offsets, const unsigned* list, float4* rgba, int* ids)
{
    __shared__ float4 s_body[kTile * kTile];  // oc.x, oc.y, oc.z, oc.oc - r*r
    __shared__ unsigned s_id[kTile * kTile];
    const int tid = threadIdx.y * kTile + threadIdx.x;
    const int tile = blockIdx.y * tiles_x + blockIdx.x;
    const int px = blockIdx.x * kTile + threadIdx.x, py = blockIdx.y * kTile + threadIdx.y;
    const bool live = px < w && py < h;
    float dx, dy;
    const V3 d = pixel_dir(cam, live ? px : 0, live ? py : 0, w, h, dx, dy);
    const float a = dot3(d, d);
    float best = -1.0f;
    unsigned who = 0xffffffffu;
    const unsigned beg = offsets[tile], end = offsets[tile + 1];
    for (unsigned base = beg; base < end; base += kTile * kTile) {
        const unsigned cnt = min(unsigned(kTile * kTile), end - base);
        __syncthreads();
        if (unsigned(tid) < cnt) {
            const unsigned b = list[base + tid];
            const V3 ctr = body_centre(pos, size_t(b));
            const float r = float(radius[b]);
            const V3 oc = mk(cam.pos[0] - ctr.x, cam.pos[1] - ctr.y, cam.pos[2] - ctr.z);
            s_body[tid] = make_float4(oc.x, oc.y, oc.z, dot3(oc, oc) - r * r);
            s_id[tid] = b;
        }
        __syncthreads();
        if (!live) continue;
        for (unsigned k = 0; k < cnt; ++k) {
            const float4 q = s_body[k];
            const float b = dot3(mk(q.x, q.y, q.z), d);
            const float disc = b * b - a * q.w;
            if (disc > 0.0f) {
                const float t = (-b - sqrtf(disc)) / a;
                const unsigned id = s_id[k];
                if (t > cam.near && (who == 0xffffffffu || t < best || (t == best && id < who))) {
                    best = t;
                    who = id;
                }
            }
        }
    }
    if (!live) return;
    const size_t pix = size_t(py) * w + px;
    if (who == 0xffffffffu) {
        rgba[pix] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (ids) ids[pix] = -1;
        return;
    }
    const V3 ctr = body_centre(pos, size_t(who));
    const float r = float(radius[who]), t = best;
    const V3 P = mk(cam.pos[0] + t * d.x, cam.pos[1] + t * d.y, cam.pos[2] + t * d.z);                 // vPosition, idx:91
    const V3 nw = mk((P.x - ctr.x) / r, (P.y - ctr.y) / r, (P.z - ctr.z) / r);
    const V3 nv = mk(dot3(mk(cam.xa[0], cam.xa[1], cam.xa[2]), nw), dot3(mk(cam.ya[0], cam.ya[1], cam.ya[2]), nw),
                     dot3(mk(cam.za[0], cam.za[1], cam.za[2]), nw));                                  // normalMatrix * normal, idx:90
    const V3 vview = mk(-(t * dx), -(t * dy), t);                                                     // -mvPosition.xyz, idx:93
    rgba[pix] = planet_fragment(P, nv, vview, type ? type[who] : 0, temp ? float(temp[who]) : 0.0f, seed ? seed[who] : 0.0f);
    if (ids) ids[pix] = int(who);
}

}  // namespace

void render_scratch_free(RenderScratch& s)
{
    cudaFree(s.counts);
    cudaFree(s.offsets);
    cudaFree(s.list);
    s = RenderScratch{};
}

template <typename Real>
cudaError_t render_launch(const Real* pos, const Real* radius, const Real* temp, const int* type, const float* seed, size_t n,
                          const RenderCamera& cam, int w, int h, RenderScratch& sc, float4* rgba, int* ids, cudaStream_t stream)
{
    const int tiles_x = (w + kTile - 1) / kTile, tiles_y = (h + kTile - 1) / kTile;
    const size_t ntiles = size_t(tiles_x) * tiles_y;
    cudaError_t e;
    if (ntiles > sc.ntiles_cap) {
        if ((e = cudaStreamSynchronize(stream)) != cudaSuccess) return e;
        cudaFree(sc.counts);
        cudaFree(sc.offsets);
        sc.counts = sc.offsets = nullptr;
        sc.ntiles_cap = 0;
        if ((e = cudaMalloc(&sc.counts, ntiles * sizeof(unsigned))) != cudaSuccess) return e;
        if ((e = cudaMalloc(&sc.offsets, (ntiles + 1) * sizeof(unsigned))) != cudaSuccess) return e;
        sc.ntiles_cap = ntiles;
    }
    if ((e = cudaMemsetAsync(sc.counts, 0, ntiles * sizeof(unsigned), stream)) != cudaSuccess) return e;
    const unsigned blocks = unsigned((n * 32 + 255) / 256);
    render_bin_kernel<Real, false><<<blocks, 256, 0, stream>>>(pos, radius, n, cam, w, h, tiles_x, sc.counts, nullptr, nullptr);
    render_scan_kernel<<<1, 1024, 0, stream>>>(sc.counts, sc.offsets, int(ntiles));
    unsigned total = 0;  // the list length decides the allocation: one 4-byte read-back per frame
    if ((e = cudaMemcpyAsync(&total, sc.offsets + ntiles, sizeof total, cudaMemcpyDeviceToHost, stream)) != cudaSuccess) return e;
    if ((e = cudaStreamSynchronize(stream)) != cudaSuccess) return e;
    if (total > sc.list_cap) {
        cudaFree(sc.list);
        sc.list = nullptr;
        sc.list_cap = 0;
        const size_t want = size_t(total) + size_t(total) / 4 + 1024;
        if ((e = cudaMalloc(&sc.list, want * sizeof(unsigned))) != cudaSuccess) return e;
        sc.list_cap = want;
    }
    if (total) render_bin_kernel<Real, true><<<blocks, 256, 0, stream>>>(pos, radius, n, cam, w, h, tiles_x, sc.counts, sc.offsets, sc.list);
    const dim3 blk(kTile, kTile), grid(tiles_x, tiles_y);
    render_tile_kernel<Real><<<grid, blk, 0, stream>>>(pos, radius, temp, type, seed, cam, w, h, tiles_x, sc.offsets, sc.list, rgba, ids);
    sc.last_list_entries = total;
    return cudaGetLastError();
}

template cudaError_t render_launch<float>(const float*, const float*, const float*, const int*, const float*, size_t,
                                          const RenderCamera&, int, int, RenderScratch&, float4*, int*, cudaStream_t);
template cudaError_t render_launch<double>(const double*, const double*, const double*, const int*, const float*, size_t,
                                           const RenderCamera&, int, int, RenderScratch&, float4*, int*, cudaStream_t);

}  // namespace nb
