/*
 * oracle/visual_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU fp64 restatement of the arithmetic parts of the reference's VisualSystem that
 * the B200 path reproduces headless (/root/reference/index.html, "idx:NNN"):
 *   - the ribbon-trail vertex loop of VisualSystem.update, idx:563-592;
 *     PINNED TO THE REFERENCE'S SOURCE TEXT: oracle/js_reference.py executes the page's own
 *     VisualSystem.update and this loop equals its Float32Array bit for bit
 *     (tests/test_reference_source.py, tests/golden/ref_js_page_galaxy_ribbons.npz);
 *   - the planet shaders, idx:83-202.  The fragment shader's arithmetic is PINNED TO ITS GLSL TEXT:
 *     oracle/glsl_reference.py executes planetFragmentShader from the page, one binary32 operation per
 *     GLSL operation, and planet_fragment / snoise below equal it bit for bit
 *     (tests/test_reference_source.py, tests/golden/ref_glsl_planet_fragments.npz).  Parity against a
 *     real WebGL driver (its own pow/normalize precision, rasterised varyings) stays unpinned.
 * THREE.Vector3 members written out (three@0.160.0): subVectors, crossVectors
 * (x = ay*bz - az*by, y = az*bx - ax*bz, z = ax*by - ay*bx), normalize =
 * multiplyScalar(1 / (length() || 1)), multiplyScalar.
 */
#include <math.h>
#include <stddef.h>

#define ORACLE_API __attribute__((visibility("default")))

/* hist[n][depth][3]: Body.history of every body, newest first (idx:564-565);
 * out[n][depth][6]: the trail mesh's `positions` Float32Array (idx:524, 582-588). */
ORACLE_API void oracle_trail_ribbons(size_t n, int depth, int trail_length, const double *hist,
                                     const double *radius, const double *camera, float *out)
{
    for (size_t b = 0; b < n; b++) {
        const double *h = hist + b * (size_t)depth * 3;
        float *posArr = out + b * (size_t)depth * 6;
        for (int i = 0; i < depth; i++) {
            const double *p = h + 3 * i;
            const double widthRatio = 1.0 - ((double)i / (trail_length - 1));        /* idx:569 */
            const double currentWidth = radius[b] * 0.8 * fmax(0, widthRatio);       /* idx:570 */
            const double tx = camera[0] - p[0], ty = camera[1] - p[1], tz = camera[2] - p[2]; /* idx:572 */
            const double *a, *c;
            if (i < depth - 1) { a = h + 3 * i; c = h + 3 * (i + 1); }               /* idx:574-575 */
            else { a = h + 3 * (i - 1); c = h + 3 * i; }                             /* idx:577 */
            const double dx = a[0] - c[0], dy = a[1] - c[1], dz = a[2] - c[2];
            double px = ty * dz - tz * dy;                                           /* idx:580 */
            double py = tz * dx - tx * dz;
            double pz = tx * dy - ty * dx;
            double len = sqrt(px * px + py * py + pz * pz);
            if (len == 0) len = 1;
            const double s = 1 / len;
            px *= s; py *= s; pz *= s;
            px *= currentWidth; py *= currentWidth; pz *= currentWidth;
            posArr[i * 6 + 0] = (float)(p[0] + px);                                  /* idx:582-584 */
            posArr[i * 6 + 1] = (float)(p[1] + py);
            posArr[i * 6 + 2] = (float)(p[2] + pz);
            posArr[i * 6 + 3] = (float)(p[0] - px);                                  /* idx:586-588 */
            posArr[i * 6 + 4] = (float)(p[1] - py);
            posArr[i * 6 + 5] = (float)(p[2] - pz);
        }
    }
}

/* ==========================================================================
 * Headless body renderer: the planet fragment shader
 * (idx:100-201) restated in C floats, fed by an analytic ray cast of each
 * body's sphere in place of the rasterised icosphere (idx:458, 545-546) and the
 * vertex shader's varyings (idx:83-98).  GLSL `float` is fp32; every operation
 * here is a single IEEE-rounded fp32 operation (compiled with -ffp-contract=off),
 * GLSL built-ins are written out with their specification formulas:
 *   mix(a,b,t) = a*(1-t) + b*t;  step(e,x) = x < e ? 0 : 1;
 *   smoothstep(e0,e1,x): t = clamp((x-e0)/(e1-e0),0,1); t*t*(3-2*t);
 *   normalize(v) = v / sqrt(dot(v,v));  dot = x*x' + y*y' + z*z' left to right;
 *   pow(x, 3.0) (idx:182) = x*x*x (GLSL leaves pow's precision to the driver).
 * The three.js pieces around the shader (ACES tone mapping, bloom, idx:711-727)
 * are not restated.  "unavailable on this box": the WebGL path itself.
 * ========================================================================== */
#include <stdint.h>
#include <string.h>

typedef struct { float x, y, z; } v3;
typedef struct { float x, y, z, w; } v4;

static float mod289f(float x) { return x - floorf(x * (1.0f / 289.0f)) * 289.0f; }              /* idx:109-110 */
static float permutef(float x) { return mod289f(((x * 34.0f) + 1.0f) * x); }                     /* idx:111 */
static float taylor_inv_sqrt(float r) { return 1.79284291400159f - 0.85373472095314f * r; }      /* idx:112 */
static float stepf(float edge, float x) { return x < edge ? 0.0f : 1.0f; }
static float dot3(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static float clampf(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }
static float smoothstepf(float e0, float e1, float x)
{
    const float t = clampf((x - e0) / (e1 - e0), 0.0f, 1.0f);
    return t * t * (3.0f - 2.0f * t);
}
static v3 normalize3(v3 a)
{
    const float l = sqrtf(dot3(a, a));
    v3 r = {a.x / l, a.y / l, a.z / l};
    return r;
}

/* snoise(vec3), idx:114-162 */
static float snoise(v3 v)
{
    const float Cx = 1.0f / 6.0f, Cy = 1.0f / 3.0f;
    const float s = dot3(v, (v3){Cy, Cy, Cy});
    v3 i = {floorf(v.x + s), floorf(v.y + s), floorf(v.z + s)};                     /* idx:117 */
    const float t = dot3(i, (v3){Cx, Cx, Cx});
    const v3 x0 = {v.x - i.x + t, v.y - i.y + t, v.z - i.z + t};                    /* idx:118 */
    const v3 g = {stepf(x0.y, x0.x), stepf(x0.z, x0.y), stepf(x0.x, x0.z)};         /* step(x0.yzx, x0.xyz) */
    const v3 l = {1.0f - g.x, 1.0f - g.y, 1.0f - g.z};
    const v3 i1 = {fminf(g.x, l.z), fminf(g.y, l.x), fminf(g.z, l.y)};              /* min(g.xyz, l.zxy) */
    const v3 i2 = {fmaxf(g.x, l.z), fmaxf(g.y, l.x), fmaxf(g.z, l.y)};
    const v3 x1 = {x0.x - i1.x + Cx, x0.y - i1.y + Cx, x0.z - i1.z + Cx};
    const v3 x2 = {x0.x - i2.x + Cy, x0.y - i2.y + Cy, x0.z - i2.z + Cy};
    const v3 x3 = {x0.x - 0.5f, x0.y - 0.5f, x0.z - 0.5f};                          /* D.yyy */
    i.x = mod289f(i.x); i.y = mod289f(i.y); i.z = mod289f(i.z);                     /* idx:127 */
    const float oz[4] = {0.0f, i1.z, i2.z, 1.0f}, oy[4] = {0.0f, i1.y, i2.y, 1.0f}, ox[4] = {0.0f, i1.x, i2.x, 1.0f};
    float p[4];
    for (int k = 0; k < 4; k++)                                                     /* idx:128-131 */
        p[k] = permutef(permutef(permutef(i.z + oz[k]) + i.y + oy[k]) + i.x + ox[k]);
    const float n_ = 0.142857142857f;
    const v3 ns = {n_ * 2.0f - 0.0f, n_ * 0.5f - 1.0f, n_ * 1.0f - 0.0f};           /* n_*D.wyz - D.xzx */
    float gx[4], gy[4], h[4];
    for (int k = 0; k < 4; k++) {
        const float j = p[k] - 49.0f * floorf(p[k] * ns.z * ns.z);                  /* idx:134 */
        const float x_ = floorf(j * ns.z);
        const float y_ = floorf(j - 7.0f * x_);
        gx[k] = x_ * ns.x + ns.y;
        gy[k] = y_ * ns.x + ns.y;
        h[k] = 1.0f - fabsf(gx[k]) - fabsf(gy[k]);
    }
    const float b0[4] = {gx[0], gx[1], gy[0], gy[1]}, b1[4] = {gx[2], gx[3], gy[2], gy[3]}; /* idx:140-141 */
    float s0[4], s1[4], sh[4];
    for (int k = 0; k < 4; k++) {
        s0[k] = floorf(b0[k]) * 2.0f + 1.0f;
        s1[k] = floorf(b1[k]) * 2.0f + 1.0f;
        sh[k] = -stepf(h[k], 0.0f);                                                 /* -step(h, vec4(0.0)) */
    }
    /* a0 = b0.xzyw + s0.xzyw*sh.xxyy ; a1 = b1.xzyw + s1.xzyw*sh.zzww */
    const float a0[4] = {b0[0] + s0[0] * sh[0], b0[2] + s0[2] * sh[0], b0[1] + s0[1] * sh[1], b0[3] + s0[3] * sh[1]};
    const float a1[4] = {b1[0] + s1[0] * sh[2], b1[2] + s1[2] * sh[2], b1[1] + s1[1] * sh[3], b1[3] + s1[3] * sh[3]};
    v3 pp[4] = {{a0[0], a0[1], h[0]}, {a0[2], a0[3], h[1]}, {a1[0], a1[1], h[2]}, {a1[2], a1[3], h[3]}};
    const v3 xs[4] = {x0, x1, x2, x3};
    float m[4], pd[4];
    for (int k = 0; k < 4; k++) {
        const float nrm = taylor_inv_sqrt(dot3(pp[k], pp[k]));                      /* idx:151-155 */
        pp[k].x *= nrm; pp[k].y *= nrm; pp[k].z *= nrm;
        m[k] = fmaxf(0.6f - dot3(xs[k], xs[k]), 0.0f);                              /* idx:156 */
        m[k] = m[k] * m[k];
        pd[k] = dot3(pp[k], xs[k]);
    }
    /* 42.0 * dot(m*m, vec4(...)): four-term dot, left to right */
    return 42.0f * ((m[0] * m[0]) * pd[0] + (m[1] * m[1]) * pd[1] + (m[2] * m[2]) * pd[2] + (m[3] * m[3]) * pd[3]);
}

/* CONFIG.materials colours (idx:71-73): color1, color2 per type rock/metal/ice */
static const float kColor1[3][3] = {{0.4f, 0.3f, 0.2f}, {0.6f, 0.6f, 0.7f}, {0.8f, 0.9f, 1.0f}};
static const float kColor2[3][3] = {{0.1f, 0.1f, 0.1f}, {0.3f, 0.3f, 0.4f}, {0.1f, 0.3f, 0.6f}};

/* planetFragmentShader main(), idx:164-201 */
static void planet_fragment(v3 vPosition, v3 vNormal, v3 vViewPosition, int type, float temperature, float seed,
                            v3 sun, float *rgba)
{
    const float n = snoise((v3){vPosition.x * 0.5f + seed, vPosition.y * 0.5f + seed, vPosition.z * 0.5f + seed});
    const float sd = seed * 2.0f;
    const float n2 = snoise((v3){vPosition.x * 2.0f + sd, vPosition.y * 2.0f + sd, vPosition.z * 2.0f + sd});
    const float detail = n * 0.6f + n2 * 0.4f;
    const float tmix = smoothstepf(-0.2f, 0.5f, detail);
    float fin[3];
    for (int k = 0; k < 3; k++) fin[k] = kColor2[type][k] * (1.0f - tmix) + kColor1[type][k] * tmix;   /* mix, idx:171 */
    const v3 lightDir = normalize3((v3){sun.x - vPosition.x, sun.y - vPosition.y, sun.z - vPosition.z});
    const v3 viewDir = normalize3(vViewPosition);
    v3 normal = normalize3(vNormal);
    const float bump = detail * 0.1f;
    normal = normalize3((v3){normal.x + bump, normal.y + bump, normal.z + bump});                        /* idx:179 */
    const float diff = fmaxf(dot3(normal, lightDir), 0.0f);
    const float f1 = fmaxf(1.0f - dot3(viewDir, normal), 0.0f);
    const float fresnel = f1 * f1 * f1;                                                                  /* idx:184 */
    const float heatIntensity = clampf(temperature / 50.0f, 0.0f, 1.0f);
    const float crack = smoothstepf(0.4f, 0.6f, fabsf(n2));
    const float heatC[3] = {1.0f, 0.3f, 0.1f}, glowC[3] = {1.0f, 0.5f, 0.2f};
    const float glow = fmaxf(0.0f, temperature - 50.0f);
    for (int k = 0; k < 3; k++) {
        const float atmosphere = kColor1[type][k] * fresnel * 0.5f;
        const float heat = heatC[k] * (1.0f - crack) * heatIntensity * 5.0f;
        const float body = glowC[k] * glow * 0.005f;
        rgba[k] = fin[k] * (diff + 0.05f) + atmosphere + heat + body;                                    /* idx:198 */
    }
    rgba[3] = 1.0f;
}

/* One fragment of planetFragmentShader and one noise sample, exported so that the restatement can be compared with the
 * shader's own text executed by oracle/glsl_reference.py (tests/test_reference_source.py). */
ORACLE_API void oracle_planet_fragment(const float *vpos, const float *vnormal, const float *vview, int type, float temperature,
                                       float seed, const float *sun, float *rgba)
{
    planet_fragment((v3){vpos[0], vpos[1], vpos[2]}, (v3){vnormal[0], vnormal[1], vnormal[2]}, (v3){vview[0], vview[1], vview[2]}, type,
                    temperature, seed, (v3){sun[0], sun[1], sun[2]}, rgba);
}

ORACLE_API float oracle_snoise(const float *v) { return snoise((v3){v[0], v[1], v[2]}); }

typedef struct {
    v3 pos, xa, ya, za; /* camera position and basis (za points from the target to the camera) */
    float tan_half, aspect;
} cam_t;

static cam_t make_camera(const double *position, const double *target, double fov_deg, int width, int height)
{
    cam_t c;
    /* Matrix4.lookAt(eye, target, up = +y): z = normalize(eye - target), x = normalize(cross(up, z)), y = cross(z, x) */
    double z[3] = {position[0] - target[0], position[1] - target[1], position[2] - target[2]};
    double l = sqrt(z[0] * z[0] + z[1] * z[1] + z[2] * z[2]);
    for (int k = 0; k < 3; k++) z[k] /= l;
    double x[3] = {1.0 * z[2] - 0.0 * z[1], 0.0 * z[0] - 0.0 * z[2], 0.0 * z[1] - 1.0 * z[0]};
    l = sqrt(x[0] * x[0] + x[1] * x[1] + x[2] * x[2]);
    for (int k = 0; k < 3; k++) x[k] /= l;
    const double y[3] = {z[1] * x[2] - z[2] * x[1], z[2] * x[0] - z[0] * x[2], z[0] * x[1] - z[1] * x[0]};
    c.pos = (v3){(float)position[0], (float)position[1], (float)position[2]};
    c.xa = (v3){(float)x[0], (float)x[1], (float)x[2]};
    c.ya = (v3){(float)y[0], (float)y[1], (float)y[2]};
    c.za = (v3){(float)z[0], (float)z[1], (float)z[2]};
    c.tan_half = (float)tan(fov_deg * M_PI / 360.0);
    c.aspect = (float)width / (float)height;
    return c;
}

/* Ray through the centre of pixel (px, py), row 0 at the top; view-space z of the direction is -1, so the ray
 * parameter t is the view-space depth. */
static v3 pixel_dir(const cam_t *c, int px, int py, int width, int height, float *dxv, float *dyv)
{
    const float ndx = (2.0f * ((float)px + 0.5f)) / (float)width - 1.0f;
    const float ndy = 1.0f - (2.0f * ((float)py + 0.5f)) / (float)height;
    const float dx = ndx * c->tan_half * c->aspect, dy = ndy * c->tan_half;
    *dxv = dx;
    *dyv = dy;
    v3 d = {c->xa.x * dx + c->ya.x * dy - c->za.x, c->xa.y * dx + c->ya.y * dy - c->za.y, c->xa.z * dx + c->ya.z * dy - c->za.z};
    return d;
}

/* Nearest intersection of the pixel ray with sphere (centre, r): returns t > near or -1. */
static float hit_sphere(const cam_t *c, v3 d, v3 centre, float r, float near)
{
    const v3 oc = {c->pos.x - centre.x, c->pos.y - centre.y, c->pos.z - centre.z};
    const float a = dot3(d, d), b = dot3(oc, d), cc = dot3(oc, oc) - r * r;
    const float disc = b * b - a * cc;
    if (!(disc > 0.0f)) return -1.0f;
    const float t = (-b - sqrtf(disc)) / a;
    return t > near ? t : -1.0f;
}

/* rgba[h][w][4] float (gl_FragColor before tone mapping; background 0,0,0,0), ids[h][w] int32 (-1: none).
 * pos_mass[4n], radius[n], temp[n] are fp32 like the GPU's uniforms; the nearest body wins, the lower index on ties. */
ORACLE_API void oracle_render(size_t n, const float *pos_mass, const float *radius, const float *temp, const int *type,
                              const float *seed, int width, int height, const double *cam_position,
                              const double *cam_target, double fov_deg, float *rgba, int32_t *ids)
{
    const cam_t c = make_camera(cam_position, cam_target, fov_deg, width, height);
    const v3 sun = {50.0f, 50.0f, 50.0f};                                            /* idx:493 */
    const float near = 0.1f;                                                         /* idx:704 */
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic, 4)
#endif
    for (int py = 0; py < height; py++) {
        for (int px = 0; px < width; px++) {
            float dxv, dyv;
            const v3 d = pixel_dir(&c, px, py, width, height, &dxv, &dyv);
            float best = -1.0f;
            long who = -1;
            for (size_t b = 0; b < n; b++) {
                const v3 ctr = {pos_mass[4 * b], pos_mass[4 * b + 1], pos_mass[4 * b + 2]};
                const float t = hit_sphere(&c, d, ctr, radius[b], near);
                if (t > 0.0f && (who < 0 || t < best)) { best = t; who = (long)b; }
            }
            float *o = rgba + 4 * ((size_t)py * width + px);
            if (ids) ids[(size_t)py * width + px] = (int32_t)who;
            if (who < 0) { o[0] = o[1] = o[2] = o[3] = 0.0f; continue; }
            const v3 ctr = {pos_mass[4 * who], pos_mass[4 * who + 1], pos_mass[4 * who + 2]};
            const v3 P = {c.pos.x + best * d.x, c.pos.y + best * d.y, c.pos.z + best * d.z};   /* vPosition, idx:91 */
            const float r = radius[who];
            const v3 nw = {(P.x - ctr.x) / r, (P.y - ctr.y) / r, (P.z - ctr.z) / r};
            const v3 nv = {dot3(c.xa, nw), dot3(c.ya, nw), dot3(c.za, nw)};                    /* normalMatrix*normal, idx:90 */
            const v3 vview = {-(best * dxv), -(best * dyv), best};                             /* -mvPosition.xyz, idx:93 */
            planet_fragment(P, nv, vview, type[who], temp[who], seed[who], sun, o);
        }
    }
}
