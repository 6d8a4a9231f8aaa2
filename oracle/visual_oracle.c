/*
 * oracle/visual_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU fp64 restatement of the arithmetic parts of the reference's VisualSystem that
 * the B200 path reproduces headless (/root/reference/index.html, "idx:NNN"):
 *   - the ribbon-trail vertex loop of VisualSystem.update, idx:563-592;
 * PARITY UNPINNED: the reference has no tests and needs a browser + WebGL to run.
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
