/*
 * oracle/nbody_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU fp64 restatement of the reference's gravity + kick/drift step
 * (/root/reference/index.html, cited below as idx:NNN).  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load this library; the product path (libnbody.so) never does.
 *
 * PARITY PINNED TO THE REFERENCE'S SOURCE TEXT, NOT TO V8: the reference ships
 * no tests, fixtures or golden vectors and no JavaScript
 * engine exists in the build container or on the GPU box.  oracle/js_reference.py
 * slices the reference's own statements out of index.html at run time and
 * executes them through a JavaScript-subset transliterator (IEEE binary64
 * arithmetic, libm pow); this restatement equals that execution BIT FOR BIT
 * (tests/test_reference_source.py, committed vectors tests/golden/ref_js_*.npz).
 * What stays unpinned is V8 itself: its Math.pow may differ from libm's by an
 * ulp on rare inputs (1e-16 relative, four orders below the 1e-12 gate).
 *
 * Arithmetic rules (so that the result equals V8's wherever pow() agrees):
 *   - double everywhere, compiled with -O2 -ffp-contract=off -fno-fast-math
 *     (JS never fuses a multiply-add);
 *   - the reference's association is kept: (f*m)*v, (x*x + y*y) + z*z,
 *     distSq + epsSq, G / pow(s, 1.5);
 *   - the reference's i<j pair loop with paired += / -= updates is kept, so
 *     every body's sum runs in ascending-j order with the self term skipped.
 *
 * Host layout (same as include/nbody.h): pos_mass[4n] = x,y,z,m per body,
 * vel[4n] = vx,vy,vz,unused, acc[4n] = ax,ay,az,unused, temp[n].
 */
#include <math.h>
#include <stddef.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ORACLE_API __attribute__((visibility("default")))

/* THREE.Vector3.addScaledVector (three@0.160.0, call sites idx:251,252,259):
 * this.x += v.x*s; this.y += v.y*s; this.z += v.z*s -- a rounded multiply
 * followed by a rounded add. */
static inline void add_scaled(double *dst, const double *v, double s)
{
    dst[0] += v[0] * s;
    dst[1] += v[1] * s;
    dst[2] += v[2] * s;
}

/* PhysicsSystem.computeGravity, idx:264-291. */
ORACLE_API void oracle_compute_gravity(size_t n, const double *pos_mass, double *acc,
                                       double G, double softening)
{
    for (size_t i = 0; i < n; i++) {            /* idx:266 acc.set(0,0,0) */
        acc[4 * i + 0] = 0.0;
        acc[4 * i + 1] = 0.0;
        acc[4 * i + 2] = 0.0;
    }
    const double epsSq = softening * softening; /* idx:269 */

    for (size_t i = 0; i < n; i++) {            /* idx:272 */
        const double *A = pos_mass + 4 * i;
        for (size_t j = i + 1; j < n; j++) {    /* idx:273 */
            const double *B = pos_mass + 4 * j;
            /* vec.subVectors(B.pos, A.pos), idx:277 */
            const double vx = B[0] - A[0];
            const double vy = B[1] - A[1];
            const double vz = B[2] - A[2];
            /* vec.lengthSq(): x*x + y*y + z*z, left to right, idx:278 */
            const double distSq = vx * vx + vy * vy + vz * vz;
            /* idx:280 */
            const double f = G / pow(distSq + epsSq, 1.5);
            /* idx:282-284: A.acc.c += f * B.mass * vec.c  == (f*B.mass)*vec.c */
            acc[4 * i + 0] += f * B[3] * vx;
            acc[4 * i + 1] += f * B[3] * vy;
            acc[4 * i + 2] += f * B[3] * vz;
            /* idx:286-288 */
            acc[4 * j + 0] -= f * A[3] * vx;
            acc[4 * j + 1] -= f * A[3] * vy;
            acc[4 * j + 2] -= f * A[3] * vz;
        }
    }
}

/* Body.updateThermal, idx:227-230, with CONFIG.heatDecay passed in (idx:67). */
static inline double update_thermal(double t, double heat_decay)
{
    t *= heat_decay;
    if (t < 0.1) t = 0.0;
    return t;
}

/* First loop of PhysicsSystem.integrate, idx:248-253. */
ORACLE_API void oracle_kick1_drift(size_t n, double *pos_mass, double *vel, const double *acc,
                                   double dt)
{
    const double halfDt = 0.5 * dt;             /* idx:248 */
    for (size_t i = 0; i < n; i++) {
        add_scaled(vel + 4 * i, acc + 4 * i, halfDt);     /* idx:251 */
        add_scaled(pos_mass + 4 * i, vel + 4 * i, dt);    /* idx:252 */
    }
}

/* Second loop of PhysicsSystem.integrate, idx:258-261.  temp may be NULL
 * (gravity-only contract); heat_decay <= 0 skips the thermal update too. */
ORACLE_API void oracle_kick2(size_t n, double *vel, const double *acc, double *temp, double dt,
                             double heat_decay)
{
    const double halfDt = 0.5 * dt;
    for (size_t i = 0; i < n; i++) {
        add_scaled(vel + 4 * i, acc + 4 * i, halfDt);     /* idx:259 */
        if (temp && heat_decay > 0.0) temp[i] = update_thermal(temp[i], heat_decay); /* idx:260 */
    }
}

/* nsteps x PhysicsSystem.integrate(dt), idx:247-262, with resolveCollisions
 * (idx:256) removed: the gravity-only step contract of SURVEY.md 8(b). */
ORACLE_API void oracle_integrate(size_t n, double *pos_mass, double *vel, double *acc,
                                 double *temp, double dt, double G, double softening,
                                 double heat_decay, int nsteps)
{
    for (int s = 0; s < nsteps; s++) {
        oracle_kick1_drift(n, pos_mass, vel, acc, dt);
        oracle_compute_gravity(n, pos_mass, acc, G, softening);
        oracle_kick2(n, vel, acc, temp, dt, heat_decay);
    }
}

/* Acceleration of a SUBSET of bodies against all n bodies, each sum in the
 * reference's order (ascending j, self skipped, idx:272-288).  For body k the
 * reference adds -(f*m_j)*(pos_k - pos_j) for j<k and +(f*m_j)*(pos_j - pos_k)
 * for j>k; both are reproduced literally.  Lets tests check full-size scenes
 * in O(k*n) instead of O(n^2). */
ORACLE_API void oracle_acc_subset(size_t n, const double *pos_mass, size_t k, const int64_t *idx,
                                  double *acc_out, double G, double softening)
{
    const double epsSq = softening * softening;
    for (size_t q = 0; q < k; q++) {
        const size_t b = (size_t)idx[q];
        const double *P = pos_mass + 4 * b;
        double ax = 0.0, ay = 0.0, az = 0.0;
        for (size_t j = 0; j < n; j++) {
            if (j == b) continue;
            const double *Q = pos_mass + 4 * j;
            if (j < b) {
                /* body b is the B side of pair (j,b): vec = pos_b - pos_j; acc_b -= (f*m_j)*vec */
                const double vx = P[0] - Q[0], vy = P[1] - Q[1], vz = P[2] - Q[2];
                const double distSq = vx * vx + vy * vy + vz * vz;
                const double f = G / pow(distSq + epsSq, 1.5);
                ax -= f * Q[3] * vx;
                ay -= f * Q[3] * vy;
                az -= f * Q[3] * vz;
            } else {
                /* body b is the A side of pair (b,j): vec = pos_j - pos_b; acc_b += (f*m_j)*vec */
                const double vx = Q[0] - P[0], vy = Q[1] - P[1], vz = Q[2] - P[2];
                const double distSq = vx * vx + vy * vy + vz * vz;
                const double f = G / pow(distSq + epsSq, 1.5);
                ax += f * Q[3] * vx;
                ay += f * Q[3] * vy;
                az += f * Q[3] * vz;
            }
        }
        acc_out[4 * q + 0] = ax;
        acc_out[4 * q + 1] = ay;
        acc_out[4 * q + 2] = az;
        acc_out[4 * q + 3] = 0.0;
    }
}

/* Diagnostics that the reference does not have (SURVEY.md 8(d) "Energy
 * definition"), consistent with the force law at idx:280:
 *   E = sum 1/2 m v^2  -  G sum_{i<j} m_i m_j / sqrt(r_ij^2 + eps^2). */
ORACLE_API void oracle_energy(size_t n, const double *pos_mass, const double *vel, double G,
                              double softening, double *kinetic, double *potential,
                              double *momentum /* [3] */)
{
    const double epsSq = softening * softening;
    double ke = 0.0, pe = 0.0, px = 0.0, py = 0.0, pz = 0.0;
    for (size_t i = 0; i < n; i++) {
        const double m = pos_mass[4 * i + 3];
        const double *v = vel + 4 * i;
        ke += 0.5 * m * (v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
        px += m * v[0];
        py += m * v[1];
        pz += m * v[2];
    }
#ifdef _OPENMP
#pragma omp parallel for reduction(+ : pe) schedule(dynamic, 64)
#endif
    for (long i = 0; i < (long)n; i++) {
        const double *A = pos_mass + 4 * i;
        double row = 0.0;
        for (size_t j = (size_t)i + 1; j < n; j++) {
            const double *B = pos_mass + 4 * j;
            const double vx = B[0] - A[0], vy = B[1] - A[1], vz = B[2] - A[2];
            row += B[3] / sqrt(vx * vx + vy * vy + vz * vz + epsSq);
        }
        pe += A[3] * row;
    }
    *kinetic = ke;
    *potential = -G * pe;
    momentum[0] = px;
    momentum[1] = py;
    momentum[2] = pz;
}

/* Courtesy multi-threaded variant: NOT the reference's path (the reference is
 * one thread, idx:873-883).  All-pairs over i with OpenMP, each body's sum in
 * the same ascending-j order as the reference, so it returns the same values
 * as oracle_compute_gravity wherever (f*m)*v is sign-symmetric (it is: the
 * negation of vec is exact).  Used to check mid-size scenes quickly. */
ORACLE_API void oracle_compute_gravity_mt(size_t n, const double *pos_mass, double *acc, double G,
                                          double softening)
{
    const double epsSq = softening * softening;
#ifdef _OPENMP
#pragma omp parallel for schedule(static)
#endif
    for (long b = 0; b < (long)n; b++) {
        const double *P = pos_mass + 4 * b;
        double ax = 0.0, ay = 0.0, az = 0.0;
        for (size_t j = 0; j < n; j++) {
            if (j == (size_t)b) continue;
            const double *Q = pos_mass + 4 * j;
            if (j < (size_t)b) {
                const double vx = P[0] - Q[0], vy = P[1] - Q[1], vz = P[2] - Q[2];
                const double distSq = vx * vx + vy * vy + vz * vz;
                const double f = G / pow(distSq + epsSq, 1.5);
                ax -= f * Q[3] * vx;
                ay -= f * Q[3] * vy;
                az -= f * Q[3] * vz;
            } else {
                const double vx = Q[0] - P[0], vy = Q[1] - P[1], vz = Q[2] - P[2];
                const double distSq = vx * vx + vy * vy + vz * vz;
                const double f = G / pow(distSq + epsSq, 1.5);
                ax += f * Q[3] * vx;
                ay += f * Q[3] * vy;
                az += f * Q[3] * vz;
            }
        }
        acc[4 * b + 0] = ax;
        acc[4 * b + 1] = ay;
        acc[4 * b + 2] = az;
    }
}

ORACLE_API int oracle_max_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
