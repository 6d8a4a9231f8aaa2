/*
 * oracle/collide_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU fp64 restatement of the stage that sits between force evaluation and
 * kick-2 in the reference step: PhysicsSystem.resolveCollisions, mergeBodies
 * and fractureBodies (/root/reference/index.html:293-443, "idx:NNN" below),
 * together with the Body constructor (idx:208-225), addBody (idx:239-245) and
 * the full integrate(dt) that calls them (idx:247-262).  SURVEY.md 8(f) row 1.
 *
 * PARITY PINNED TO THE REFERENCE'S SOURCE TEXT, NOT TO V8 (see nbody_oracle.c):
 * the reference has no tests or golden vectors for this stage and no JS engine
 * exists here; oracle/js_reference.py executes the reference's own
 * resolveCollisions / mergeBodies / fractureBodies text and this restatement
 * equals it bit for bit through merges and fractures, new ids and random draws
 * included (tests/test_reference_source.py).  The restatement is
 * statement by statement, in the reference's association, with the
 * THREE.Vector3 members it calls written out (three@0.160.0, un-vendored):
 *   normalize()       = multiplyScalar(1 / (length() || 1))
 *   length()          = sqrt(x*x + y*y + z*z)
 *   dot(v)            = x*v.x + y*v.y + z*v.z
 *   multiplyScalar(s) = x*=s, y*=s, z*=s;  add / sub component-wise.
 *
 * Two substitutions, both forced by running outside a browser:
 *   - Math.random (unseeded in the reference) is a sequential splitmix64 stream
 *     owned by the world, one draw wherever the reference draws one, INCLUDING
 *     the draw the Body constructor spends on `id` (idx:210);
 *   - Body.id (a random base-36 string, idx:210) is that draw itself (a double
 *     in [0,1)); the contact map is keyed by the unordered id pair exactly like
 *     `[A.id, B.id].sort().join('-')` (idx:311).
 * `visuals.*` calls (idx:358, 384, 408, 441) have no physics effect and are
 * counted, not executed.
 */
#include <math.h>
#include <stddef.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORACLE_API __attribute__((visibility("default")))

void oracle_compute_gravity(size_t n, const double *pos_mass, double *acc, double G,
                            double softening); /* nbody_oracle.c */

typedef struct {
    double id;
    double mass, invMass;
    double pos[3], vel[3], acc[3];
    double radius, temperature;
    int type; /* 0 rock, 1 metal, 2 ice (CONFIG.materials, idx:70-74) */
} obody;

typedef struct {
    double lo, hi; /* unordered id pair */
    double time;
    int seen; /* currentFrameContacts membership, idx:299,312 */
} ocontact;

typedef struct oracle_world {
    obody *b;
    size_t n, cap;
    ocontact *c; /* PhysicsSystem.contactPairs, idx:236 */
    size_t nc, ccap;
    uint64_t rng;
    /* CONFIG fields read by this stage (idx:57, 62-64, 67) */
    size_t maxBodies;
    double fractureThreshold, minFragmentMass, mergeTime, heatDecay;
    /* bookkeeping for tests */
    long n_merge, n_fracture, n_overlap, n_flash, n_explosion;
} oracle_world;

static const double kDensity[3] = {1.0, 3.0, 0.5};

/* Math.random stand-in: sequential splitmix64 -> [0,1) with 53 bits. */
static double rnd(oracle_world *w)
{
    uint64_t z = (w->rng += 0x9E3779B97F4A7C15ULL);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    z ^= z >> 31;
    return (double)(z >> 11) * (1.0 / 9007199254740992.0);
}

ORACLE_API oracle_world *oracle_world_create(uint64_t seed, size_t maxBodies)
{
    oracle_world *w = (oracle_world *)calloc(1, sizeof *w);
    w->rng = seed;
    w->maxBodies = maxBodies;      /* CONFIG.maxBodies = 300, idx:57 */
    w->fractureThreshold = 25.0;   /* idx:62 */
    w->minFragmentMass = 0.2;      /* idx:63 */
    w->mergeTime = 0.5;            /* idx:64 */
    w->heatDecay = 0.998;          /* idx:67 */
    return w;
}

ORACLE_API void oracle_world_seed(oracle_world *w, uint64_t seed) { w->rng = seed; }

ORACLE_API void oracle_world_destroy(oracle_world *w)
{
    if (!w) return;
    free(w->b);
    free(w->c);
    free(w);
}

ORACLE_API void oracle_world_config(oracle_world *w, double fractureThreshold,
                                    double minFragmentMass, double mergeTime, double heatDecay)
{
    w->fractureThreshold = fractureThreshold;
    w->minFragmentMass = minFragmentMass;
    w->mergeTime = mergeTime;
    w->heatDecay = heatDecay;
}

/* new Body(mass, pos, vel, type, temp), idx:209-225 */
static obody make_body(oracle_world *w, double mass, const double *pos, const double *vel, int type,
                       double temp)
{
    obody o;
    memset(&o, 0, sizeof o);
    o.id = rnd(w);                                   /* idx:210 */
    o.mass = mass;
    o.invMass = mass > 0 ? 1 / mass : 0;             /* idx:212 */
    memcpy(o.pos, pos, sizeof o.pos);
    memcpy(o.vel, vel, sizeof o.vel);
    o.type = type;
    o.radius = pow((3 * mass) / (4 * M_PI * kDensity[type]), 1.0 / 3); /* idx:221 */
    o.temperature = temp;
    return o;
}

/* PhysicsSystem.addBody, idx:239-245 */
static void add_body(oracle_world *w, const obody *o)
{
    if (w->n >= w->maxBodies && w->n > 0) {          /* bodies.shift(), idx:240-242 */
        memmove(w->b, w->b + 1, (w->n - 1) * sizeof *w->b);
        w->n--;
    }
    if (w->n == w->cap) {
        w->cap = w->cap ? 2 * w->cap : 64;
        w->b = (obody *)realloc(w->b, w->cap * sizeof *w->b);
    }
    w->b[w->n++] = *o;
}

ORACLE_API void oracle_world_add(oracle_world *w, double mass, const double *pos, const double *vel,
                                 int type, double temp)
{
    obody o = make_body(w, mass, pos, vel, type, temp);
    add_body(w, &o);
}

ORACLE_API void oracle_world_clear(oracle_world *w) /* resetScenario, idx:745-746 */
{
    w->n = 0;
    w->nc = 0;
}

ORACLE_API size_t oracle_world_size(const oracle_world *w) { return w->n; }
ORACLE_API size_t oracle_world_contacts(const oracle_world *w) { return w->nc; }

/* contactPairs as (lower id, higher id, accumulated time) triples, in storage order; the Math.random state */
ORACLE_API void oracle_world_get_contacts(const oracle_world *w, double *lo_hi_time)
{
    for (size_t k = 0; k < w->nc; k++) {
        lo_hi_time[3 * k] = w->c[k].lo;
        lo_hi_time[3 * k + 1] = w->c[k].hi;
        lo_hi_time[3 * k + 2] = w->c[k].time;
    }
}
ORACLE_API uint64_t oracle_world_rng(const oracle_world *w) { return w->rng; }

ORACLE_API void oracle_world_counters(const oracle_world *w, long *out /* [5] */)
{
    out[0] = w->n_merge;
    out[1] = w->n_fracture;
    out[2] = w->n_overlap;
    out[3] = w->n_flash;
    out[4] = w->n_explosion;
}

/* Copy the state out in the host layout of include/nbody.h (+ radius, type, id). */
ORACLE_API void oracle_world_get(const oracle_world *w, double *pos_mass, double *vel, double *acc,
                                 double *temp, double *radius, int *type, double *id)
{
    for (size_t i = 0; i < w->n; i++) {
        const obody *o = &w->b[i];
        for (int k = 0; k < 3; k++) {
            if (pos_mass) pos_mass[4 * i + k] = o->pos[k];
            if (vel) vel[4 * i + k] = o->vel[k];
            if (acc) acc[4 * i + k] = o->acc[k];
        }
        if (pos_mass) pos_mass[4 * i + 3] = o->mass;
        if (vel) vel[4 * i + 3] = 0.0;
        if (acc) acc[4 * i + 3] = 0.0;
        if (temp) temp[i] = o->temperature;
        if (radius) radius[i] = o->radius;
        if (type) type[i] = o->type;
        if (id) id[i] = o->id;
    }
}

/* contactPairs lookup by `[A.id, B.id].sort().join('-')`, idx:311 */
static ocontact *contact_find(oracle_world *w, double ida, double idb)
{
    const double lo = ida < idb ? ida : idb, hi = ida < idb ? idb : ida;
    for (size_t k = 0; k < w->nc; k++)
        if (w->c[k].lo == lo && w->c[k].hi == hi) return &w->c[k];
    return NULL;
}

static ocontact *contact_set(oracle_world *w, double ida, double idb, double t)
{
    ocontact *e = contact_find(w, ida, idb);
    if (!e) {
        if (w->nc == w->ccap) {
            w->ccap = w->ccap ? 2 * w->ccap : 64;
            w->c = (ocontact *)realloc(w->c, w->ccap * sizeof *w->c);
        }
        e = &w->c[w->nc++];
        e->lo = ida < idb ? ida : idb;
        e->hi = ida < idb ? idb : ida;
    }
    e->time = t;
    return e;
}

static void contact_delete(oracle_world *w, ocontact *e)
{
    *e = w->c[--w->nc]; /* Map order has no effect on results */
}

typedef struct {
    obody *v;
    size_t n, cap;
} blist;

static void blist_push(blist *l, const obody *o)
{
    if (l->n == l->cap) {
        l->cap = l->cap ? 2 * l->cap : 16;
        l->v = (obody *)realloc(l->v, l->cap * sizeof *l->v);
    }
    l->v[l->n++] = *o;
}

/* PhysicsSystem.mergeBodies, idx:392-409 */
static void merge_bodies(oracle_world *w, const obody *A, const obody *B, blist *newBodies)
{
    const double newMass = A->mass + B->mass;
    const double inv = 1 / newMass;
    double newVel[3], newPos[3];
    for (int k = 0; k < 3; k++) {
        newVel[k] = (A->vel[k] * A->mass + B->vel[k] * B->mass) * inv; /* idx:394-396 */
        newPos[k] = (A->pos[k] * A->mass + B->pos[k] * B->mass) * inv; /* idx:398-400 */
    }
    const double newTemp = (A->temperature * A->mass + B->temperature * B->mass) / newMass; /* idx:402 */
    const int type = A->mass > B->mass ? A->type : B->type;                                 /* idx:403 */
    obody m = make_body(w, newMass, newPos, newVel, type, newTemp);
    blist_push(newBodies, &m);
    w->n_flash++; /* visuals.triggerFlash, idx:408 */
    w->n_merge++;
}

/* PhysicsSystem.fractureBodies, idx:411-443 */
static void fracture_bodies(oracle_world *w, const obody *A, const obody *B, double energy,
                            blist *newBodies)
{
    const double totalMass = A->mass + B->mass;
    const double invT = 1 / totalMass;
    double com[3], baseVel[3];
    for (int k = 0; k < 3; k++) {
        com[k] = (A->pos[k] * A->mass + B->pos[k] * B->mass) * invT;     /* idx:413 */
        baseVel[k] = (A->vel[k] * A->mass + B->vel[k] * B->mass) * invT; /* idx:414-415 */
    }
    const double severity = fmin(energy / w->fractureThreshold, 5.0);   /* idx:417 */
    const int numFragments = (int)floor(3 + rnd(w) * 3 * severity);     /* idx:418 */
    double remainingMass = totalMass;

    for (int k = 0; k < numFragments; k++) {
        if (remainingMass < w->minFragmentMass) break;                  /* idx:422 */
        double fragMass = remainingMass * (0.3 + rnd(w) * 0.4);         /* idx:424 */
        if (k == numFragments - 1) fragMass = remainingMass;            /* idx:425 */
        if (fragMass < w->minFragmentMass) continue;                    /* idx:427 */
        remainingMass -= fragMass;

        double dir[3];
        dir[0] = rnd(w) - 0.5;                                          /* idx:430 */
        dir[1] = rnd(w) - 0.5;
        dir[2] = rnd(w) - 0.5;
        double len = sqrt(dir[0] * dir[0] + dir[1] * dir[1] + dir[2] * dir[2]);
        if (len == 0) len = 1;
        const double s = 1 / len;
        dir[0] *= s;
        dir[1] *= s;
        dir[2] *= s;
        const double off = (A->radius + B->radius) * 0.5 * rnd(w);      /* idx:431 */
        double fragPos[3], fragVel[3];
        for (int q = 0; q < 3; q++) fragPos[q] = com[q] + dir[q] * off; /* idx:432 */
        const double ejectSpeed = sqrt(energy / totalMass) * (0.5 + rnd(w)); /* idx:433 */
        for (int q = 0; q < 3; q++) fragVel[q] = baseVel[q] + dir[q] * ejectSpeed; /* idx:434 */
        const double fragTemp = fmax(A->temperature, B->temperature) + (energy / totalMass) * 0.1; /* idx:435 */
        const int type = A->mass > B->mass ? A->type : B->type;
        obody f = make_body(w, fragMass, fragPos, fragVel, type, fragTemp);
        blist_push(newBodies, &f);
        w->n_explosion++; /* visuals.spawnExplosion, idx:441 */
    }
    w->n_fracture++;
}

/* PhysicsSystem.resolveCollisions(dt), idx:293-390 */
ORACLE_API void oracle_world_resolve_collisions(oracle_world *w, double dt)
{
    blist newBodies = {0, 0, 0};
    const size_t N = w->n;
    unsigned char *removed = (unsigned char *)calloc(N ? N : 1, 1);
    size_t n_removed = 0;
    for (size_t k = 0; k < w->nc; k++) w->c[k].seen = 0;

    for (size_t i = 0; i < N; i++) {
        if (removed[i]) continue;                                        /* idx:302 */
        for (size_t j = i + 1; j < N; j++) {
            if (removed[j]) continue;                                    /* idx:304 */
            obody *A = &w->b[i], *B = &w->b[j];
            double vec[3];
            for (int k = 0; k < 3; k++) vec[k] = B->pos[k] - A->pos[k];  /* idx:309 */
            const double distSq = vec[0] * vec[0] + vec[1] * vec[1] + vec[2] * vec[2]; /* idx:310 */
            const double minDist = A->radius + B->radius;                /* idx:311 */
            if (!(distSq < minDist * minDist)) continue;                 /* idx:313 */
            w->n_overlap++;

            ocontact *e = contact_find(w, A->id, B->id);
            const double prev = e && e->time ? e->time : 0;              /* get(pairId) || 0 */
            const double newContactTime = prev + dt;                     /* idx:318 */
            e = contact_set(w, A->id, B->id, newContactTime);
            e->seen = 1;

            const double dist = sqrt(distSq);                            /* idx:321 */
            double normal[3], relVel[3];
            {
                double len = sqrt(vec[0] * vec[0] + vec[1] * vec[1] + vec[2] * vec[2]);
                if (len == 0) len = 1;                                   /* length() || 1 */
                const double s = 1 / len;
                for (int k = 0; k < 3; k++) normal[k] = vec[k] * s;      /* idx:322 */
            }
            for (int k = 0; k < 3; k++) relVel[k] = B->vel[k] - A->vel[k]; /* idx:324 */
            const double van = relVel[0] * normal[0] + relVel[1] * normal[1] + relVel[2] * normal[2];
            if (!(van < 0)) continue;                                    /* idx:327 */

            const double restitution = 0.2;
            const double invSum = A->invMass + B->invMass;
            const double jImpulse = -(1 + restitution) * van / invSum;   /* idx:329 */
            double impulse[3];
            for (int k = 0; k < 3; k++) impulse[k] = normal[k] * jImpulse;

            const double mu = (A->mass * B->mass) / (A->mass + B->mass); /* idx:332 */
            const double impactEnergy = 0.5 * mu * van * van;            /* idx:333 */
            A->temperature += (impactEnergy / A->mass) * 0.2;            /* idx:335 */
            B->temperature += (impactEnergy / B->mass) * 0.2;
            const double Q = impactEnergy / (A->mass + B->mass);         /* idx:338 */

            if (newContactTime > w->mergeTime && Q < w->fractureThreshold * 2) { /* idx:340 */
                merge_bodies(w, A, B, &newBodies);
                if (!removed[i]) { removed[i] = 1; n_removed++; }
                if (!removed[j]) { removed[j] = 1; n_removed++; }
                contact_delete(w, contact_find(w, A->id, B->id));
                continue; /* the row of A goes on even though A is now removed, idx:345 */
            }

            const int isFracture = Q > w->fractureThreshold;             /* idx:348 */
            const double cs = (minDist - dist) / invSum * 0.8;           /* idx:350 */
            for (int k = 0; k < 3; k++) {
                const double corr = normal[k] * cs;
                A->pos[k] -= corr * A->invMass;                          /* idx:351 */
                B->pos[k] += corr * B->invMass;                          /* idx:352 */
            }

            if (isFracture && (A->mass > w->minFragmentMass || B->mass > w->minFragmentMass)) {
                fracture_bodies(w, A, B, impactEnergy, &newBodies);      /* idx:355 */
                if (!removed[i]) { removed[i] = 1; n_removed++; }
                if (!removed[j]) { removed[j] = 1; n_removed++; }
                w->n_flash++;                                            /* idx:358 */
                contact_delete(w, contact_find(w, A->id, B->id));
            } else {
                double tangent[3];
                for (int k = 0; k < 3; k++) {
                    A->vel[k] -= impulse[k] * A->invMass;                /* idx:361 */
                    B->vel[k] += impulse[k] * B->invMass;                /* idx:362 */
                }
                for (int k = 0; k < 3; k++) tangent[k] = relVel[k] - normal[k] * van; /* idx:364 */
                double len = sqrt(tangent[0] * tangent[0] + tangent[1] * tangent[1] + tangent[2] * tangent[2]);
                if (len == 0) len = 1;
                const double s = 1 / len;
                for (int k = 0; k < 3; k++) tangent[k] *= s;
                const double rt = relVel[0] * tangent[0] + relVel[1] * tangent[1] + relVel[2] * tangent[2];
                const double jt = -rt * 0.5 / invSum;                    /* idx:365 */
                for (int k = 0; k < 3; k++) {
                    const double ti = tangent[k] * jt;                   /* idx:366 */
                    A->vel[k] -= ti * A->invMass;                        /* idx:368 */
                    B->vel[k] += ti * B->invMass;                        /* idx:369 */
                }
            }
        }
    }

    for (size_t k = 0; k < w->nc;) {                                     /* idx:376-380 */
        if (!w->c[k].seen) contact_delete(w, &w->c[k]);
        else k++;
    }

    if (n_removed > 0) {                                                 /* idx:382-385 */
        size_t o = 0;
        for (size_t i = 0; i < N; i++)
            if (!removed[i]) w->b[o++] = w->b[i];
        w->n = o;
    }
    for (size_t k = 0; k < newBodies.n; k++) add_body(w, &newBodies.v[k]); /* idx:387-389 */
    free(newBodies.v);
    free(removed);
}

/* PhysicsSystem.integrate(dt), idx:247-262, COMPLETE (collisions included). */
ORACLE_API void oracle_world_integrate(oracle_world *w, double dt, double G, double softening,
                                       int collisions, int nsteps)
{
    for (int s = 0; s < nsteps; s++) {
        const double halfDt = 0.5 * dt;
        for (size_t i = 0; i < w->n; i++) {                              /* idx:250-253 */
            obody *o = &w->b[i];
            for (int k = 0; k < 3; k++) o->vel[k] += o->acc[k] * halfDt;
            for (int k = 0; k < 3; k++) o->pos[k] += o->vel[k] * dt;
        }
        {                                                                /* idx:255 */
            const size_t n = w->n;
            double *pm = (double *)malloc(4 * (n ? n : 1) * sizeof(double));
            double *acc = (double *)malloc(4 * (n ? n : 1) * sizeof(double));
            for (size_t i = 0; i < n; i++) {
                memcpy(pm + 4 * i, w->b[i].pos, 3 * sizeof(double));
                pm[4 * i + 3] = w->b[i].mass;
            }
            oracle_compute_gravity(n, pm, acc, G, softening);
            for (size_t i = 0; i < n; i++) memcpy(w->b[i].acc, acc + 4 * i, 3 * sizeof(double));
            free(pm);
            free(acc);
        }
        if (collisions) oracle_world_resolve_collisions(w, dt);          /* idx:256 */
        for (size_t i = 0; i < w->n; i++) {                              /* idx:258-261 */
            obody *o = &w->b[i];
            for (int k = 0; k < 3; k++) o->vel[k] += o->acc[k] * halfDt;
            o->temperature *= w->heatDecay;                              /* idx:228 */
            if (o->temperature < 0.1) o->temperature = 0;                /* idx:229 */
        }
    }
}

/* Broad phase alone: every pair i<j, in the reference's visiting order
 * (idx:301-305), whose spheres overlap by the reference's test (idx:309-313)
 * widened by `margin`: distSq < (r_i + r_j + margin)^2.  Returns the number
 * of pairs found; writes at most cap of them as (i,j) uint32 pairs. */
ORACLE_API size_t oracle_overlap_pairs(size_t n, const double *pos_mass, const double *radius,
                                       double margin, uint32_t *pairs, size_t cap)
{
    size_t cnt = 0;
    for (size_t i = 0; i < n; i++) {
        const double *A = pos_mass + 4 * i;
        for (size_t j = i + 1; j < n; j++) {
            const double *B = pos_mass + 4 * j;
            const double vx = B[0] - A[0], vy = B[1] - A[1], vz = B[2] - A[2];
            const double distSq = vx * vx + vy * vy + vz * vz;
            const double minDist = radius[i] + radius[j] + margin;
            if (distSq < minDist * minDist) {
                if (cnt < cap) {
                    pairs[2 * cnt] = (uint32_t)i;
                    pairs[2 * cnt + 1] = (uint32_t)j;
                }
                cnt++;
            }
        }
    }
    return cnt;
}

/* Same set, OpenMP over rows (courtesy, for mid-size scenes); rows are
 * concatenated in order so the output order is unchanged. */
ORACLE_API size_t oracle_overlap_pairs_mt(size_t n, const double *pos_mass, const double *radius,
                                          double margin, uint32_t *pairs, size_t cap)
{
    size_t *row_cnt = (size_t *)calloc(n + 1, sizeof(size_t));
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic, 64)
#endif
    for (long i = 0; i < (long)n; i++) {
        const double *A = pos_mass + 4 * i;
        size_t c = 0;
        for (size_t j = (size_t)i + 1; j < n; j++) {
            const double *B = pos_mass + 4 * j;
            const double vx = B[0] - A[0], vy = B[1] - A[1], vz = B[2] - A[2];
            const double distSq = vx * vx + vy * vy + vz * vz;
            const double minDist = radius[i] + radius[j] + margin;
            if (distSq < minDist * minDist) c++;
        }
        row_cnt[i + 1] = c;
    }
    for (size_t i = 0; i < n; i++) row_cnt[i + 1] += row_cnt[i];
    const size_t total = row_cnt[n];
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic, 64)
#endif
    for (long i = 0; i < (long)n; i++) {
        if (row_cnt[i + 1] == row_cnt[i]) continue;
        const double *A = pos_mass + 4 * i;
        size_t o = row_cnt[i];
        for (size_t j = (size_t)i + 1; j < n; j++) {
            const double *B = pos_mass + 4 * j;
            const double vx = B[0] - A[0], vy = B[1] - A[1], vz = B[2] - A[2];
            const double distSq = vx * vx + vy * vy + vz * vz;
            const double minDist = radius[i] + radius[j] + margin;
            if (distSq < minDist * minDist) {
                if (o < cap) {
                    pairs[2 * o] = (uint32_t)i;
                    pairs[2 * o + 1] = (uint32_t)j;
                }
                o++;
            }
        }
    }
    free(row_cnt);
    return total;
}
