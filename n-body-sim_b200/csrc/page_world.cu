// The page's COMPLETE step on the device, for the page's own sizes (response included):
// PhysicsSystem.integrate(dt) with resolveCollisions, mergeBodies and fractureBodies
// (/root/reference/index.html:247-262, 293-443; "idx:NNN" below), the Body constructor (idx:208-225) and addBody
// (idx:239-245) -- one CTA, all state in shared memory, nsteps calls per launch, nothing returns to the host in between.
//
// This translation unit is compiled with -fmad=false: every statement of the collision stage is the reference's own
// association evaluated in IEEE binary64 (+ - * / sqrt are correctly rounded on the device), so given the same input state
// the stage is BIT-IDENTICAL to the test suite's CPU restatement (itself pinned to the page's executed text) -- ids,
// random draws, merges, fragments and contact times included.  The one library function of the stage, Math.pow(x, 1/3) in
// the Body constructor (idx:221), is evaluated to a correctly rounded result in double-double arithmetic (pow_third below;
// the device's own pow() is a 2-ulp routine).  computeGravity inside the step is this repository's fp64 pair arithmetic
// (force_f64.cuh MATH 2, explicit fused multiply-adds), as everywhere else: 1e-12 of the reference, not bit-identical.
//
// The reference's double loop is order dependent: a response moves the two bodies (idx:351-352), and a later overlap test
// sees the moved positions.  The kernel keeps the reference's order exactly without walking all N^2/2 pairs one by one:
//   1. all pairs that overlap NOW are found in parallel (one warp per row) and listed in the reference's visiting order;
//   2. the list is walked in order; every entry is RE-TESTED with the positions of that moment (all threads evaluate the
//      same shared-memory values, so the walk needs no barrier until a pair really touches);
//   3. whenever a response has moved a body, that body is tested in parallel against all others, and every pair that
//      overlaps now and whose turn is still to come joins the list at its place.
// A pair that overlaps when its turn comes either overlapped at the start or started to after the last move of one of its
// bodies before its turn -- both are in the list -- so the walk visits exactly the pairs the reference's loop acts on.
#include "capi_ctx.cuh"

#include "force_f64.cuh"

namespace {

constexpr int kWorldThreads = 1024;
constexpr int kWorldMax = NBODY_WORLD_MAX_BODIES;  // bodies (shared-memory arrays)
constexpr int kWorldMaxContacts = 2048;            // PhysicsSystem.contactPairs entries
constexpr int kWorldMaxCand = 2048;                // overlapping pairs per resolveCollisions call
constexpr int kWorldMaxNew = 8192;                 // bodies created by one call (merge: 1, fracture: <= 18 per event)

struct NewBody {
    double mass, pos[3], vel[3], temp, radius, id;
    int type;
};

struct WorldParams {
    double4* pos[2];  // the context's position buffers (x, y, z, m); [0] is current on entry and on exit
    double4* vel;
    double4* acc;
    double* temp;
    double* radius;
    double* id;
    int* type;
    double* c_lo;     // contactPairs: unordered id pair, accumulated time
    double* c_hi;
    double* c_time;
    NewBody* fresh;   // scratch
    int* state;       // [0] n, [1] contacts, [2] error (1 contacts full, 2 candidate list full, 3 new-body scratch full, 4 too many bodies)
    unsigned long long* rng;
    long long* counters;  // merges, fractures, overlaps, flashes, explosions (visuals.* calls are counted, not executed)
    int cap_rows;     // rows of the context's arrays to keep consistent (pads beyond n)
    int max_bodies;   // CONFIG.maxBodies, idx:57
    int nsteps, collisions;
    double dt, half_dt, G, eps2, heat_decay;
    double fracture_threshold, min_fragment_mass, merge_time;  // idx:62-64
    double pad_x;
    double* rec;      // nbody_world_frame: the frame's record in pinned, device-mapped host memory (or null)
};

// The record nbody_world_frame reads after ONE synchronise: sixteen 64-bit words (n, contacts, error, the random stream's
// state, the five counters), then pos_mass[4n] | vel[4n] | acc[4n] | temp[n] | id[n] | radius[n] | type[n] (as doubles).
constexpr int kWorldRecHeader = 16;
constexpr size_t world_rec_bytes(size_t rows) { return (kWorldRecHeader + 16 * rows) * sizeof(double); }

struct Smem {
    double px[kWorldMax], py[kWorldMax], pz[kWorldMax];
    double vx[kWorldMax], vy[kWorldMax], vz[kWorldMax];
    double ax[kWorldMax], ay[kWorldMax], az[kWorldMax];
    double m[kWorldMax], rad[kWorldMax], temp[kWorldMax], id[kWorldMax];
    double c_lo[kWorldMaxContacts], c_hi[kWorldMaxContacts], c_time[kWorldMaxContacts];
    uint2 cand[kWorldMaxCand];
    int type[kWorldMax];
    int row_count[kWorldMax];
    unsigned char removed[kWorldMax];
    unsigned char c_seen[kWorldMaxContacts];
    int n, nc, ncand, n_new, err, found, n_add;
    int kept_total, changed;
    int warp_tot[32];
    uint2 add[64];
    unsigned long long rng;
    long long counters[5];
    double part[3 * kWorldThreads];  // computeGravity: partial sums of the threads that share a body
};

// Math.random stand-in: sequential splitmix64 -> [0,1) with 53 bits (the stream the test oracles draw from).
__device__ double rnd(Smem& S)
{
    unsigned long long z = (S.rng += 0x9E3779B97F4A7C15ULL);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    z ^= z >> 31;
    return double(z >> 11) * (1.0 / 9007199254740992.0);
}

// Math.pow(x, 1/3) as the page evaluates it: the exponent is the double nearest 1/3, c = 1/3 - 2^-54/3, so
// x^c = cbrt(x) * exp(-(2^-54/3) ln x).  cbrt(x) to ~2^-100 by one Newton step in double-double from the library seed,
// the correction factor (|.| < 1e-14) to first order with the second-order term added; one final rounding.
__device__ double pow_third(double x)
{
    if (!(x > 0.0)) return x == 0.0 ? 0.0 : __longlong_as_double(0x7ff8000000000000LL);
    if (!(x < 1.7976931348623157e308)) return x;
    const double y0 = cbrt(x);
    const double ph = y0 * y0, pl = __fma_rn(y0, y0, -ph);                 // y0^2 = ph + pl exactly
    const double qh = ph * y0, ql = __fma_rn(ph, y0, -qh) + pl * y0;       // y0^3 = qh + ql (to 2^-104)
    const double r = (x - qh) - ql;                                         // x - y0^3
    const double yl = r / (3.0 * ph);                                       // Newton: cbrt(x) = y0 + yl
    const double e = -(1.0 / 54043195528445952.0) * log(x);                 // -(2^-54 / 3) ln x
    const double t = yl + y0 * (e + 0.5 * e * e);
    return y0 + t;
}

__device__ __forceinline__ bool overlap(const Smem& S, int i, int j)
{
    NB_CHECK(i >= 0 && j >= 0 && i < S.n && j < S.n && S.n <= kWorldMax);
    const double vx = S.px[j] - S.px[i], vy = S.py[j] - S.py[i], vz = S.pz[j] - S.pz[i];  // vec.subVectors(B.pos, A.pos), idx:309
    const double distSq = vx * vx + vy * vy + vz * vz;                                    // idx:310
    const double minDist = S.rad[i] + S.rad[j];                                           // idx:311
    return distSq < minDist * minDist;                                                    // idx:313
}

// new Body(mass, pos, vel, type, temp), idx:209-225, appended to the call's newBodies list
__device__ void make_body(Smem& S, const WorldParams& p, double mass, const double* pos, const double* vel, int type, double temp)
{
    if (S.n_new >= kWorldMaxNew) {
        S.err = 3;
        return;
    }
    NewBody& o = p.fresh[S.n_new++];
    o.id = rnd(S);  // idx:210
    o.mass = mass;
    for (int k = 0; k < 3; ++k) {
        o.pos[k] = pos[k];
        o.vel[k] = vel[k];
    }
    o.type = type;
    const double density = type == 1 ? 3.0 : (type == 2 ? 0.5 : 1.0);          // CONFIG.materials, idx:70-74
    o.radius = pow_third((3 * mass) / (4 * 3.141592653589793 * density));      // idx:221
    o.temp = temp;
}

// PhysicsSystem.mergeBodies, idx:392-409
__device__ void merge_bodies(Smem& S, const WorldParams& p, int a, int b)
{
    const double newMass = S.m[a] + S.m[b];
    const double inv = 1 / newMass;
    const double va[3] = {S.vx[a], S.vy[a], S.vz[a]}, vb[3] = {S.vx[b], S.vy[b], S.vz[b]};
    const double pa[3] = {S.px[a], S.py[a], S.pz[a]}, pb[3] = {S.px[b], S.py[b], S.pz[b]};
    double newVel[3], newPos[3];
    for (int k = 0; k < 3; ++k) {
        newVel[k] = (va[k] * S.m[a] + vb[k] * S.m[b]) * inv;  // idx:394-396
        newPos[k] = (pa[k] * S.m[a] + pb[k] * S.m[b]) * inv;  // idx:398-400
    }
    const double newTemp = (S.temp[a] * S.m[a] + S.temp[b] * S.m[b]) / newMass;  // idx:402
    const int type = S.m[a] > S.m[b] ? S.type[a] : S.type[b];                     // idx:403
    make_body(S, p, newMass, newPos, newVel, type, newTemp);
    S.counters[3]++;  // visuals.triggerFlash, idx:408
    S.counters[0]++;
}

// PhysicsSystem.fractureBodies, idx:411-443
__device__ void fracture_bodies(Smem& S, const WorldParams& p, int a, int b, double energy)
{
    const double totalMass = S.m[a] + S.m[b];
    const double invT = 1 / totalMass;
    const double va[3] = {S.vx[a], S.vy[a], S.vz[a]}, vb[3] = {S.vx[b], S.vy[b], S.vz[b]};
    const double pa[3] = {S.px[a], S.py[a], S.pz[a]}, pb[3] = {S.px[b], S.py[b], S.pz[b]};
    double com[3], baseVel[3];
    for (int k = 0; k < 3; ++k) {
        com[k] = (pa[k] * S.m[a] + pb[k] * S.m[b]) * invT;      // idx:413
        baseVel[k] = (va[k] * S.m[a] + vb[k] * S.m[b]) * invT;  // idx:414-415
    }
    const double severity = fmin(energy / p.fracture_threshold, 5.0);  // idx:417
    const int numFragments = int(floor(3 + rnd(S) * 3 * severity));    // idx:418
    double remainingMass = totalMass;
    for (int k = 0; k < numFragments; ++k) {
        if (remainingMass < p.min_fragment_mass) break;          // idx:422
        double fragMass = remainingMass * (0.3 + rnd(S) * 0.4);  // idx:424
        if (k == numFragments - 1) fragMass = remainingMass;     // idx:425
        if (fragMass < p.min_fragment_mass) continue;            // idx:427
        remainingMass -= fragMass;
        double dir[3];
        dir[0] = rnd(S) - 0.5;  // idx:430
        dir[1] = rnd(S) - 0.5;
        dir[2] = rnd(S) - 0.5;
        double len = sqrt(dir[0] * dir[0] + dir[1] * dir[1] + dir[2] * dir[2]);
        if (len == 0) len = 1;
        const double s = 1 / len;
        dir[0] *= s;
        dir[1] *= s;
        dir[2] *= s;
        const double off = (S.rad[a] + S.rad[b]) * 0.5 * rnd(S);  // idx:431
        double fragPos[3], fragVel[3];
        for (int q = 0; q < 3; ++q) fragPos[q] = com[q] + dir[q] * off;            // idx:432
        const double ejectSpeed = sqrt(energy / totalMass) * (0.5 + rnd(S));       // idx:433
        for (int q = 0; q < 3; ++q) fragVel[q] = baseVel[q] + dir[q] * ejectSpeed;  // idx:434
        const double fragTemp = fmax(S.temp[a], S.temp[b]) + (energy / totalMass) * 0.1;  // idx:435
        const int type = S.m[a] > S.m[b] ? S.type[a] : S.type[b];
        make_body(S, p, fragMass, fragPos, fragVel, type, fragTemp);
        S.counters[4]++;  // visuals.spawnExplosion, idx:441
    }
    S.counters[1]++;
}

__device__ void contact_delete(Smem& S, int k)
{
    NB_CHECK(k >= 0 && k < S.nc && S.nc <= kWorldMaxContacts);
    --S.nc;  // Map order has no effect on results
    S.c_lo[k] = S.c_lo[S.nc];
    S.c_hi[k] = S.c_hi[S.nc];
    S.c_time[k] = S.c_time[S.nc];
    S.c_seen[k] = S.c_seen[S.nc];
}

// The body of the reference's `if (distSq < minDist * minDist)` for pair (i, j), idx:314-372; one thread.
// `entry` = index of the pair's contactPairs entry or -1.  Returns true when the positional correction moved the two bodies.
__device__ bool respond(Smem& S, const WorldParams& p, int i, int j, int entry)
{
    S.counters[2]++;
    const double vec[3] = {S.px[j] - S.px[i], S.py[j] - S.py[i], S.pz[j] - S.pz[i]};
    const double distSq = vec[0] * vec[0] + vec[1] * vec[1] + vec[2] * vec[2];
    const double minDist = S.rad[i] + S.rad[j];
    const double prev = entry >= 0 && S.c_time[entry] != 0.0 ? S.c_time[entry] : 0;  // get(pairId) || 0, idx:317
    const double newContactTime = prev + p.dt;                                        // idx:318
    if (entry < 0) {
        if (S.nc >= kWorldMaxContacts) {
            S.err = 1;
            return false;
        }
        entry = S.nc++;
        const double ida = S.id[i], idb = S.id[j];
        S.c_lo[entry] = ida < idb ? ida : idb;  // [A.id, B.id].sort().join('-'), idx:314
        S.c_hi[entry] = ida < idb ? idb : ida;
    }
    S.c_time[entry] = newContactTime;  // idx:319
    S.c_seen[entry] = 1;               // currentFrameContacts.add, idx:315

    const double dist = sqrt(distSq);  // idx:321
    double normal[3], relVel[3];
    {
        double len = sqrt(vec[0] * vec[0] + vec[1] * vec[1] + vec[2] * vec[2]);
        if (len == 0) len = 1;  // length() || 1
        const double s = 1 / len;
        for (int k = 0; k < 3; ++k) normal[k] = vec[k] * s;  // idx:322
    }
    relVel[0] = S.vx[j] - S.vx[i];  // idx:324
    relVel[1] = S.vy[j] - S.vy[i];
    relVel[2] = S.vz[j] - S.vz[i];
    const double van = relVel[0] * normal[0] + relVel[1] * normal[1] + relVel[2] * normal[2];  // idx:325
    if (!(van < 0)) return false;                                                              // idx:327

    const double invA = S.m[i] > 0 ? 1 / S.m[i] : 0, invB = S.m[j] > 0 ? 1 / S.m[j] : 0;  // Body.invMass, idx:212
    const double restitution = 0.2;
    const double invSum = invA + invB;
    const double jImpulse = -(1 + restitution) * van / invSum;  // idx:329
    double impulse[3];
    for (int k = 0; k < 3; ++k) impulse[k] = normal[k] * jImpulse;
    const double mu = (S.m[i] * S.m[j]) / (S.m[i] + S.m[j]);  // idx:332
    const double impactEnergy = 0.5 * mu * van * van;         // idx:333
    S.temp[i] += (impactEnergy / S.m[i]) * 0.2;               // idx:335
    S.temp[j] += (impactEnergy / S.m[j]) * 0.2;
    const double Q = impactEnergy / (S.m[i] + S.m[j]);        // idx:338

    if (newContactTime > p.merge_time && Q < p.fracture_threshold * 2) {  // idx:340
        merge_bodies(S, p, i, j);
        S.removed[i] = 1;
        S.removed[j] = 1;
        contact_delete(S, entry);
        return false;  // (the row of A goes on even though A is now removed, idx:345)
    }

    const bool isFracture = Q > p.fracture_threshold;    // idx:348
    const double cs = (minDist - dist) / invSum * 0.8;   // idx:350
    {
        const double cx = normal[0] * cs, cy = normal[1] * cs, cz = normal[2] * cs;
        S.px[i] -= cx * invA;  // idx:351
        S.py[i] -= cy * invA;
        S.pz[i] -= cz * invA;
        S.px[j] += cx * invB;  // idx:352
        S.py[j] += cy * invB;
        S.pz[j] += cz * invB;
    }
    if (isFracture && (S.m[i] > p.min_fragment_mass || S.m[j] > p.min_fragment_mass)) {
        fracture_bodies(S, p, i, j, impactEnergy);  // idx:355
        S.removed[i] = 1;
        S.removed[j] = 1;
        S.counters[3]++;  // idx:358
        contact_delete(S, entry);
    } else {
        S.vx[i] -= impulse[0] * invA;  // idx:361
        S.vy[i] -= impulse[1] * invA;
        S.vz[i] -= impulse[2] * invA;
        S.vx[j] += impulse[0] * invB;  // idx:362
        S.vy[j] += impulse[1] * invB;
        S.vz[j] += impulse[2] * invB;
        double tangent[3];
        for (int k = 0; k < 3; ++k) tangent[k] = relVel[k] - normal[k] * van;  // idx:364
        double len = sqrt(tangent[0] * tangent[0] + tangent[1] * tangent[1] + tangent[2] * tangent[2]);
        if (len == 0) len = 1;
        const double s = 1 / len;
        for (int k = 0; k < 3; ++k) tangent[k] *= s;
        const double rt = relVel[0] * tangent[0] + relVel[1] * tangent[1] + relVel[2] * tangent[2];
        const double jt = -rt * 0.5 / invSum;  // idx:365
        const double tx = tangent[0] * jt, ty = tangent[1] * jt, tz = tangent[2] * jt;  // idx:366
        S.vx[i] -= tx * invA;  // idx:368
        S.vy[i] -= ty * invA;
        S.vz[i] -= tz * invA;
        S.vx[j] += tx * invB;  // idx:369
        S.vy[j] += ty * invB;
        S.vz[j] += tz * invB;
    }
    return true;
}

__device__ __forceinline__ bool pair_after(uint2 a, int i, int j) { return int(a.x) > i || (int(a.x) == i && int(a.y) > j); }

// PhysicsSystem.resolveCollisions(dt), idx:293-390; the whole CTA.
__device__ void resolve_collisions(Smem& S, const WorldParams& p)
{
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int N = S.n;
    if (tid < N) S.removed[tid] = 0;
    for (int k = tid; k < S.nc; k += kWorldThreads) S.c_seen[k] = 0;
    if (tid == 0) {
        S.ncand = 0;
        S.n_new = 0;
    }
    __syncthreads();

    // 1. the pairs that overlap now, in the reference's visiting order: one warp per row, two passes (count, fill)
    for (int i = warp; i < N; i += kWorldThreads / 32) {
        int cnt = 0;
        for (int j0 = i + 1; j0 < N; j0 += 32) {
            const int j = j0 + lane;
            cnt += __popc(__ballot_sync(0xffffffffu, j < N && overlap(S, i, j)));
        }
        if (lane == 0) S.row_count[i] = cnt;
    }
    __syncthreads();
    {  // exclusive prefix over the rows: warp scans, then a scan of the 32 warp totals
        const int own = tid < N ? S.row_count[tid] : 0;
        int v = own;
        for (int o = 1; o < 32; o <<= 1) {
            const int u = __shfl_up_sync(0xffffffffu, v, o);
            if (lane >= o) v += u;
        }
        if (lane == 31) S.warp_tot[warp] = v;
        __syncthreads();
        if (warp == 0) {
            int w = S.warp_tot[lane];
            for (int o = 1; o < 32; o <<= 1) {
                const int u = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o) w += u;
            }
            S.warp_tot[lane] = w;
        }
        __syncthreads();
        if (tid < N) S.row_count[tid] = v - own + (warp > 0 ? S.warp_tot[warp - 1] : 0);
        if (tid == 0) {
            S.ncand = S.warp_tot[31];
            if (S.ncand > kWorldMaxCand) S.err = 2;
        }
    }
    __syncthreads();
    if (S.err) return;
    for (int i = warp; i < N; i += kWorldThreads / 32) {
        int at = S.row_count[i];
        for (int j0 = i + 1; j0 < N; j0 += 32) {
            const int j = j0 + lane;
            const unsigned hit = __ballot_sync(0xffffffffu, j < N && overlap(S, i, j));
            NB_CHECK(at + __popc(hit) <= kWorldMaxCand);
            if (hit & (1u << lane)) S.cand[at + __popc(hit & ((1u << lane) - 1u))] = make_uint2(unsigned(i), unsigned(j));
            at += __popc(hit);
        }
    }
    __syncthreads();

    // 2. the walk.  Every thread reads the same shared values, so the branches below are uniform.
    int row = -1;
    bool row_skip = false;
    for (int c = 0; c < S.ncand; ++c) {
        const int i = int(S.cand[c].x), j = int(S.cand[c].y);
        if (i != row) {
            row = i;
            row_skip = S.removed[i] != 0;  // `if (removedIndices.has(i)) continue;` is evaluated when the row starts, idx:302
        }
        if (row_skip || S.removed[j]) continue;  // idx:304
        if (!overlap(S, i, j)) continue;         // idx:313, with the positions of this moment
        // contactPairs.get(pairId): all threads look
        __syncthreads();
        if (tid == 0) S.found = -1;
        __syncthreads();
        {
            const double ida = S.id[i], idb = S.id[j];
            const double lo = ida < idb ? ida : idb, hi = ida < idb ? idb : ida;
            for (int k = tid; k < S.nc; k += kWorldThreads)
                if (S.c_lo[k] == lo && S.c_hi[k] == hi) S.found = k;
        }
        __syncthreads();
        if (tid == 0) {
            S.n_add = respond(S, p, i, j, S.found) ? 0 : -1;
        }
        __syncthreads();
        if (S.err) return;
        if (S.n_add < 0) continue;  // nothing moved
        // 3. the two bodies moved: every pair with one of them that overlaps now and is still to come joins the list
        for (int y = tid; y < N; y += kWorldThreads) {
            for (int w = 0; w < 2; ++w) {
                const int x = w ? j : i;
                if (y == x || S.removed[y] || (w && y == i)) continue;  // ((i, j) itself is behind us; (j, i) is the same pair)
                const uint2 pr = make_uint2(unsigned(x < y ? x : y), unsigned(x < y ? y : x));
                if (!pair_after(pr, i, j) || !overlap(S, int(pr.x), int(pr.y))) continue;
                const int k = atomicAdd(&S.n_add, 1);
                if (k < 64) S.add[k] = pr;
            }
        }
        __syncthreads();
        if (tid == 0) {
            if (S.n_add > 64) S.err = 2;
            for (int a = 0; a < S.n_add && !S.err; ++a) {  // sorted insert behind the cursor, no duplicates (a handful per call)
                const uint2 pr = S.add[a];
                int at = c + 1;
                while (at < S.ncand && (S.cand[at].x < pr.x || (S.cand[at].x == pr.x && S.cand[at].y < pr.y))) ++at;
                if (at < S.ncand && S.cand[at].x == pr.x && S.cand[at].y == pr.y) continue;
                if (S.ncand >= kWorldMaxCand) {
                    S.err = 2;
                    break;
                }
                for (int q = S.ncand; q > at; --q) S.cand[q] = S.cand[q - 1];
                S.cand[at] = pr;
                S.ncand++;
            }
        }
        __syncthreads();
        if (S.err) return;
    }
    __syncthreads();

    // contacts not seen this call go (idx:376-380)
    if (tid == 0) {
        for (int k = 0; k < S.nc;) {
            if (!S.c_seen[k]) contact_delete(S, k);
            else ++k;
        }
    }
    __syncthreads();

    // bodies.filter (idx:382-385), then addBody for every new body with the 300-body shift (idx:239-245, 387-389):
    // the list becomes the LAST min(total, maxBodies) entries of (kept bodies ++ new bodies)
    const int kept_total = __syncthreads_count(tid < N && !S.removed[tid]);
    if (kept_total == N && S.n_new == 0) return;  // (uniform: nothing was removed or created)
    int kept_before = 0, keep = 0;
    double row13[13];
    int rtype = 0;
    if (tid < N) {
        keep = !S.removed[tid];
        for (int k = 0; k < tid; ++k) kept_before += !S.removed[k];
    }
    const int total = kept_total + S.n_new;
    const int drop = total > p.max_bodies ? total - p.max_bodies : 0;
    if (total - drop > kWorldMax) {
        if (tid == 0) S.err = 4;
        __syncthreads();
        return;
    }
    if (tid < N && keep) {
        row13[0] = S.px[tid]; row13[1] = S.py[tid]; row13[2] = S.pz[tid];
        row13[3] = S.vx[tid]; row13[4] = S.vy[tid]; row13[5] = S.vz[tid];
        row13[6] = S.ax[tid]; row13[7] = S.ay[tid]; row13[8] = S.az[tid];
        row13[9] = S.m[tid]; row13[10] = S.rad[tid]; row13[11] = S.temp[tid]; row13[12] = S.id[tid];
        rtype = S.type[tid];
    }
    __syncthreads();
    if (tid < N && keep && kept_before >= drop) {
        const int d = kept_before - drop;
        S.px[d] = row13[0]; S.py[d] = row13[1]; S.pz[d] = row13[2];
        S.vx[d] = row13[3]; S.vy[d] = row13[4]; S.vz[d] = row13[5];
        S.ax[d] = row13[6]; S.ay[d] = row13[7]; S.az[d] = row13[8];
        S.m[d] = row13[9]; S.rad[d] = row13[10]; S.temp[d] = row13[11]; S.id[d] = row13[12];
        S.type[d] = rtype;
    }
    for (int k = tid; k < S.n_new; k += kWorldThreads) {
        const int at = kept_total + k - drop;
        if (at < 0) continue;
        NB_CHECK(at < kWorldMax);
        const NewBody o = p.fresh[k];
        S.px[at] = o.pos[0]; S.py[at] = o.pos[1]; S.pz[at] = o.pos[2];
        S.vx[at] = o.vel[0]; S.vy[at] = o.vel[1]; S.vz[at] = o.vel[2];
        S.ax[at] = 0.0; S.ay[at] = 0.0; S.az[at] = 0.0;  // this.acc = new THREE.Vector3(0, 0, 0), idx:217
        S.m[at] = o.mass; S.rad[at] = o.radius; S.temp[at] = o.temp; S.id[at] = o.id;
        S.type[at] = o.type;
    }
    __syncthreads();
    if (tid == 0) S.n = total - drop;
    __syncthreads();
}

// computeGravity (idx:264-291) for the bodies in shared memory: floor(1024 / n) threads share a body's j range, the partial
// sums are added in a fixed order.  MATH 2 pair arithmetic (force_f64.cuh).
__device__ void compute_gravity(Smem& S, const WorldParams& p, double* part /* [3][1024] */)
{
    const int tid = threadIdx.x, n = S.n;
    const int split = n > 0 ? kWorldThreads / n : 1;  // >= 1
    const int i = tid % (n > 0 ? n : 1), part_k = tid / (n > 0 ? n : 1);
    double sx = 0.0, sy = 0.0, sz = 0.0;
    if (n > 0 && part_k < split) {
        const int len = (n + split - 1) / split, j0 = part_k * len, j1 = j0 + len < n ? j0 + len : n;
        const double xi = S.px[i], yi = S.py[i], zi = S.pz[i];
        for (int j = j0; j < j1; ++j) {
            const double dx = S.px[j] - xi, dy = S.py[j] - yi, dz = S.pz[j] - zi;
            const double s = __fma_rn(dz, dz, __fma_rn(dy, dy, __fma_rn(dx, dx, p.eps2)));
            double w = nb::inv_cube_f64(s) * S.m[j];
            if (j == i) w = 0.0;  // the reference never evaluates the self pair (idx:273)
            sx = __fma_rn(w, dx, sx);
            sy = __fma_rn(w, dy, sy);
            sz = __fma_rn(w, dz, sz);
        }
        part[tid] = sx;
        part[kWorldThreads + tid] = sy;
        part[2 * kWorldThreads + tid] = sz;
    }
    __syncthreads();
    if (tid < n) {
        double tx = 0.0, ty = 0.0, tz = 0.0;
        for (int k = 0; k < split; ++k) {
            tx += part[k * n + tid];
            ty += part[kWorldThreads + k * n + tid];
            tz += part[2 * kWorldThreads + k * n + tid];
        }
        S.ax[tid] = tx * p.G;
        S.ay[tid] = ty * p.G;
        S.az[tid] = tz * p.G;
    }
    __syncthreads();
}

__global__ void __launch_bounds__(kWorldThreads, 1) world_step_kernel(const WorldParams p)
{
    extern __shared__ __align__(16) unsigned char raw[];
    Smem& S = *reinterpret_cast<Smem*>(raw);
    double* part = S.part;
    const int tid = threadIdx.x;
    if (tid == 0) {
        S.n = p.state[0];
        S.nc = p.state[1];
        S.err = 0;
        S.rng = *p.rng;
        for (int k = 0; k < 5; ++k) S.counters[k] = p.counters[k];
    }
    __syncthreads();
    if (tid < S.n) {
        const double4 q = p.pos[0][tid], v = p.vel[tid], a = p.acc[tid];
        S.px[tid] = q.x; S.py[tid] = q.y; S.pz[tid] = q.z; S.m[tid] = q.w;
        S.vx[tid] = v.x; S.vy[tid] = v.y; S.vz[tid] = v.z;
        S.ax[tid] = a.x; S.ay[tid] = a.y; S.az[tid] = a.z;
        S.temp[tid] = p.temp[tid];
        S.rad[tid] = p.radius[tid];
        S.id[tid] = p.id[tid];
        S.type[tid] = p.type[tid];
    }
    for (int k = tid; k < S.nc; k += kWorldThreads) {
        S.c_lo[k] = p.c_lo[k];
        S.c_hi[k] = p.c_hi[k];
        S.c_time[k] = p.c_time[k];
    }
    __syncthreads();

    for (int step = 0; step < p.nsteps && !S.err; ++step) {
        if (tid < S.n) {  // idx:250-253
            S.vx[tid] += S.ax[tid] * p.half_dt;
            S.vy[tid] += S.ay[tid] * p.half_dt;
            S.vz[tid] += S.az[tid] * p.half_dt;
            S.px[tid] += S.vx[tid] * p.dt;
            S.py[tid] += S.vy[tid] * p.dt;
            S.pz[tid] += S.vz[tid] * p.dt;
        }
        __syncthreads();
        compute_gravity(S, p, part);             // idx:255
        if (p.collisions) resolve_collisions(S, p);  // idx:256
        __syncthreads();
        if (S.err) break;
        if (tid < S.n) {  // idx:258-261
            S.vx[tid] += S.ax[tid] * p.half_dt;
            S.vy[tid] += S.ay[tid] * p.half_dt;
            S.vz[tid] += S.az[tid] * p.half_dt;
            double t = S.temp[tid] * p.heat_decay;  // Body.updateThermal, idx:227-230
            if (t < 0.1) t = 0;
            S.temp[tid] = t;
        }
        __syncthreads();
    }

    // back to the context's arrays; rows beyond n become pads again (massless, far away, every pad somewhere else)
    const int n = S.n;
    for (int r = tid; r < p.cap_rows; r += kWorldThreads) {
        double4 q, v, a;
        double t = 0.0, rd = 0.0, idv = 0.0;
        int ty = 0;
        if (r < n) {
            q = make_double4(S.px[r], S.py[r], S.pz[r], S.m[r]);
            v = make_double4(S.vx[r], S.vy[r], S.vz[r], 0.0);
            a = make_double4(S.ax[r], S.ay[r], S.az[r], 0.0);
            t = S.temp[r];
            rd = S.rad[r];
            idv = S.id[r];
            ty = S.type[r];
        } else {
            q = make_double4(p.pad_x * double(1 + (r - n)), 0.0, 0.0, 0.0);
            v = a = make_double4(0.0, 0.0, 0.0, 0.0);
        }
        p.pos[0][r] = q;
        p.pos[1][r] = q;
        p.vel[r] = v;
        p.acc[r] = a;
        p.temp[r] = t;
        p.radius[r] = rd;
        if (r < kWorldMax) {
            p.id[r] = idv;
            p.type[r] = ty;
        }
    }
    for (int k = tid; k < S.nc; k += kWorldThreads) {
        p.c_lo[k] = S.c_lo[k];
        p.c_hi[k] = S.c_hi[k];
        p.c_time[k] = S.c_time[k];
    }
    if (tid == 0) {
        p.state[0] = S.n;
        p.state[1] = S.nc;
        p.state[2] = S.err;
        *p.rng = S.rng;
        for (int k = 0; k < 5; ++k) p.counters[k] = S.counters[k];
    }
    if (p.rec) {  // what the caller reads each frame goes straight to its side of the bus: no copy to queue, one synchronise
        if (tid == 0) {
            long long* h = reinterpret_cast<long long*>(p.rec);
            h[0] = S.n;
            h[1] = S.nc;
            h[2] = S.err;
            h[3] = static_cast<long long>(S.rng);
            for (int k = 0; k < 5; ++k) h[4 + k] = S.counters[k];
        }
        double* b = p.rec + kWorldRecHeader;
        for (int r = tid; r < n; r += kWorldThreads) {
            reinterpret_cast<double4*>(b)[r] = make_double4(S.px[r], S.py[r], S.pz[r], S.m[r]);
            reinterpret_cast<double4*>(b + 4 * size_t(n))[r] = make_double4(S.vx[r], S.vy[r], S.vz[r], 0.0);
            reinterpret_cast<double4*>(b + 8 * size_t(n))[r] = make_double4(S.ax[r], S.ay[r], S.az[r], 0.0);
            b[12 * size_t(n) + r] = S.temp[r];
            b[13 * size_t(n) + r] = S.id[r];
            b[14 * size_t(n) + r] = S.rad[r];
            b[15 * size_t(n) + r] = double(S.type[r]);
        }
    }
}

__global__ void pow_third_kernel(const double* x, double* y, int n)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < n) y[k] = pow_third(x[k]);
}

struct World {
    double* id = nullptr;
    int* type = nullptr;
    double *c_lo = nullptr, *c_hi = nullptr, *c_time = nullptr;
    NewBody* fresh = nullptr;
    int* state = nullptr;
    unsigned long long* rng = nullptr;
    long long* counters = nullptr;
    double* rec_host = nullptr;  // nbody_world_frame's record: pinned and device-mapped
    int max_bodies = 300;
    double fracture_threshold = 25.0, min_fragment_mass = 0.2, merge_time = 0.5;  // CONFIG, idx:62-64
    bool bodies_set = false;
};

World* world_of(nbody_ctx* c) { return static_cast<World*>(c->world_state); }

}  // namespace

void world_free(nbody_ctx* c)
{
    World* w = world_of(c);
    if (!w) return;
    cudaFree(w->id);
    cudaFree(w->type);
    cudaFree(w->c_lo);
    cudaFree(w->c_hi);
    cudaFree(w->c_time);
    cudaFree(w->fresh);
    cudaFree(w->state);
    cudaFree(w->rng);
    cudaFree(w->counters);
    if (w->rec_host) cudaFreeHost(w->rec_host);
    delete w;
    c->world_state = nullptr;
}

extern "C" int nbody_world_enable(nbody_ctx* c, unsigned long long seed, size_t max_bodies)
{
    DeviceGuard device_guard_;
    if (!c) return NBODY_EINVAL;
    if (c->prec != NBODY_F64 || c->world != 1) return fail(c, NBODY_EINVAL, "world: the device collision stage needs an fp64 context on one GPU");
    if (max_bodies == 0 || max_bodies > size_t(kWorldMax) || max_bodies > c->capacity)
        return fail(c, NBODY_EINVAL, "world: max_bodies=%zu outside 1..min(%d, capacity=%zu)", max_bodies, kWorldMax, c->capacity);
    Shard& s = c->shards[0];
    CU(c, cudaSetDevice(s.device));
    World* w = world_of(c);
    if (!w) {
        w = new (std::nothrow) World();
        if (!w) return NBODY_ENOMEM;
        c->world_state = w;
        CU(c, cudaMalloc(&w->id, kWorldMax * sizeof(double)));
        CU(c, cudaMalloc(&w->type, kWorldMax * sizeof(int)));
        CU(c, cudaMalloc(&w->c_lo, kWorldMaxContacts * sizeof(double)));
        CU(c, cudaMalloc(&w->c_hi, kWorldMaxContacts * sizeof(double)));
        CU(c, cudaMalloc(&w->c_time, kWorldMaxContacts * sizeof(double)));
        CU(c, cudaMalloc(&w->fresh, kWorldMaxNew * sizeof(NewBody)));
        CU(c, cudaMalloc(&w->state, 4 * sizeof(int)));
        CU(c, cudaMalloc(&w->rng, sizeof(unsigned long long)));
        CU(c, cudaMalloc(&w->counters, 5 * sizeof(long long)));
        CU(c, cudaHostAlloc(&w->rec_host, world_rec_bytes(kWorldMax), cudaHostAllocMapped | cudaHostAllocPortable));
        CU(c, cudaFuncSetAttribute(world_step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(sizeof(Smem))));
    }
    w->max_bodies = int(max_bodies);
    w->bodies_set = false;
    CU(c, cudaMemsetAsync(w->state, 0, 4 * sizeof(int), s.stream));
    CU(c, cudaMemsetAsync(w->counters, 0, 5 * sizeof(long long), s.stream));
    CU(c, cudaMemcpyAsync(w->rng, &seed, sizeof seed, cudaMemcpyHostToDevice, s.stream));
    CU(c, cudaStreamSynchronize(s.stream));
    return NBODY_OK;
}

extern "C" int nbody_world_config(nbody_ctx* c, double fracture_threshold, double min_fragment_mass, double merge_time)
{
    DeviceGuard device_guard_;
    if (!c || !world_of(c)) return c ? fail(c, NBODY_EINVAL, "world: not enabled") : NBODY_EINVAL;
    World* w = world_of(c);
    w->fracture_threshold = fracture_threshold;
    w->min_fragment_mass = min_fragment_mass;
    w->merge_time = merge_time;
    return NBODY_OK;
}

extern "C" int nbody_world_set_bodies(nbody_ctx* c, const double* id, const int* type, const double* radius)
{
    DeviceGuard device_guard_;
    int rc = check_ready(c, "world_set_bodies");
    if (rc != NBODY_OK) return rc;
    World* w = world_of(c);
    if (!w) return fail(c, NBODY_EINVAL, "world: not enabled");
    if (!id || !type || !radius) return fail(c, NBODY_EINVAL, "world_set_bodies: id, type and radius are required");
    if (c->n > size_t(w->max_bodies)) return fail(c, NBODY_EINVAL, "world: %zu bodies uploaded, max_bodies=%d", c->n, w->max_bodies);
    rc = nbody_set_radius(c, radius);
    if (rc != NBODY_OK) return rc;
    Shard& s = c->shards[0];
    CU(c, cudaSetDevice(s.device));
    const int n = int(c->n);
    CU(c, cudaMemcpyAsync(w->id, id, c->n * sizeof(double), cudaMemcpyHostToDevice, s.stream));
    CU(c, cudaMemcpyAsync(w->type, type, c->n * sizeof(int), cudaMemcpyHostToDevice, s.stream));
    CU(c, cudaMemcpyAsync(w->state, &n, sizeof n, cudaMemcpyHostToDevice, s.stream));
    CU(c, cudaMemsetAsync(w->state + 1, 0, sizeof(int), s.stream));  // resetScenario: physics = a new contact map (idx:745-746)
    CU(c, cudaStreamSynchronize(s.stream));
    w->bodies_set = true;
    return NBODY_OK;
}

namespace {

int launch_world(nbody_ctx* c, const char* who, double dt, double G, double softening, int nsteps, int collisions, double* rec)
{
    int rc = check_ready(c, who);
    if (rc != NBODY_OK) return rc;
    World* w = world_of(c);
    if (!w || !w->bodies_set) return fail(c, NBODY_EINVAL, "%s before world_enable / world_set_bodies", who);
    if (nsteps < 0 || softening < 0.0) return fail(c, NBODY_EINVAL, "%s: nsteps=%d softening=%g", who, nsteps, softening);
    if (nsteps == 0 && !rec) return NBODY_OK;
    Shard& s = c->shards[0];
    CU(c, cudaSetDevice(s.device));
    WorldParams p;
    p.pos[0] = static_cast<double4*>(s.pos[c->cur]);
    p.pos[1] = static_cast<double4*>(s.pos[c->cur ^ 1]);
    p.vel = static_cast<double4*>(s.vel);
    p.acc = static_cast<double4*>(s.acc);
    p.temp = static_cast<double*>(s.temp);
    p.radius = static_cast<double*>(s.radius);
    p.id = w->id;
    p.type = w->type;
    p.c_lo = w->c_lo;
    p.c_hi = w->c_hi;
    p.c_time = w->c_time;
    p.fresh = w->fresh;
    p.state = w->state;
    p.rng = w->rng;
    p.counters = w->counters;
    p.cap_rows = int(std::min(c->cap_pad, round_up(size_t(w->max_bodies), pad_unit_of(c))));
    p.max_bodies = w->max_bodies;
    p.nsteps = nsteps;
    p.collisions = collisions != 0;
    p.dt = dt;
    p.half_dt = 0.5 * dt;
    p.G = G;
    p.eps2 = softening * softening;
    p.heat_decay = c->heat_decay;
    p.fracture_threshold = w->fracture_threshold;
    p.min_fragment_mass = w->min_fragment_mass;
    p.merge_time = w->merge_time;
    p.pad_x = 1e150;
    p.rec = rec;
    world_step_kernel<<<1, kWorldThreads, sizeof(Smem), s.stream>>>(p);
    c->launches++;
    CUK(c, "world_step_kernel", s.device);
    c->near_valid = false;
    c->world_dirty = true;
    return NBODY_OK;
}

const char* world_error_text(long long e)
{
    static const char* what[] = {"", "contact map full", "overlap list full", "too many new bodies in one call", "body list outgrew the device arrays"};
    return what[e > 0 && e < 5 ? e : 0];
}

}  // namespace

extern "C" int nbody_world_step(nbody_ctx* c, double dt, double G, double softening, int nsteps, int collisions)
{
    DeviceGuard device_guard_;
    return launch_world(c, "world_step", dt, G, softening, nsteps, collisions, nullptr);
}

int world_adopt_count(nbody_ctx* c, size_t n);  // nbody_capi.cu

// The page's frame in one call (idx:873-883: subSteps x integrate, then visuals.update reads every body): the step kernel
// itself stores what the caller reads into a pinned, device-mapped record, so the frame costs one launch and ONE
// synchronise -- nbody_world_step + nbody_world_sync + nbody_download is one launch, three synchronises and nine copies.
extern "C" int nbody_world_frame(nbody_ctx* c, double dt, double G, double softening, int nsteps, int collisions, size_t cap, size_t* n_out,
                                 double* pos_mass, double* vel, double* acc, double* temp, double* id, int* type, double* radius,
                                 long long counters[5], size_t* n_contacts, unsigned long long* rng_state)
{
    DeviceGuard device_guard_;
    if (!c) return NBODY_EINVAL;
    World* w = world_of(c);
    if (!w) return fail(c, NBODY_EINVAL, "world_frame: world not enabled");
    int rc = launch_world(c, "world_frame", dt, G, softening, nsteps, collisions, w->rec_host);
    if (rc != NBODY_OK) return rc;
    Shard& s = c->shards[0];
    CU(c, cudaStreamSynchronize(s.stream));
    const long long* h = reinterpret_cast<const long long*>(w->rec_host);
    if (h[2]) return fail(c, NBODY_ENOMEM, "world_frame: %s", world_error_text(h[2]));
    rc = world_adopt_count(c, size_t(h[0]));
    if (rc != NBODY_OK) return rc;
    c->world_dirty = false;
    const size_t n = c->n;
    if (n_out) *n_out = n;
    if (n_contacts) *n_contacts = size_t(h[1]);
    if (rng_state) *rng_state = static_cast<unsigned long long>(h[3]);
    if (counters) memcpy(counters, h + 4, 5 * sizeof(long long));
    if ((pos_mass || vel || acc || temp || id || type || radius) && n > cap)
        return fail(c, NBODY_ENOMEM, "world_frame: the device holds %zu bodies, the caller's arrays %zu", n, cap);
    const double* b = w->rec_host + kWorldRecHeader;
    if (pos_mass) memcpy(pos_mass, b, 4 * n * sizeof(double));
    if (vel) memcpy(vel, b + 4 * n, 4 * n * sizeof(double));
    if (acc) memcpy(acc, b + 8 * n, 4 * n * sizeof(double));
    if (temp) memcpy(temp, b + 12 * n, n * sizeof(double));
    if (id) memcpy(id, b + 13 * n, n * sizeof(double));
    if (radius) memcpy(radius, b + 14 * n, n * sizeof(double));
    if (type)
        for (size_t k = 0; k < n; ++k) type[k] = int(b[15 * n + k]);
    return NBODY_OK;
}

// The body count follows the device (merges, fractures): everything nbody_upload derives from n is re-derived
// (world_adopt_count, nbody_capi.cu).

extern "C" int nbody_world_sync(nbody_ctx* c, size_t* n_out, double* id, int* type, double* radius, long long counters[5], size_t* n_contacts,
                                unsigned long long* rng_state)
{
    DeviceGuard device_guard_;
    int rc = check_ready(c, "world_sync");
    if (rc != NBODY_OK) return rc;
    World* w = world_of(c);
    if (!w) return fail(c, NBODY_EINVAL, "world: not enabled");
    Shard& s = c->shards[0];
    CU(c, cudaSetDevice(s.device));
    int st[4] = {0, 0, 0, 0};
    CU(c, cudaMemcpyAsync(st, w->state, sizeof st, cudaMemcpyDeviceToHost, s.stream));
    CU(c, cudaStreamSynchronize(s.stream));
    if (st[2]) return fail(c, NBODY_ENOMEM, "world_step: %s", world_error_text(st[2]));
    if (c->world_dirty) {
        rc = world_adopt_count(c, size_t(st[0]));
        if (rc != NBODY_OK) return rc;
        c->world_dirty = false;
    }
    const size_t n = c->n;
    if (n_out) *n_out = n;
    if (n_contacts) *n_contacts = size_t(st[1]);
    if (id) CU(c, cudaMemcpyAsync(id, w->id, n * sizeof(double), cudaMemcpyDeviceToHost, s.stream));
    if (type) CU(c, cudaMemcpyAsync(type, w->type, n * sizeof(int), cudaMemcpyDeviceToHost, s.stream));
    if (radius) CU(c, cudaMemcpyAsync(radius, s.radius, n * sizeof(double), cudaMemcpyDeviceToHost, s.stream));
    if (counters) CU(c, cudaMemcpyAsync(counters, w->counters, 5 * sizeof(long long), cudaMemcpyDeviceToHost, s.stream));
    if (rng_state) CU(c, cudaMemcpyAsync(rng_state, w->rng, sizeof(unsigned long long), cudaMemcpyDeviceToHost, s.stream));
    CU(c, cudaStreamSynchronize(s.stream));
    return NBODY_OK;
}

extern "C" int nbody_world_contacts(nbody_ctx* c, double* lo_hi_time /* 3 * cap */, size_t cap, size_t* count)
{
    DeviceGuard device_guard_;
    int rc = check_ready(c, "world_contacts");
    if (rc != NBODY_OK) return rc;
    World* w = world_of(c);
    if (!w || !count) return fail(c, NBODY_EINVAL, "world_contacts: world not enabled or count missing");
    Shard& s = c->shards[0];
    CU(c, cudaSetDevice(s.device));
    int st[2] = {0, 0};
    CU(c, cudaMemcpyAsync(st, w->state, sizeof st, cudaMemcpyDeviceToHost, s.stream));
    CU(c, cudaStreamSynchronize(s.stream));
    *count = size_t(st[1]);
    if (*count > cap) return fail(c, NBODY_ENOMEM, "world_contacts: %zu entries, room for %zu", *count, cap);
    if (*count && lo_hi_time) {
        std::vector<double> a(*count), b(*count), t(*count);
        CU(c, cudaMemcpy(a.data(), w->c_lo, *count * sizeof(double), cudaMemcpyDeviceToHost));
        CU(c, cudaMemcpy(b.data(), w->c_hi, *count * sizeof(double), cudaMemcpyDeviceToHost));
        CU(c, cudaMemcpy(t.data(), w->c_time, *count * sizeof(double), cudaMemcpyDeviceToHost));
        for (size_t k = 0; k < *count; ++k) {
            lo_hi_time[3 * k] = a[k];
            lo_hi_time[3 * k + 1] = b[k];
            lo_hi_time[3 * k + 2] = t[k];
        }
    }
    return NBODY_OK;
}

// Math.pow(x, 1/3) as the device evaluates it for new bodies (pow_third above), for the tests: y[k] = x[k]^(1/3), host arrays.
extern "C" int nbody_selftest_pow_third(nbody_ctx* c, const double* x, double* y, size_t n)
{
    DeviceGuard device_guard_;
    if (!c || !x || !y || n == 0 || n > (1u << 24)) return NBODY_EINVAL;
    Shard& s = c->shards[0];
    CU(c, cudaSetDevice(s.device));
    double *dx = nullptr, *dy = nullptr;
    CU(c, cudaMalloc(&dx, n * sizeof(double)));
    CU(c, cudaMalloc(&dy, n * sizeof(double)));
    CU(c, cudaMemcpy(dx, x, n * sizeof(double), cudaMemcpyHostToDevice));
    pow_third_kernel<<<unsigned((n + 255) / 256), 256, 0, s.stream>>>(dx, dy, int(n));
    CUK(c, "pow_third_kernel", s.device);
    CU(c, cudaStreamSynchronize(s.stream));
    CU(c, cudaMemcpy(y, dy, n * sizeof(double), cudaMemcpyDeviceToHost));
    cudaFree(dx);
    cudaFree(dy);
    return NBODY_OK;
}
