/*
 * include/nbody.h -- C ABI of libnbody.so, the B200 (sm_100a) all-pairs
 * gravity + kick/drift engine that stands behind the reference's step contract.
 *
 * The reference (/root/reference/index.html, "idx:NNN" below) has no FFI: its
 * boundary is the in-file JS method PhysicsSystem.integrate(dt) (idx:247-262)
 * which mutates physics.bodies[*].pos/vel/acc/temperature in place, reads G
 * and softening from the global CONFIG at call time (idx:268-269) and is
 * called subSteps times per frame by animate() (idx:877-879).  Each entry
 * point below names the reference statement(s) it replaces.  The bindings a
 * maintainer adds on the reference side (N-API addon + JS shim, or ctypes)
 * are shown in INTEGRATION.md.
 *
 * Conventions
 *   - plain pointers and sizes only; no C++ or torch types cross the ABI;
 *   - every function returns NBODY_OK (0) or a negative nbody_status; no
 *     exception crosses; nbody_last_error() gives the message;
 *   - host arrays are caller-owned, borrowed for the duration of the call;
 *     element type follows the context precision (float or double);
 *       pos_mass[4n] = x,y,z,m      (Body.pos + Body.mass, idx:211,214)
 *       vel[4n]      = vx,vy,vz,-   (Body.vel, idx:216)
 *       acc[4n]      = ax,ay,az,-   (Body.acc, idx:217; state across steps)
 *       temp[n]                     (Body.temperature, idx:223)
 *   - a context is not thread-safe; different contexts are independent;
 *   - step/kick/force calls are asynchronous on the context's stream;
 *     download, energy and sync block until queued work is complete;
 *   - there is no CPU fallback: without a CUDA device every call fails with
 *     NBODY_ECUDA.
 */
#ifndef NBODY_H_
#define NBODY_H_

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define NBODY_API __attribute__((visibility("default")))
#else
#define NBODY_API
#endif

typedef struct nbody_ctx nbody_ctx; /* opaque: device buffers, streams, NCCL communicators */

typedef enum { NBODY_F32 = 0, NBODY_F64 = 1 } nbody_prec;

typedef enum {
    NBODY_OK = 0,
    NBODY_EINVAL = -1, /* bad argument or call order */
    NBODY_ECUDA = -2,  /* CUDA runtime/driver error (includes "no device") */
    NBODY_ENCCL = -3,  /* NCCL error or NCCL library not loadable */
    NBODY_ENOMEM = -4, /* host or device allocation failed (or an output buffer too small) */
    NBODY_EIO = -5     /* snapshot file could not be read/written or is malformed */
} nbody_status;

#define NBODY_NCCL_ID_BYTES 128

/* ---- lifetime ---------------------------------------------------------- */

/* One host process driving n_gpus devices (the N-API host is one process):
 * i-bodies are sharded evenly by contiguous range (the pairs-once force passes share their work-unit table by cost
 * instead and add the devices' exact accumulators), positions are exchanged
 * every step: by default the kernels that move bodies store each
 * new position straight into every peer's buffer over NVLink (peer access / CUDA
 * IPC) and a one-warp epoch barrier follows; NBODY_EXCHANGE=nccl in the
 * environment, or a peer that cannot be mapped, selects ncclAllGather instead.  device_ids may be NULL (devices 0..n_gpus-1).
 * capacity = largest n that upload() will accept.  Replaces `new
 * PhysicsSystem()` (idx:233-237, 734). */
NBODY_API int nbody_create(nbody_ctx** out, nbody_prec prec, int n_gpus, const int* device_ids,
                           size_t capacity);

/* One process per GPU (torchrun-style launch): this process is `rank` of
 * `world` and drives `device`.  nccl_id is the NBODY_NCCL_ID_BYTES blob made by
 * nbody_nccl_unique_id() on rank 0 and broadcast by the caller (any transport);
 * it may be NULL when world == 1. */
NBODY_API int nbody_create_rank(nbody_ctx** out, nbody_prec prec, int rank, int world, int device,
                                const void* nccl_id, size_t capacity);
NBODY_API int nbody_nccl_unique_id(void* id_out /* NBODY_NCCL_ID_BYTES */);

NBODY_API void nbody_destroy(nbody_ctx* ctx);

/* ---- state ------------------------------------------------------------- */

/* Replace the whole body list, as resetScenario / addBody do (idx:745,
 * 239-245).  acc == NULL => acc = 0, like a new Body (idx:217).  temp == NULL
 * => temperatures 0 and never updated.  In rank mode every rank passes the
 * SAME full arrays and the call is a COLLECTIVE (every rank must make it): with peer mappings each rank copies only
 * the slice it owns to its device and packs it into every peer's position buffers, between two cross-rank barriers
 * (a peer a step ahead or behind reads and writes this rank's buffers). */
NBODY_API int nbody_upload(nbody_ctx* ctx, size_t n, const void* pos_mass, const void* vel,
                           const void* acc, const void* temp);

/* Copy state back for the readers after the step (visuals.update, idx:881).
 * Any pointer may be NULL.  In rank mode pos_mass is complete and vel/acc/temp
 * are written only for the rank's own range [nbody_range()]. */
NBODY_API int nbody_download(nbody_ctx* ctx, void* pos_mass, void* vel, void* acc, void* temp);

/* As nbody_download, but pos_mass too is written for the own range only ([i_begin, i_end) of nbody_get_info; every other
 * row of the caller's array is left alone): in rank mode the ranks together read each body back once instead of once per
 * rank.  With one process driving all devices it equals nbody_download. */
NBODY_API int nbody_download_own(nbody_ctx* ctx, void* pos_mass, void* vel, void* acc, void* temp);

/* ---- the hot path ------------------------------------------------------ */

/* nsteps x PhysicsSystem.integrate(dt) with resolveCollisions removed
 * (idx:247-262 minus idx:256), G and softening as read from CONFIG at
 * idx:268-269.  nsteps = CONFIG.subSteps reproduces animate()'s inner loop
 * (idx:877-879).  Temperatures decay by CONFIG.heatDecay = 0.998 per step when
 * a temp array was uploaded (idx:227-230, 260); see nbody_set_heat_decay.
 * Deviations and notes.  (1) fp32, softening == 0, two DISTINCT bodies at one point: the page yields NaN there
 * (idx:280: G / Math.pow(0, 1.5) = Infinity, times 0) and one step later every body is NaN; the fp32 kernels clamp r^2
 * and the pair contributes 0, the run stays finite.  The fp64 mode follows the page.  (2) fp32 systems of 32,768 bodies
 * and more with softening > 0 are stepped by the pairs-once kernel (every pair evaluated once for both bodies, as the
 * page's i<j loop does, exact fixed-point accumulation): a NaN or infinite coordinate turns EVERY acceleration of that
 * step into NaN (the page reaches the same state one step later).  Results do not depend on the number of GPUs. */
NBODY_API int nbody_step(nbody_ctx* ctx, double dt, double G, double softening, int nsteps);

/* One frame of the JS shim in one call (animate(), idx:877-881): upload the caller's arrays
 * (as nbody_upload: a new body list), nsteps x integrate(dt), and write pos_mass, vel, acc (and temp, when not
 * NULL) back in place for visuals.update -- the state ALSO stays on the device for nbody_contacts,
 * nbody_history_push, nbody_render.  Same results as the three calls; for systems of up to 4,096 bodies on one
 * device the arrays travel as one pinned record each way and the call synchronises once. */
NBODY_API int nbody_step_host(nbody_ctx* ctx, size_t n, void* pos_mass, void* vel, void* acc, void* temp,
                              double dt, double G, double softening, int nsteps);

/* Split form, so a host can keep the reference's own collision stage between
 * force evaluation and kick-2 (idx:255-259). */
NBODY_API int nbody_kick1_drift(nbody_ctx* ctx, double dt);             /* idx:248-253 */
NBODY_API int nbody_force(nbody_ctx* ctx, double G, double softening);  /* idx:264-291 */
NBODY_API int nbody_kick2(nbody_ctx* ctx, double dt, double heat_decay /* <=0: no thermal */); /* idx:258-261 */

NBODY_API int nbody_set_heat_decay(nbody_ctx* ctx, double heat_decay); /* CONFIG.heatDecay, idx:67 */

/* ---- collision broad phase ------------------------------ */

/* Body.radius (idx:221) of the n uploaded bodies, element type per the context
 * precision; enables the contact stage.  Every upload() drops the radii (a new
 * body list has new radii).  While radii are set, fp32 force sweeps also track
 * each body's nearest neighbour (one ALU-pipe FMNMX per j-pair), so that
 * nbody_contacts called right after nbody_force / the last step of nbody_step
 * only scans bodies that can be touching something.  NULL switches it off.
 * In rank mode every rank passes the same full array. */
NBODY_API int nbody_set_radius(nbody_ctx* ctx, const void* radius);

/* The O(N^2) overlap search of resolveCollisions (idx:301-313), on the device:
 * every pair i<j with |pos_j - pos_i|^2 < (r_i + r_j + margin)^2, the deciding
 * comparison evaluated in fp64 in the reference's association on the state as
 * stored, written as (i,j) uint32 pairs in the order the reference visits them
 * (ascending i, then ascending j).  margin = 0 is the reference's test; a
 * positive margin lets a host run the order-dependent response (idx:314-389,
 * which moves bodies, idx:351-352) on a superset.  *count = pairs found; if it
 * exceeds cap nothing is written and NBODY_ENOMEM is returned (call again with
 * room for *count).  Synchronous.  In rank mode each rank gets the pairs its own
 * bodies' scans found; the union over ranks is the complete set (a pair with a
 * dominant body can appear on two ranks: sort and de-duplicate after merging). */
NBODY_API int nbody_contacts(nbody_ctx* ctx, double margin, unsigned* pairs /* 2*cap */, size_t cap,
                             size_t* count);

/* What the last nbody_contacts did: out = {bodies flagged by the force sweep's
 * nearest-neighbour tracking, bodies scanned by the pair kernel, pairs found,
 * how: 0 every body was scanned against every other, 1 the flagged bodies against
 * every other, 2 the flagged bodies against each other (under half the bodies flagged)}. */
NBODY_API int nbody_contact_stats(nbody_ctx* ctx, long long out[4]);

/* ---- the page's COMPLETE step on the device (response included) ---- */

#define NBODY_WORLD_MAX_BODIES 1024

/* Turns an fp64 context on one GPU into the page's `physics` object (idx:233-237): besides positions, velocities,
 * accelerations, temperatures and radii the device then keeps Body.id and Body.type of every body, PhysicsSystem.contactPairs
 * and a seeded Math.random stream, and nbody_world_step runs PhysicsSystem.integrate(dt) WITH resolveCollisions, mergeBodies
 * and fractureBodies (idx:247-262, 293-443) without returning to the host: one CTA, the state in shared memory, the
 * collision stage bit-identical to the reference's statements in IEEE binary64 (csrc/page_world.cu).
 * seed: the Math.random stand-in is a sequential splitmix64 stream (the oracles' stream); max_bodies = CONFIG.maxBodies
 * (idx:57), at most NBODY_WORLD_MAX_BODIES and the context's capacity. */
NBODY_API int nbody_world_enable(nbody_ctx* ctx, unsigned long long seed, size_t max_bodies);
/* CONFIG.fractureThreshold, .minFragmentMass, .mergeTime (idx:62-64; defaults 25, 0.2, 0.5). */
NBODY_API int nbody_world_config(nbody_ctx* ctx, double fracture_threshold, double min_fragment_mass, double merge_time);
/* After nbody_upload: Body.id (here the Math.random draw the constructor spent, idx:210: a double in [0,1)), Body.type
 * (0 rock, 1 metal, 2 ice) and Body.radius (idx:221) of the n bodies; empties the contact map like resetScenario (idx:745). */
NBODY_API int nbody_world_set_bodies(nbody_ctx* ctx, const double* id, const int* type, const double* radius);
/* nsteps x PhysicsSystem.integrate(dt), complete (collisions != 0) or with idx:256 removed (collisions == 0).
 * Asynchronous; merges and fractures change the body count ON THE DEVICE: call nbody_world_sync before any other entry point. */
NBODY_API int nbody_world_step(nbody_ctx* ctx, double dt, double G, double softening, int nsteps, int collisions);
/* Synchronises; the context adopts the device's body count (nbody_download etc. then work on the new list).  Any output may
 * be NULL: *n, id[n], type[n], radius[n], counters = {merges, fractures, overlapping pairs acted on, visuals.triggerFlash
 * calls, visuals.spawnExplosion calls} since nbody_world_enable, the size of the contact map, the random stream's state. */
NBODY_API int nbody_world_sync(nbody_ctx* ctx, size_t* n, double* id, int* type, double* radius, long long counters[5],
                               size_t* n_contacts, unsigned long long* rng_state);
/* The page's frame in one call (idx:873-883: subSteps x integrate(dt), then visuals.update reads every body): nbody_world_step
 * (nsteps >= 0) followed by everything nbody_world_sync and nbody_download would return, through ONE kernel launch and ONE
 * synchronise -- the step kernel stores the frame's record straight into pinned host memory.  `cap` = the rows the caller's
 * arrays hold (pos_mass, vel, acc: 4 doubles per row); any output may be NULL; NBODY_ENOMEM when the device's body list has
 * outgrown cap (the context has adopted the new count all the same: call again with larger arrays and nsteps = 0). */
NBODY_API int nbody_world_frame(nbody_ctx* ctx, double dt, double G, double softening, int nsteps, int collisions, size_t cap,
                                size_t* n, double* pos_mass, double* vel, double* acc, double* temp, double* id, int* type,
                                double* radius, long long counters[5], size_t* n_contacts, unsigned long long* rng_state);
/* PhysicsSystem.contactPairs as (lower id, higher id, accumulated contact time) triples, in no particular order. */
NBODY_API int nbody_world_contacts(nbody_ctx* ctx, double* lo_hi_time /* 3*cap */, size_t cap, size_t* count);

/* Test hook: y[k] = Math.pow(x[k], 1/3) as nbody_world_step evaluates Body.radius of merged bodies and fragments (idx:221):
 * a double-double evaluation, correctly rounded (the device's pow() is a 2-ulp routine).  Host arrays. */
NBODY_API int nbody_selftest_pow_third(nbody_ctx* ctx, const double* x, double* y, size_t n);

/* Test hook, host logic only (works without a device): the work-unit table of the fp32 pairs-once pass for n bodies
 * (bodies_per_thread 4 or 8, `run` position tiles of 512 bodies per unit) and the contiguous shares of `world` ranks,
 * balanced by tile visits.  units_out receives (i-tile, first tile, tile count, diagonal flag) per unit. */
NBODY_API int nbody_selftest_sym32_units(size_t n, int bodies_per_thread, int run, int world, unsigned* units_out /* 4*cap */,
                                         size_t cap, size_t* count, unsigned* split_out /* world + 1 */);
/* ... and of its warp-unit form (csrc/force_f32_symw.cuh; bodies_per_thread 2, 4 or 8): i-blocks of 32 x bodies_per_thread
 * bodies against runs of `run` groups of 64 bodies; (i-block, first group, group count, diagonal flag). */
NBODY_API int nbody_selftest_symw_units(size_t n, int bodies_per_thread, int run, int world, unsigned* units_out /* 4*cap */,
                                        size_t cap, size_t* count, unsigned* split_out /* world + 1 */);

/* ---- an entry point that allocates nothing ---- */

#define NBODY_STATIC_MAX_BODIES 16384

/* No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Needs no context.
 * acc[i] = G * sum_{j != i} m_j (p_j - p_i) / (|p_j - p_i|^2 + softening^2)^1.5 for n <= NBODY_STATIC_MAX_BODIES bodies in fp32, by the
 * pairs-once force pass with exact fixed-point accumulators (csrc/force_f32_symw.cuh): what nbody_force + nbody_download(acc)
 * return, within the fp32 tolerance of the tests (2e-5 of the largest acceleration).  softening > 0.
 * pos_mass_dev: 4n floats (x, y, z, m) and acc_dev: 4n floats (ax, ay, az, 0), both DEVICE pointers on the CURRENT device;
 * the four kernels are queued on `stream` (a cudaStream_t, NULL = the default stream) and the call returns without
 * synchronising.  It calls no cudaMalloc / cudaFree and no allocating library routine: all scratch memory (0.76 MiB: packed
 * positions, accumulators, unit table) is ONE static device array that every GPU which loads the library holds until the
 * process exits.  Every call on a device shares that array: calls that overlap on the GPU (two streams, two threads) race;
 * keep them ordered.  n > NBODY_STATIC_MAX_BODIES, NULL pointers or softening <= 0: NBODY_EINVAL; the reason is written to
 * err (err_cap bytes, may be NULL).  A non-finite coordinate turns every acceleration into NaN, as in nbody_step. */
NBODY_API int nbody_accel_f32_static(const float* pos_mass_dev, float* acc_dev, size_t n, double G, double softening, void* stream,
                                     char* err, size_t err_cap);

/* ---- renderer feed: position history and ribbon trails ---- */

/* Keep Body.history on the device: the last `depth` positions of every body,
 * newest first (CONFIG.trailLength = 80, idx:58, 540, 564-565).  depth = 0 frees
 * the ring; 2 <= depth <= 256 otherwise.  The ring is (re)filled with `depth`
 * copies of the current position at the first push after it was enabled or after
 * an upload that changed n, as a Body seen for the first time is (idx:540); an
 * upload with the same n keeps it.  Kept on device 0 (every device holds all
 * positions). */
NBODY_API int nbody_history_enable(nbody_ctx* ctx, int depth);

/* One visuals.update() worth of history: unshift the current position, drop the
 * oldest (idx:564-565).  Asynchronous on the context's stream. */
NBODY_API int nbody_history_push(nbody_ctx* ctx);

/* The body list changed (merge, fracture, spawn: idx:383-389, 835) and was uploaded again, but a
 * Body object carries its history with it: body i of the current list continues the ring of
 * body source[i] of the list the ring was last pushed for; source[i] < 0 marks a body the
 * renderer has not seen, whose ring the next push fills with its position (idx:540).  Without
 * this call an upload that changes n starts every ring over.  Reads of the history are refused
 * between a remap and the next push. */
NBODY_API int nbody_history_remap(nbody_ctx* ctx, const int* source);

/* out[n][depth][3] (context precision) = Body.history of every body, newest first. */
NBODY_API int nbody_history_download(nbody_ctx* ctx, void* out);

/* The ribbon loop of VisualSystem.update (idx:567-592) for every body: out is
 * float[n][depth][6], the `positions` Float32Array of each trail mesh (idx:524,
 * 582-588): vertex pair i = history[i] +/- perp, perp = normalize(cross(camera -
 * history[i], dir_i)) * radius * 0.8 * max(0, 1 - i/(depth-1)).  Needs
 * nbody_set_radius.  fp64 contexts reproduce the JS arithmetic and its float
 * rounding exactly. */
NBODY_API int nbody_trail_ribbons(nbody_ctx* ctx, const double camera[3], float* out);

/* ---- headless body renderer ------------------------------- */

/* Per-body material inputs of the planet shader: Body.type (0 rock, 1 metal, 2 ice:
 * CONFIG.materials colours, idx:70-74) and the `seed` uniform each mesh material draws
 * at creation (Math.random()*100, idx:492).  Either may be NULL (rock / 0).  Dropped by
 * every upload(), like the radii. */
NBODY_API int nbody_set_materials(nbody_ctx* ctx, const int* type, const float* seed);

/* Render every body as the page's planet shaders would shade it (planetVertexShader +
 * planetFragmentShader, idx:83-202), with no graphics API: each sphere (centre pos,
 * radius from nbody_set_radius, idx:545-546) is ray-cast from a perspective camera
 * (THREE.PerspectiveCamera(fov_deg, width/height, 0.1, 3000) at cam_position looking at
 * cam_target, +y up; the page uses fov 45, position (0,80,150), target 0, idx:704-705),
 * the nearest body per pixel wins (lower index on a tie), and the fragment shader runs
 * on the hit point with the body's temperature (the uploaded temp array), type and seed.
 * rgba: float[height][width][4] = gl_FragColor BEFORE three.js's tone mapping and bloom
 * (library passes, not restated); background 0,0,0,0; row 0 is the top of the image.
 * ids: int[height][width] = body index or -1; may be NULL.  rgba may be NULL too: the
 * frame is then rendered and left on the device.  All arithmetic is fp32, one rounded
 * operation per GLSL operation (no FMA contraction), in every context precision. */
NBODY_API int nbody_render(nbody_ctx* ctx, int width, int height, const double cam_position[3],
                           const double cam_target[3], double fov_deg, float* rgba, int* ids);

/* ---- snapshots (the reference saves nothing) ---- */

/* Raw little-endian dump of the state.  64-byte header:
 *   0  char     magic[8] = "NBODYSN1"
 *   8  uint64   n
 *   16 uint64   step            (caller's step counter)
 *   24 float64  dt, 32 float64 G, 40 float64 softening
 *   48 uint32   precision       (0 = float32, 1 = float64: element type of the arrays)
 *   52 uint32   flags           (bit 0: temp present, bit 1: radius present)
 *   56 uint64   reserved (0)
 * then pos_mass[4n], vel[4n], acc[4n], [temp[n]], [radius[n]] in the host layout
 * above.  save synchronises and downloads; load uploads (n <= capacity, same
 * precision) and returns the header fields through the non-NULL pointers. */
NBODY_API int nbody_snapshot_save(nbody_ctx* ctx, const char* path, unsigned long long step, double dt,
                                  double G, double softening);
NBODY_API int nbody_snapshot_load(nbody_ctx* ctx, const char* path, unsigned long long* step, double* dt,
                                  double* G, double* softening);

/* ---- diagnostics (not in the reference) -------------------- */

/* E_kin = sum 1/2 m v^2, E_pot = -G sum_{i<j} m_i m_j / sqrt(r^2 + eps^2),
 * momentum = sum m v; evaluated in fp64 on the device whatever the state
 * precision.  In rank mode each rank returns the same totals. */
NBODY_API int nbody_energy(nbody_ctx* ctx, double G, double softening, double* kinetic,
                           double* potential, double momentum[3]);

NBODY_API int nbody_sync(nbody_ctx* ctx);
NBODY_API const char* nbody_last_error(const nbody_ctx* ctx); /* valid until next call on ctx */
NBODY_API const char* nbody_strerror(int status);
NBODY_API const char* nbody_version(void);
NBODY_API int nbody_device_count(void); /* >= 0, or negative nbody_status */

/* ---- plumbing for hosts that own streams / want timings ------------------ */

/* Run device 0's (or the rank's) work on a caller-owned CUDA stream
 * (cudaStream_t / CUstream passed as void*, e.g. torch's current stream);
 * NULL restores the context's own stream. */
NBODY_API int nbody_set_stream(nbody_ctx* ctx, void* cuda_stream);

/* Bracket every force-kernel launch with CUDA events on its own stream.
 * nbody_read_timing synchronises, returns launches and summed milliseconds on
 * device 0 (or the rank's device) since the last read, and resets both. */
NBODY_API int nbody_enable_timing(nbody_ctx* ctx, int enable);
NBODY_API int nbody_read_timing(nbody_ctx* ctx, long long* launches, double* force_ms);

typedef struct {
    size_t n;            /* bodies uploaded */
    size_t n_padded;     /* rounded up to the i-tile; pads have mass 0 */
    size_t i_begin;      /* first body this process/device 0 owns */
    size_t i_end;        /* one past the last */
    int n_gpus;          /* devices in this context (1 in rank mode) */
    int world;           /* ranks (1 unless nbody_create_rank) */
    int rank;
    int sm_count;        /* of device 0 / the rank's device */
    int j_chunks;        /* j-range split of the force kernel for this n */
    int i_tile;          /* bodies per CTA */
    long long kernel_launches; /* kernels launched by this context so far */
    int exchange;        /* position exchange: 0 none (1 GPU), 1 ncclAllGather, 2 fused peer stores over NVLink */
    int uniform_mass;    /* 0: general fp32 force kernel.  1: every uploaded mass is equal and the kernel skips the per-pair
                            mass multiply.  1+k: all but k (<= 64) bodies share one mass; same kernel, k corrections per body */
    int f64_symmetric;   /* fp64, one GPU: > 0 when the force pass evaluates every pair once and applies it to both bodies,
                            as the reference's i<j loop does (idx:272-288); the value is the j-blocks per CTA.  0: one-sided */
    int f32_symmetric;   /* fp32: > 0 when force passes (softening > 0, no radii set) evaluate every pair once and apply it to both
                            bodies through exact fixed-point accumulators (force_f32_sym.cuh); the value is the position tiles
                            (512 bodies) per work unit.  0: the one-sided sweep */
    int f32_sym_bodies_per_thread; /* i-bodies per thread of that pass (4 or 8), 0 when it is off */
    int f32_sym_last_pass;         /* 1 when the last force pass actually was the pairs-once one (softening == 0 and set radii take the sweep) */
    long long upload_h2d_bytes;    /* host->device bytes THIS process copied in the last nbody_upload (with peer mappings every device
                                      copies only its own slice of pos_mass and packs it into every peer's buffers) */
    int f32_sym_warp_units;        /* 1 when that pass is the warp-unit form (force_f32_symw.cuh): every warp its own worker, units of
                                      (32 x bodies_per_thread i-bodies) x (f32_symmetric groups of 64 bodies) */
} nbody_info;
NBODY_API int nbody_get_info(nbody_ctx* ctx, nbody_info* out);

#ifdef __cplusplus
}
#endif
#endif /* NBODY_H_ */
