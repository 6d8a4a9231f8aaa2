// Host-side state of libnbody.so shared by its translation units (nbody_capi.cu: lifetime, state,
// the hot path; capi_contacts.cu: collision broad phase; capi_visual.cu: renderer feed, snapshots,
// renderer glue).  One nbody_ctx owns one Shard per device it drives.
#pragma once
#include "../../include/nbody.h"

#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <type_traits>
#include <vector>

#include "common.cuh"
#include "nccl_dyn.h"
#include "render.cuh"

using namespace nb;

struct Shard {
    int device = 0;
    int rank = 0;
    cudaStream_t own_stream = nullptr;
    cudaStream_t stream = nullptr;
    void* arena = nullptr;              // the one allocation peers map: pos[0] | pos[1] | accfix | flags, sym_ctl
    size_t arena_bytes = 0, off_pos1 = 0, off_acc = 0, off_flags = 0;
    void* pos[2] = {nullptr, nullptr};  // full padded pos_mass, double-buffered
    void* vel = nullptr;                // local range
    void* acc = nullptr;
    void* temp = nullptr;
    void* partial = nullptr;
    size_t partial_bytes = 0;
    unsigned* counters = nullptr;
    unsigned* medium_arrivals = nullptr; // grid-barrier counter of medium_step_kernel (medium.cuh)
    double* sym_part_i = nullptr;       // fp64 pairs-once pass (force_f64_sym.cuh): per-chunk i-side sums
    double* sym_part_j = nullptr;       // ... and per-i-block j-side sums
    size_t sym_bytes_i = 0, sym_bytes_j = 0;
    void* stage = nullptr;              // AoS staging for upload/download (4*cap_pad elements)
    void* io_dev = nullptr;             // small systems on one device: the whole state as one record, [pos_mass 4n | vel 4n | acc 4n | temp n]
    void* io_host = nullptr;            // ... its pinned, device-mapped host twin: one copy each way per call, none at all for the smallest systems
    cudaEvent_t io_done = nullptr;      // the last device-side read of io_host (an upload returns before it has run)
    double* energy_out = nullptr;
    size_t energy_rows = 0;
    // collision broad phase (contacts.cuh)
    void* radius = nullptr;             // full padded array, Real, replicated like the masses
    float* near = nullptr;              // local: nearest-neighbour rsqrt from the NEAR force kernel
    unsigned* cand = nullptr;           // local: candidate body indices
    unsigned* cand_count = nullptr;
    unsigned* cand_all = nullptr;       // several devices: the candidates of all of them, concatenated in shard order
    float4* cand_rec = nullptr;         // dense (x, y, z, r) records of those candidates (contacts.cuh, candidate x candidate)
    size_t cand_rec_cap = 0;
    unsigned* big = nullptr;            // the big bodies of this shard's own range (global indices)
    unsigned n_big = 0;
    uint2* pairs = nullptr;
    size_t pair_cap = 0;
    unsigned long long* pair_count = nullptr;
    // position history ring (trails.cuh), first shard only: every shard holds all positions
    void* hist = nullptr;               // Real4 [depth][hist_stride]
    void* hist_alt = nullptr;           // second ring, the target of nbody_history_remap (the two swap)
    int* hist_src = nullptr;            // source index per body of the last remap (< 0: new to the ring)
    float* ribbons = nullptr;           // float [n][depth][6] staging for nbody_trail_ribbons
    size_t ribbons_bytes = 0;
    // headless renderer (render.cu), first shard only
    int* mat_type = nullptr;            // Body.type per body (0 rock, 1 metal, 2 ice)
    float* mat_seed = nullptr;          // the `seed` uniform of each body's material (idx:492)
    // fp32 pairs-once pass (force_f32_sym.cuh)
    unsigned long long* accfix = nullptr;  // [3][cap_pad] fixed-point acceleration sums of ALL bodies, slot order
    unsigned* sym_ctl = nullptr;           // the unit counter of the persistent CTAs; cleared before every launch
    void* sym_units = nullptr;             // the unit table (SymUnit), the same on every device
    size_t sym_units_cap = 0;
    double* mass_sums = nullptr;           // per-block sums of |m| from pack_pos_kernel (the bound behind the fixed-point scale)
    void* peer_acc[kMaxPeers] = {};        // every rank's accfix / sym_ctl as mapped on this device
    void* peer_ctl[kMaxPeers] = {};
    MassOutlier* outliers = nullptr;    // fp32: bodies whose mass differs from the common one (force_f32.cuh)
    unsigned* outlier_count = nullptr;  // how many pack_pos_kernel found
    void* temp_full = nullptr;          // all temperatures gathered on this device (multi-GPU contexts)
    RenderScratch render_scratch;
    float4* image = nullptr;
    int* image_ids = nullptr;
    size_t image_pixels = 0;
    ncclComm_t comm = nullptr;
    // peer-store exchange (exchange.cuh): every rank's two position buffers and flag array as mapped on this device
    void* peer_pos[kMaxPeers][2] = {};
    PeerFlags peer_flags = {};
    unsigned long long* flags = nullptr; // this rank's flag array (kMaxPeers entries)
    unsigned* xerr = nullptr;            // exchange timeout flag
    std::vector<void*> ipc_opened;       // IPC mappings to close (rank mode)
    size_t i_begin = 0;                 // set by upload
    size_t n_local = 0;                 // padded
    size_t n_local_real = 0;            // bodies of [i_begin, i_begin+n_local) that are < n
    // force-kernel timing (shard 0 only)
    std::vector<cudaEvent_t> ev;
    size_t ev_used = 0;
};


struct nbody_ctx {
    nbody_prec prec = NBODY_F32;
    size_t esz = 4;
    size_t capacity = 0;
    size_t cap_pad = 0;
    size_t n = 0, n_pad = 0;   // bodies uploaded; padded for even i-sharding (pads have mass 0)
    size_t n_j = 0;            // j extent = n rounded up to kJUnit: independent of the GPU count
    int cur = 0;
    int world = 1;       // total ranks (devices across all processes)
    bool rank_mode = false;
    int my_rank = 0;     // rank of shards[0] in rank mode
    std::vector<Shard> shards;
    bool has_temp = false;
    bool uniform = false;           // fp32: every uploaded mass is the same value -> the 11-op force kernel
    double uniform_mass = 0.0;
    int f64_math = 2;               // NBODY_F64_MATH at create
    bool f64_sym_allowed = false;   // NBODY_F64_SYM at create (one GPU, MATH 2): pairs evaluated once (force_f64_sym.cuh)
    int f64_sym_r = 4;              // i-bodies per thread of that pass (fewer for small systems)
    int f64_sym_k = 0;              // j-blocks per CTA of that pass for the current upload; 0: the one-sided kernel
    int f64_R_forced = 0;          // NBODY_F64_R at create: 1 or 2 i-bodies per thread in the fp64 kernel, 0 = by body count
    // fp32 pairs-once pass (force_f32_sym.cuh)
    bool sym32_allowed = false;     // NBODY_F32_SYM at create (default on): fp32, softening > 0, no nearest-neighbour tracking
    int sym32_R = 4;                // i-bodies per thread for the current upload: 8 with equal masses, 4 with general ones
    int sym32_R_forced = 0;         // NBODY_F32_SYM_R at create: 4 or 8 whatever the masses
    size_t sym32_min_n = 0;         // smallest system that takes it (NBODY_F32_SYM_MIN)
    int sym32_run = 0;              // position tiles per unit for the current upload; 0: no table (one-sided sweep)
    size_t sym32_table_n = 0;       // the body count / R / run the device table was built for
    int sym32_table_R = 0, sym32_table_run = 0;
    std::vector<unsigned> sym32_split;  // unit ranges of the ranks: rank r takes [split[r], split[r+1])
    double sum_abs_mass = 0.0;      // of the current upload (the equal-mass sweep counts every body as 1)
    int sym32_slots[2][2] = {};    // resident CTAs per device of the general / equal-mass kernel (grid size of the persistent launch)
    bool sym32_used = false;        // the last force pass was the pairs-once one
    bool sym32_fine = false;        // the current table is the warp-unit one (force_f32_symw.cuh): units of 64-body groups
    bool sym32_table_fine = false;
    size_t symw_min_n = 0, symw_max_n = 0;  // body counts that always take the warp-unit form (NBODY_F32_SYMW_MIN / _MAX)
    int symw_R_forced = 0;          // NBODY_F32_SYMW_R at create: 2, 4 or 8
    bool symw_general_large = true; // NBODY_F32_SYMW_GENERAL at create: general masses keep the warp units above symw_max_n
    bool sym_coop = false;          // cooperative launches work here and NBODY_SYM_COOP is not 0: sym_coop.cuh runs mid-size nbody_step calls
    size_t sym_coop_max_n = 0;      // ... up to this many bodies (NBODY_SYM_COOP_MAX)
    int sym_coop_slots[2][4] = {};  // resident CTAs of sym_coop_step_kernel [equal masses][R = 2, 4, -, 8]
    int symw_slots[2][4] = {};      // resident CTAs per device of the warp-unit kernels [equal masses][R = 2, 4, 6, 8]
    size_t medium_max_n = 0;        // largest system that kernel takes (fp32 5,120, fp64 4,096; NBODY_MEDIUM_MAX lowers it)
    int medium_ctas = 0;            // > 0: systems of up to medium_max_n bodies run all steps of a call in one cooperative launch
                                    // of this many CTAs (medium.cuh; NBODY_MEDIUM at create, one GPU, needs cooperative launch)
    size_t small_max_n = 512;       // largest system the one-cluster kernel takes (NBODY_SMALL_MAX at create, <= kSmallMaxN)
    bool small_path = true;         // NBODY_SMALL at create: one-cluster multi-step kernel for n <= kSmallMaxN (small.cuh)
    bool detect_masses = false;     // upload in flight: pack_pos_kernel compares every mass with uniform_mass
    void* step_rec_out = nullptr;   // nbody_step_host -> launch_small_steps: where the cluster kernel also writes the final state
    bool upload_sharded = false;    // the last upload packed every shard's own slice into every rank's buffers (several devices, peer mappings)
    long long upload_h2d_bytes = 0; // host->device bytes this process copied in the last upload
    bool upload_deferred = false;   // the last upload took the small-system path and returned without a synchronise
    bool detect_pending = false;    // ... so nobody has read the detection count yet: resolve_mass_detection() does
                                    // before the first consumer of `uniform` (a force launch of the large kernels, nbody_get_info)
    std::vector<MassOutlier> outliers;  // ... except these few (at most kMaxOutliers), corrected per i-body
    bool has_radius = false;        // nbody_set_radius was called for the current body list
    bool near_valid = false;        // Shard::near describes the CURRENT positions (last launch was a NEAR force pass)
    double near_eps2 = 0.0;         // softening^2 of that pass
    double r_max = 0.0;             // largest radius
    double r_split = 0.0;           // bodies with a larger radius are "big": scanned against everybody (contacts.cuh)
    double r_small_max = 0.0;       // largest radius that is not big
    long long cstat[4] = {0, 0, 0, 0};  // last nbody_contacts: flagged, scanned, pairs, 0 full scan / 1 candidates x all / 2 candidates x candidates
    int hist_depth = 0;             // 0: no ring
    size_t hist_prev_n = 0;         // an upload changed n: the count the ring still describes, until the next push or remap
    bool hist_fresh_pending = false;  // a remap left rows of new bodies for the next push to fill
    size_t hist_stride = 0;         // bodies per slot (= cap_pad)
    size_t hist_n = 0;              // body count the ring was filled for (0: refill on the next push)
    int hist_head = 0;              // slot of the newest sample
    bool has_materials = false;     // nbody_set_materials was called for the current body list
    bool p2p = false;               // peer-store exchange instead of ncclAllGather
    unsigned long long epoch = 0;   // exchange barrier epoch (same sequence on every rank)
    double heat_decay = 0.998;  // CONFIG.heatDecay, /root/reference/index.html:67
    int R = 4;           // i-bodies per thread (fp32)
    int i_tile = 512;
    int j_unit = kJUnit;  // bodies; granularity of the j extent and of the j-chunks
    int j_chunks = 1;
    size_t chunk_len = 0;
    int sm_count = 0;
    long long launches = 0;
    bool timing = false;
    long long t_launches = 0;
    double t_ms = 0.0;
    void* world_state = nullptr;    // the page's complete step on the device (page_world.cu): ids, types, contact map, random stream
    bool world_dirty = false;       // a world step ran: the body count on the device may differ from n until nbody_world_sync
    std::string err;
};

namespace {

// Every entry point leaves the calling thread's current device as it found it (a context drives several devices; a host
// such as PyTorch creates its streams on whatever cudaGetDevice answers).
struct DeviceGuard {
    int dev = -1;
    DeviceGuard() { if (cudaGetDevice(&dev) != cudaSuccess) { dev = -1; cudaGetLastError(); } }
    ~DeviceGuard() { if (dev >= 0) cudaSetDevice(dev); }
};

int fail(nbody_ctx* c, int code, const char* fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (c) c->err = buf;
    return code;
}

#define CU(ctx, call)                                                                         \
    do {                                                                                      \
        cudaError_t e_ = (call);                                                              \
        if (e_ != cudaSuccess)                                                                \
            return fail(ctx, e_ == cudaErrorMemoryAllocation ? NBODY_ENOMEM : NBODY_ECUDA,    \
                        "%s: %s", #call, cudaGetErrorString(e_));                             \
    } while (0)

// after a kernel launch: name the kernel and the device in the message
#define CUK(ctx, what, dev)                                                                   \
    do {                                                                                      \
        cudaError_t e_ = cudaGetLastError();                                                  \
        if (e_ != cudaSuccess) return fail(ctx, NBODY_ECUDA, "%s on device %d: %s", what, int(dev), cudaGetErrorString(e_)); \
    } while (0)

#define NC(ctx, call)                                                                         \
    do {                                                                                      \
        ncclResult_t r_ = (call);                                                             \
        if (r_ != 0)                                                                          \
            return fail(ctx, NBODY_ENCCL, "%s: %s", #call,                                    \
                        nccl().GetErrorString ? nccl().GetErrorString(r_) : "nccl error");    \
    } while (0)

size_t round_up(size_t a, size_t b) { return (a + b - 1) / b * b; }

int env_int(const char* name, int dflt)
{
    const char* s = getenv(name);
    return (s && *s) ? atoi(s) : dflt;
}

size_t gcd_sz(size_t a, size_t b) { return b ? gcd_sz(b, a % b) : a; }

// n is padded to whole i-tiles per shard AND to whole j units: lcm(i_tile * world, j_unit)
// (fp32: the i-tile of the pairs-once pass when that is the larger one)
inline size_t pad_unit_of(const nbody_ctx* c)
{
    const size_t tile = c->prec == NBODY_F32 && c->sym32_allowed ? size_t(std::max(c->i_tile, kF32Threads * 8)) : size_t(c->i_tile);
    const size_t a = tile * size_t(c->world);
    return a / gcd_sz(a, size_t(c->j_unit)) * size_t(c->j_unit);
}

inline int check_ready(nbody_ctx* c, const char* what)
{
    if (!c) return NBODY_EINVAL;
    cudaGetLastError();  // a non-sticky error another library of the process left behind is not this call's (a fault of ours is sticky and stays)
    if (c->n == 0) return fail(c, NBODY_EINVAL, "%s before upload", what);
    return NBODY_OK;
}

}  // namespace
