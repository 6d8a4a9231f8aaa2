// libnbody.so -- implementation of include/nbody.h.
// One nbody_ctx owns one Shard per device it drives.  In "rank mode"
// (nbody_create_rank) the process holds a single Shard that is rank r of a
// world of processes; otherwise it holds n_gpus Shards (ranks 0..n_gpus-1 of a
// world inside this process).  Both run the same per-step schedule; only the
// NCCL communicator construction differs.
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
#include <vector>

#include "common.cuh"
#include "contacts.cuh"
#include "exchange.cuh"
#include "force_f32.cuh"
#include "force_f64.cuh"
#include "integrate.cuh"
#include "nccl_dyn.h"
#include "render.cuh"
#include "small.cuh"
#include "trails.cuh"

using namespace nb;

#ifndef NB_UNIFORM_UNROLL
#define NB_UNIFORM_UNROLL 16
#endif
constexpr int kUniUnroll = NB_UNIFORM_UNROLL;  // inner-loop unroll of the equal-mass force kernel

namespace {

struct Shard {
    int device = 0;
    int rank = 0;
    cudaStream_t own_stream = nullptr;
    cudaStream_t stream = nullptr;
    void* pos[2] = {nullptr, nullptr};  // full padded pos_mass, double-buffered
    void* vel = nullptr;                // local range
    void* acc = nullptr;
    void* temp = nullptr;
    void* partial = nullptr;
    size_t partial_bytes = 0;
    unsigned* counters = nullptr;
    void* stage = nullptr;              // AoS staging for upload/download (4*cap_pad elements)
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

}  // namespace

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
    int f64_R_forced = 0;          // NBODY_F64_R at create: 1 or 2 i-bodies per thread in the fp64 kernel, 0 = by body count
    bool small_path = true;         // NBODY_SMALL at create: one-cluster multi-step kernel for n <= kSmallMaxN (small.cuh)
    bool detect_masses = false;     // upload in flight: pack_pos_kernel compares every mass with uniform_mass
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
    std::string err;
};

namespace {

int check_exchange(nbody_ctx* c);  // defined with the exchange code below

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

// fp64 force kernel: two i-bodies per thread pay from ~32k bodies up (N=65,536: 961 -> 974 G interactions/s with
// equal masses, 906 -> 948 without) and cost below (N=16,384: 304 -> 326 us/step): profiles/r1e_f64_r_uniform.jsonl
constexpr size_t kF64TwoBodyMinN = 32768;
int f64_bodies_per_thread(const nbody_ctx* c, size_t n)
{
    if (c->f64_math != 2) return 1;
    if (c->f64_R_forced == 1 || c->f64_R_forced == 2) return c->f64_R_forced;
    return n >= kF64TwoBodyMinN ? 2 : 1;
}

// n is padded to whole i-tiles per shard AND to whole j units: lcm(i_tile * world, j_unit)
size_t pad_unit(const nbody_ctx* c)
{
    const size_t a = size_t(c->i_tile) * size_t(c->world);
    return a / gcd_sz(a, size_t(c->j_unit)) * size_t(c->j_unit);
}

// j-range decomposition: a pure function of n (through n_j) and the i-tile, never of
// the number of GPUs, so partial sums are added in the same order on 1/2/4/8 GPUs.
// Enough (i-tile, j-chunk) units that the tail of the last wave stays ~1-3% of the launch
// even when 8 GPUs share the i-tiles: 148 SMs x 3 resident CTAs x 256 (measured: with x48
// the N=1M kernel lost 4% at 8 GPUs to a 6.3-wave grid).
void choose_chunks(nbody_ctx* c)
{
    const size_t tiles = (c->n_j + c->i_tile - 1) / c->i_tile;
    // fp64: an interaction costs ~8x the fp32 one, so far fewer units already balance the SMs, and each
    // unit amortises its ring start-up and its share of the partial-sum pass over more j
    // (N=16,384: 256 chunks of 64 bodies 374 us/step, 86 chunks of 192 bodies 332 us; profiles/r1e_f64_sweep.jsonl)
    const size_t want_units = size_t(env_int("NBODY_WANT_UNITS", c->prec == NBODY_F64 ? 148 * 3 * 16 : 148 * 3 * 256));
    const size_t max_chunks = c->n_j / c->j_unit;
    size_t len;
    const int forced = env_int("NBODY_CHUNK_LEN", 0);
    if (forced > 0) {
        len = round_up(size_t(forced), size_t(c->j_unit));
    } else {
        size_t s = (want_units + tiles - 1) / tiles;
        if (s < 1) s = 1;
        if (s > max_chunks) s = max_chunks;
        len = round_up((c->n_j + s - 1) / s, size_t(c->j_unit));
    }
    if (len > c->n_j) len = c->n_j;
    c->chunk_len = len;
    c->j_chunks = int((c->n_j + len - 1) / len);
}

int alloc_shard(nbody_ctx* c, Shard& s)
{
    CU(c, cudaSetDevice(s.device));
    CU(c, cudaStreamCreateWithFlags(&s.own_stream, cudaStreamNonBlocking));
    s.stream = s.own_stream;
    const size_t local_cap = c->cap_pad / c->world;
    const size_t v4 = 4 * c->esz;
    CU(c, cudaMalloc(&s.pos[0], c->cap_pad * v4));
    CU(c, cudaMalloc(&s.pos[1], c->cap_pad * v4));
    CU(c, cudaMalloc(&s.stage, c->cap_pad * v4));
    CU(c, cudaMalloc(&s.vel, local_cap * v4));
    CU(c, cudaMalloc(&s.acc, local_cap * v4));
    CU(c, cudaMalloc(&s.temp, local_cap * c->esz));
    const size_t ntiles = local_cap / c->i_tile + 1;
    CU(c, cudaMalloc(&s.counters, ntiles * sizeof(unsigned)));
    CU(c, cudaMemset(s.counters, 0, ntiles * sizeof(unsigned)));
    CU(c, cudaMalloc(&s.flags, kMaxPeers * sizeof(unsigned long long)));
    CU(c, cudaMemset(s.flags, 0, kMaxPeers * sizeof(unsigned long long)));
    CU(c, cudaMalloc(&s.xerr, sizeof(unsigned)));
    CU(c, cudaMemset(s.xerr, 0, sizeof(unsigned)));
    s.energy_rows = (local_cap + kEnergyThreads - 1) / kEnergyThreads + 1;
    CU(c, cudaMalloc(&s.energy_out, s.energy_rows * 5 * sizeof(double)));
    CU(c, cudaMalloc(&s.radius, c->cap_pad * c->esz));
    CU(c, cudaMalloc(&s.outliers, kMaxOutliers * sizeof(MassOutlier)));
    CU(c, cudaMalloc(&s.outlier_count, sizeof(unsigned)));
    CU(c, cudaMalloc(&s.near, local_cap * sizeof(float)));
    CU(c, cudaMalloc(&s.cand, local_cap * sizeof(unsigned)));
    CU(c, cudaMalloc(&s.cand_count, sizeof(unsigned)));
    CU(c, cudaMalloc(&s.pair_count, sizeof(unsigned long long)));
    return NBODY_OK;
}

// The fp64 force kernel's variants (force_f64.cuh): MATH 0/1 exist for one body per thread and general masses only.
template <typename F> void pick_f64_kernel(int math, bool zero_soft, bool uniform, int R, F&& f)
{
    if (math == 0) return zero_soft ? f(force_f64_kernel<0, true>) : f(force_f64_kernel<0, false>);
    if (math == 1) return zero_soft ? f(force_f64_kernel<1, true>) : f(force_f64_kernel<1, false>);
    if (R == 2) {
        if (uniform) return zero_soft ? f(force_f64_kernel<2, true, true, 2>) : f(force_f64_kernel<2, false, true, 2>);
        return zero_soft ? f(force_f64_kernel<2, true, false, 2>) : f(force_f64_kernel<2, false, false, 2>);
    }
    if (uniform) return zero_soft ? f(force_f64_kernel<2, true, true, 1>) : f(force_f64_kernel<2, false, true, 1>);
    return zero_soft ? f(force_f64_kernel<2, true, false, 1>) : f(force_f64_kernel<2, false, false, 1>);
}

template <typename F> int for_each_f64_kernel(F&& f, nbody_ctx* c)
{
    for (int math = 0; math < 3; ++math)
        for (int z = 0; z < 2; ++z)
            for (int u = 0; u < 2; ++u)
                for (int R = 1; R <= 2; ++R) {
                    cudaError_t e = cudaSuccess;
                    pick_f64_kernel(math, z != 0, u != 0, R, [&](auto kernel) { e = f(kernel); });
                    if (e != cudaSuccess) return fail(c, NBODY_ECUDA, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
                }
    return NBODY_OK;
}

int set_kernel_attrs(nbody_ctx* c)
{
    const int sm32 = int(force_f32_smem_bytes());
    const int sm64 = int(force_f64_smem_bytes());
    CU(c, cudaFuncSetAttribute(force_f32_kernel<2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm32));
    CU(c, cudaFuncSetAttribute(force_f32_kernel<2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm32));
    CU(c, cudaFuncSetAttribute(force_f32_kernel<4, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm32));
    CU(c, cudaFuncSetAttribute(force_f32_kernel<4, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm32));
    CU(c, cudaFuncSetAttribute(force_f32_kernel<4, false, kUniUnroll, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm32));
    CU(c, cudaFuncSetAttribute(force_f32_kernel<4, true, kUniUnroll, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm32));
    CU(c, cudaFuncSetAttribute(force_f32_near_kernel<4, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm32));
    CU(c, cudaFuncSetAttribute(force_f32_near_kernel<4, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm32));
    CU(c, cudaFuncSetAttribute(force_f32_near_kernel<4, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm32));
    CU(c, cudaFuncSetAttribute(force_f32_near_kernel<4, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm32));
    CU(c, cudaFuncSetAttribute(small_step_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(small_smem_bytes<float>())));
    CU(c, cudaFuncSetAttribute(small_step_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(small_smem_bytes<double>())));
    return for_each_f64_kernel([&](auto kernel) { return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, sm64); }, c);
}

int create_common(nbody_ctx** out, nbody_prec prec, int world, size_t capacity, nbody_ctx** made)
{
    if (!out) return NBODY_EINVAL;
    *out = nullptr;
    if ((prec != NBODY_F32 && prec != NBODY_F64) || world < 1 || capacity == 0) return NBODY_EINVAL;
    nbody_ctx* c = new (std::nothrow) nbody_ctx();
    if (!c) return NBODY_ENOMEM;
    c->prec = prec;
    c->esz = prec == NBODY_F32 ? 4 : 8;
    c->world = world;
    c->capacity = capacity;
    if (prec == NBODY_F32) {
        c->R = env_int("NBODY_R", 4);
        if (c->R != 2 && c->R != 4) c->R = 4;
        c->i_tile = kF32Threads * c->R;
    } else {
        c->f64_math = env_int("NBODY_F64_MATH", 2);
        c->f64_R_forced = env_int("NBODY_F64_R", 0);
        c->R = f64_bodies_per_thread(c, capacity);  // for cap_pad: the largest i-tile an upload may pick
        c->i_tile = kThreads * c->R;
        c->j_unit = kJUnitF64;
    }
    c->small_path = env_int("NBODY_SMALL", 1) != 0 && !(prec == NBODY_F64 && c->f64_math != 2);  // (its pair arithmetic is MATH 2's)
    c->cap_pad = round_up(capacity, pad_unit(c));
    *made = c;
    return NBODY_OK;
}

}  // namespace

// ============================================================================
// lifetime
// ============================================================================

extern "C" int nbody_device_count(void)
{
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return NBODY_ECUDA;
    }
    return n;
}

extern "C" const char* nbody_version(void) { return "nbody-b200 0.1 (sm_100a)"; }

extern "C" const char* nbody_strerror(int status)
{
    switch (status) {
    case NBODY_OK: return "ok";
    case NBODY_EINVAL: return "invalid argument";
    case NBODY_ECUDA: return "CUDA error";
    case NBODY_ENCCL: return "NCCL error";
    case NBODY_ENOMEM: return "out of memory";
    case NBODY_EIO: return "file I/O error";
    default: return "unknown status";
    }
}

extern "C" const char* nbody_last_error(const nbody_ctx* ctx) { return ctx ? ctx->err.c_str() : ""; }

extern "C" void nbody_destroy(nbody_ctx* c)
{
    if (!c) return;
    for (Shard& s : c->shards) {
        cudaSetDevice(s.device);
        if (s.own_stream) cudaStreamSynchronize(s.own_stream);
        if (s.comm && nccl().ok()) nccl().CommDestroy(s.comm);
        for (cudaEvent_t e : s.ev) cudaEventDestroy(e);
        for (void* m : s.ipc_opened) cudaIpcCloseMemHandle(m);
        cudaFree(s.flags);
        cudaFree(s.xerr);
        cudaFree(s.pos[0]);
        cudaFree(s.pos[1]);
        cudaFree(s.stage);
        cudaFree(s.vel);
        cudaFree(s.acc);
        cudaFree(s.temp);
        cudaFree(s.partial);
        cudaFree(s.counters);
        cudaFree(s.energy_out);
        cudaFree(s.radius);
        cudaFree(s.near);
        cudaFree(s.cand);
        cudaFree(s.cand_count);
        cudaFree(s.cand_all);
        cudaFree(s.cand_rec);
        cudaFree(s.big);
        cudaFree(s.pairs);
        cudaFree(s.pair_count);
        cudaFree(s.hist);
        cudaFree(s.hist_alt);
        cudaFree(s.hist_src);
        cudaFree(s.ribbons);
        cudaFree(s.mat_type);
        cudaFree(s.mat_seed);
        cudaFree(s.outliers);
        cudaFree(s.outlier_count);
        cudaFree(s.temp_full);
        render_scratch_free(s.render_scratch);
        cudaFree(s.image);
        cudaFree(s.image_ids);
        if (s.own_stream) cudaStreamDestroy(s.own_stream);
    }
    delete c;
}

static int finish_create(nbody_ctx* c, nbody_ctx** out, int rc)
{
    if (rc != NBODY_OK) {
        // keep the message reachable: callers get no ctx on failure, so print it
        fprintf(stderr, "libnbody: create failed: %s\n", c->err.c_str());
        nbody_destroy(c);
        return rc;
    }
    *out = c;
    return NBODY_OK;
}

// Map every rank's position buffers and flag array on every device (exchange.cuh).
// Single process: peer access between the context's devices.  Rank mode: CUDA IPC
// handles, gathered once over the NCCL communicator that already exists.  If any rank
// cannot map a peer, ALL ranks fall back to ncclAllGather (c->p2p stays false).
static int setup_peer_exchange(nbody_ctx* c)
{
    const char* mode = getenv("NBODY_EXCHANGE");
    if ((mode && strcmp(mode, "nccl") == 0) || c->world > kMaxPeers) return NBODY_OK;
    if (!c->rank_mode) {
        for (Shard& a : c->shards)
            for (Shard& b : c->shards) {
                if (a.device == b.device) continue;
                int can = 0;
                CU(c, cudaDeviceCanAccessPeer(&can, a.device, b.device));
                if (!can) return NBODY_OK;
            }
        for (Shard& a : c->shards) {
            CU(c, cudaSetDevice(a.device));
            for (Shard& b : c->shards) {
                if (a.device != b.device) {
                    cudaError_t e = cudaDeviceEnablePeerAccess(b.device, 0);
                    if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
                    else if (e != cudaSuccess) return fail(c, NBODY_ECUDA, "cudaDeviceEnablePeerAccess: %s", cudaGetErrorString(e));
                }
                a.peer_pos[b.rank][0] = b.pos[0];
                a.peer_pos[b.rank][1] = b.pos[1];
                a.peer_flags.flags[b.rank] = b.flags;
            }
        }
        c->p2p = true;
        return NBODY_OK;
    }
    struct Blob {
        cudaIpcMemHandle_t pos0, pos1, flags;
        int ok;
    };
    Shard& s = c->shards[0];
    CU(c, cudaSetDevice(s.device));
    std::vector<Blob> all(c->world);
    Blob mine;
    memset(&mine, 0, sizeof mine);
    mine.ok = cudaIpcGetMemHandle(&mine.pos0, s.pos[0]) == cudaSuccess && cudaIpcGetMemHandle(&mine.pos1, s.pos[1]) == cudaSuccess &&
              cudaIpcGetMemHandle(&mine.flags, s.flags) == cudaSuccess;
    cudaGetLastError();
    Blob* dev = nullptr;
    CU(c, cudaMalloc(&dev, sizeof(Blob) * c->world));
    auto gather = [&](const Blob& b) -> int {
        CU(c, cudaMemcpyAsync(dev + c->my_rank, &b, sizeof(Blob), cudaMemcpyHostToDevice, s.stream));
        NC(c, nccl().AllGather(dev + c->my_rank, dev, sizeof(Blob), ncclUint8_, s.comm, s.stream));
        CU(c, cudaMemcpyAsync(all.data(), dev, sizeof(Blob) * c->world, cudaMemcpyDeviceToHost, s.stream));
        CU(c, cudaStreamSynchronize(s.stream));
        return NBODY_OK;
    };
    int rc = gather(mine);
    if (rc != NBODY_OK) return rc;
    bool ok = true;
    for (int r = 0; r < c->world; ++r) ok = ok && all[r].ok;
    if (ok) {
        for (int r = 0; r < c->world && ok; ++r) {
            if (r == c->my_rank) {
                s.peer_pos[r][0] = s.pos[0];
                s.peer_pos[r][1] = s.pos[1];
                s.peer_flags.flags[r] = s.flags;
                continue;
            }
            void* ptr[3] = {nullptr, nullptr, nullptr};
            const cudaIpcMemHandle_t* h[3] = {&all[r].pos0, &all[r].pos1, &all[r].flags};
            for (int k = 0; k < 3 && ok; ++k) {
                if (cudaIpcOpenMemHandle(&ptr[k], *h[k], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
                    cudaGetLastError();
                    ok = false;
                } else {
                    s.ipc_opened.push_back(ptr[k]);
                }
            }
            s.peer_pos[r][0] = ptr[0];
            s.peer_pos[r][1] = ptr[1];
            s.peer_flags.flags[r] = static_cast<unsigned long long*>(ptr[2]);
        }
    }
    mine.ok = ok;  // second round: did EVERY rank map every peer?
    rc = gather(mine);
    cudaFree(dev);
    if (rc != NBODY_OK) return rc;
    for (int r = 0; r < c->world; ++r) ok = ok && all[r].ok;
    c->p2p = ok;
    return NBODY_OK;
}

static int create_body(nbody_ctx* c, const int* devs, int n_local_shards, const void* nccl_id)
{
    int ndev = 0;
    CU(c, cudaGetDeviceCount(&ndev));
    if (ndev < 1) return fail(c, NBODY_ECUDA, "no CUDA device");
    c->shards.resize(n_local_shards);
    for (int k = 0; k < n_local_shards; ++k) {
        Shard& s = c->shards[k];
        s.device = devs[k];
        s.rank = c->rank_mode ? c->my_rank : k;
        if (s.device < 0 || s.device >= ndev) return fail(c, NBODY_EINVAL, "device %d out of range (%d visible)", s.device, ndev);
        cudaDeviceProp prop;
        CU(c, cudaGetDeviceProperties(&prop, s.device));
        if (prop.major < 10)
            return fail(c, NBODY_ECUDA, "device %d is sm_%d%d; this library is built for sm_100a only", s.device, prop.major, prop.minor);
        if (k == 0) c->sm_count = prop.multiProcessorCount;
        int rc = alloc_shard(c, s);
        if (rc != NBODY_OK) return rc;
        rc = set_kernel_attrs(c);
        if (rc != NBODY_OK) return rc;
    }
    if (c->world > 1) {
        if (!nccl().ok()) return fail(c, NBODY_ENCCL, "%s", nccl().error.c_str());
        if (c->rank_mode) {
            if (!nccl_id) return fail(c, NBODY_EINVAL, "nccl_id is required when world > 1");
            ncclUniqueId id;
            memcpy(&id, nccl_id, sizeof id);
            CU(c, cudaSetDevice(c->shards[0].device));
            NC(c, nccl().CommInitRank(&c->shards[0].comm, c->world, id, c->my_rank));
        } else {
            std::vector<ncclComm_t> comms(n_local_shards);
            NC(c, nccl().CommInitAll(comms.data(), n_local_shards, devs));
            for (int k = 0; k < n_local_shards; ++k) c->shards[k].comm = comms[k];
        }
        return setup_peer_exchange(c);
    }
    return NBODY_OK;
}

extern "C" int nbody_create(nbody_ctx** out, nbody_prec prec, int n_gpus, const int* device_ids, size_t capacity)
{
    nbody_ctx* c = nullptr;
    int rc = create_common(out, prec, n_gpus, capacity, &c);
    if (rc != NBODY_OK) return rc;
    std::vector<int> devs(n_gpus);
    for (int k = 0; k < n_gpus; ++k) devs[k] = device_ids ? device_ids[k] : k;
    return finish_create(c, out, create_body(c, devs.data(), n_gpus, nullptr));
}

extern "C" int nbody_create_rank(nbody_ctx** out, nbody_prec prec, int rank, int world, int device,
                                 const void* nccl_id, size_t capacity)
{
    if (rank < 0 || rank >= world) return NBODY_EINVAL;
    nbody_ctx* c = nullptr;
    int rc = create_common(out, prec, world, capacity, &c);
    if (rc != NBODY_OK) return rc;
    c->rank_mode = true;
    c->my_rank = rank;
    return finish_create(c, out, create_body(c, &device, 1, nccl_id));
}

extern "C" int nbody_nccl_unique_id(void* id_out)
{
    if (!id_out) return NBODY_EINVAL;
    if (!nccl().ok()) {
        fprintf(stderr, "libnbody: %s\n", nccl().error.c_str());
        return NBODY_ENCCL;
    }
    ncclUniqueId id;
    if (nccl().GetUniqueId(&id) != 0) return NBODY_ENCCL;
    memcpy(id_out, &id, sizeof id);
    return NBODY_OK;
}

extern "C" int nbody_set_heat_decay(nbody_ctx* c, double heat_decay)
{
    if (!c) return NBODY_EINVAL;
    c->heat_decay = heat_decay;
    return NBODY_OK;
}

extern "C" int nbody_set_stream(nbody_ctx* c, void* cuda_stream)
{
    if (!c) return NBODY_EINVAL;
    Shard& s = c->shards[0];
    CU(c, cudaSetDevice(s.device));
    CU(c, cudaStreamSynchronize(s.stream));
    s.stream = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : s.own_stream;
    return NBODY_OK;
}

extern "C" int nbody_sync(nbody_ctx* c)
{
    if (!c) return NBODY_EINVAL;
    for (Shard& s : c->shards) {
        CU(c, cudaSetDevice(s.device));
        CU(c, cudaStreamSynchronize(s.stream));
    }
    return check_exchange(c);
}

extern "C" int nbody_get_info(nbody_ctx* c, nbody_info* o)
{
    if (!c || !o) return NBODY_EINVAL;
    memset(o, 0, sizeof *o);
    o->n = c->n;
    o->n_padded = c->n_pad;
    o->i_begin = std::min(c->shards[0].i_begin, c->n);
    const Shard& last = c->shards.back();
    o->i_end = std::min(last.i_begin + last.n_local, c->n);
    o->n_gpus = int(c->shards.size());
    o->world = c->world;
    o->rank = c->rank_mode ? c->my_rank : 0;
    o->sm_count = c->sm_count;
    o->j_chunks = c->j_chunks;
    o->i_tile = c->i_tile;
    o->kernel_launches = c->launches;
    o->exchange = c->world == 1 ? 0 : (c->p2p ? 2 : 1);
    o->uniform_mass = c->uniform ? 1 + int(c->outliers.size()) : 0;
    return NBODY_OK;
}

// ============================================================================
// state
// ============================================================================

namespace {

template <typename Real> int upload_t(nbody_ctx* c, size_t n, const Real* pm, const Real* vel, const Real* acc, const Real* temp)
{
    const size_t v4 = 4 * sizeof(Real);
    for (Shard& s : c->shards) {
        CU(c, cudaSetDevice(s.device));
        s.n_local = c->n_pad / c->world;
        s.i_begin = size_t(s.rank) * s.n_local;
        const size_t lo = std::min(s.i_begin, n), hi = std::min(s.i_begin + s.n_local, n);
        s.n_local_real = hi - lo;
        // positions + masses: host AoS -> staging -> device layout (both buffers)
        CU(c, cudaMemcpyAsync(s.stage, pm, n * v4, cudaMemcpyHostToDevice, s.stream));
        const unsigned blocks = unsigned((c->n_pad + 255) / 256);
        const bool detect = c->detect_masses && &s == &c->shards[0];  // every shard packs the same array: one looks
        if (detect) CU(c, cudaMemsetAsync(s.outlier_count, 0, sizeof(unsigned), s.stream));
        pack_pos_kernel<Real><<<blocks, 256, 0, s.stream>>>(static_cast<const Real*>(s.stage), static_cast<Real*>(s.pos[0]),
                                                           static_cast<Real*>(s.pos[1]), n, c->n_pad, Real(sizeof(Real) == 8 ? 1e150 : 1e18), Real(c->uniform_mass),
                                                           detect ? s.outlier_count : nullptr, s.outliers);
        c->launches++;
        CU(c, cudaGetLastError());
        CU(c, cudaMemsetAsync(s.vel, 0, s.n_local * v4, s.stream));
        CU(c, cudaMemsetAsync(s.acc, 0, s.n_local * v4, s.stream));
        CU(c, cudaMemsetAsync(s.temp, 0, s.n_local * sizeof(Real), s.stream));
        if (s.n_local_real) {
            CU(c, cudaMemcpyAsync(s.vel, vel + 4 * lo, s.n_local_real * v4, cudaMemcpyHostToDevice, s.stream));
            if (acc) CU(c, cudaMemcpyAsync(s.acc, acc + 4 * lo, s.n_local_real * v4, cudaMemcpyHostToDevice, s.stream));
            if (temp) CU(c, cudaMemcpyAsync(s.temp, temp + lo, s.n_local_real * sizeof(Real), cudaMemcpyHostToDevice, s.stream));
        }
        // per-chunk partial sums
        const size_t need = c->j_chunks > 1 ? size_t(c->j_chunks) * s.n_local * v4 : 0;
        if (need > s.partial_bytes) {
            CU(c, cudaStreamSynchronize(s.stream));
            cudaFree(s.partial);
            s.partial = nullptr;
            s.partial_bytes = 0;
            CU(c, cudaMalloc(&s.partial, need));
            s.partial_bytes = need;
        }
    }
    // host buffers are borrowed for the duration of the call only
    for (Shard& s : c->shards) {
        CU(c, cudaSetDevice(s.device));
        CU(c, cudaStreamSynchronize(s.stream));
    }
    return NBODY_OK;
}

template <typename Real> int download_t(nbody_ctx* c, Real* pm, Real* vel, Real* acc, Real* temp)
{
    const size_t v4 = 4 * sizeof(Real);
    if (pm && c->n) {
        Shard& s = c->shards[0];  // every shard holds the complete position array
        CU(c, cudaSetDevice(s.device));
        const unsigned blocks = unsigned((c->n + 255) / 256);
        unpack_pos_kernel<Real><<<blocks, 256, 0, s.stream>>>(static_cast<const Real*>(s.pos[c->cur]), static_cast<Real*>(s.stage), c->n);
        c->launches++;
        CU(c, cudaGetLastError());
        CU(c, cudaMemcpyAsync(pm, s.stage, c->n * v4, cudaMemcpyDeviceToHost, s.stream));
    }
    for (Shard& s : c->shards) {
        CU(c, cudaSetDevice(s.device));
        const size_t lo = std::min(s.i_begin, c->n);
        if (s.n_local_real) {
            if (vel) CU(c, cudaMemcpyAsync(vel + 4 * lo, s.vel, s.n_local_real * v4, cudaMemcpyDeviceToHost, s.stream));
            if (acc) CU(c, cudaMemcpyAsync(acc + 4 * lo, s.acc, s.n_local_real * v4, cudaMemcpyDeviceToHost, s.stream));
            if (temp) CU(c, cudaMemcpyAsync(temp + lo, s.temp, s.n_local_real * sizeof(Real), cudaMemcpyDeviceToHost, s.stream));
        }
    }
    for (Shard& s : c->shards) {
        CU(c, cudaSetDevice(s.device));
        CU(c, cudaStreamSynchronize(s.stream));
    }
    return check_exchange(c);
}

// After upload_t's final synchronise: what pack_pos_kernel found.  Up to kMaxOutliers exceptions keep
// the equal-mass sweep (they are corrected per i-body); the list is sorted so that the order of the
// corrections, and with it every bit of the result, does not depend on the order the atomics landed in.
int finish_mass_detection(nbody_ctx* c)
{
    Shard& s0 = c->shards[0];
    CU(c, cudaSetDevice(s0.device));
    unsigned count = 0;
    CU(c, cudaMemcpy(&count, s0.outlier_count, sizeof(unsigned), cudaMemcpyDeviceToHost));
    // (exceptions are corrected in the fp32 epilogue only: an fp64 context takes the sweep for exactly equal masses)
    const size_t max_out = c->prec == NBODY_F64 ? 0 : std::min<size_t>(size_t(std::max(0, env_int("NBODY_OUTLIERS", kMaxOutliers))), kMaxOutliers);
    if (count > max_out) return NBODY_OK;  // too many different masses: the general kernel
    c->uniform = true;
    if (count == 0) return NBODY_OK;
    c->outliers.resize(count);
    CU(c, cudaMemcpy(c->outliers.data(), s0.outliers, count * sizeof(MassOutlier), cudaMemcpyDeviceToHost));
    std::sort(c->outliers.begin(), c->outliers.end(), [](const MassOutlier& a, const MassOutlier& b) { return a.index < b.index; });
    for (Shard& s : c->shards) {
        CU(c, cudaSetDevice(s.device));
        CU(c, cudaMemcpy(s.outliers, c->outliers.data(), count * sizeof(MassOutlier), cudaMemcpyHostToDevice));
    }
    return NBODY_OK;
}

}  // namespace

extern "C" int nbody_upload(nbody_ctx* c, size_t n, const void* pos_mass, const void* vel, const void* acc, const void* temp)
{
    if (!c) return NBODY_EINVAL;
    if (n == 0 || n > c->capacity) return fail(c, NBODY_EINVAL, "upload: n=%zu outside 1..capacity=%zu", n, c->capacity);
    if (!pos_mass || !vel) return fail(c, NBODY_EINVAL, "upload: pos_mass and vel are required");
    c->n = n;
    if (c->prec == NBODY_F64) {
        c->R = f64_bodies_per_thread(c, n);
        c->i_tile = kThreads * c->R;
    }
    c->n_pad = round_up(n, pad_unit(c));
    c->n_j = round_up(n, size_t(c->j_unit));
    c->cur = 0;
    c->has_temp = temp != nullptr;
    c->uniform = false;
    c->outliers.clear();
    c->detect_masses = false;
    const bool f32_sweep = c->prec == NBODY_F32 && c->R == 4, f64_sweep = c->prec == NBODY_F64 && c->f64_math == 2;
    if ((f32_sweep || f64_sweep) && env_int("NBODY_UNIFORM", 1) != 0) {
        // Equal-mass sweep (force_f32.cuh / force_f64.cuh, UNIFORM): the mass most bodies share, if there is one,
        // by majority vote over an evenly spaced sample; pack_pos_kernel then compares every mass with it.
        const size_t samples = std::min<size_t>(n, 1024);
        double m0 = 0.0;
        size_t votes = 0;
        for (size_t k = 0; k < samples; ++k) {
            const size_t at = 4 * (k * n / samples) + 3;
            const double m = f32_sweep ? double(static_cast<const float*>(pos_mass)[at]) : static_cast<const double*>(pos_mass)[at];
            if (votes == 0) m0 = m;
            votes += (m == m0) ? 1 : size_t(-1);
        }
        c->uniform_mass = m0;
        c->detect_masses = m0 > 0.0 && std::isfinite(m0);
    }
    c->has_radius = false;  // a new body list: radii must be set again (Body.radius follows mass, idx:221)
    if (c->hist_n != n) {  // history ring: refilled at the next push (trails.cuh) unless nbody_history_remap carries it over
        if (c->hist_n) c->hist_prev_n = c->hist_n;  // kept over further uploads: the ring itself is not touched
        c->hist_n = 0;
    }
    c->has_materials = false;
    c->near_valid = false;
    choose_chunks(c);
    int rc = c->prec == NBODY_F32
                 ? upload_t<float>(c, n, (const float*)pos_mass, (const float*)vel, (const float*)acc, (const float*)temp)
                 : upload_t<double>(c, n, (const double*)pos_mass, (const double*)vel, (const double*)acc, (const double*)temp);
    if (rc == NBODY_OK && c->detect_masses) rc = finish_mass_detection(c);
    c->detect_masses = false;
    return rc;
}

extern "C" int nbody_download(nbody_ctx* c, void* pos_mass, void* vel, void* acc, void* temp)
{
    if (!c) return NBODY_EINVAL;
    if (c->n == 0) return fail(c, NBODY_EINVAL, "download before upload");
    return c->prec == NBODY_F32 ? download_t<float>(c, (float*)pos_mass, (float*)vel, (float*)acc, (float*)temp)
                                : download_t<double>(c, (double*)pos_mass, (double*)vel, (double*)acc, (double*)temp);
}

// ============================================================================
// the hot path
// ============================================================================

namespace {

constexpr size_t kEventPool = 2048;  // pairs

int fold_timing(nbody_ctx* c, Shard& s)
{
    if (s.ev_used == 0) return NBODY_OK;
    CU(c, cudaSetDevice(s.device));
    CU(c, cudaStreamSynchronize(s.stream));
    for (size_t k = 0; k < s.ev_used; k += 2) {
        float ms = 0.f;
        CU(c, cudaEventElapsedTime(&ms, s.ev[k], s.ev[k + 1]));
        c->t_ms += ms;
        c->t_launches++;
    }
    s.ev_used = 0;
    return NBODY_OK;
}

// Complete the exchange of the slices just written to pos[which].
// Peer-store mode: the update kernel already stored every slice on every rank; queue
// the epoch barrier.  Otherwise: ncclAllGather of the own slices, in place.
int exchange_positions(nbody_ctx* c, int which)
{
    if (c->world == 1) return NBODY_OK;
    if (c->p2p) {
        c->epoch++;
        const unsigned long long timeout_ns = (unsigned long long)env_int("NBODY_EXCHANGE_TIMEOUT_S", 60) * 1000000000ull;
        for (Shard& s : c->shards) {
            CU(c, cudaSetDevice(s.device));
            exchange_barrier_kernel<<<1, 32, 0, s.stream>>>(s.peer_flags, s.rank, c->world, c->epoch, timeout_ns, s.xerr);
            c->launches++;
            CU(c, cudaGetLastError());
        }
        return NBODY_OK;
    }
    const size_t v4 = 4 * c->esz;
    if (!c->rank_mode) NC(c, nccl().GroupStart());
    for (Shard& s : c->shards) {
        if (!c->rank_mode) CU(c, cudaSetDevice(s.device));
        char* base = static_cast<char*>(s.pos[which]);
        const size_t bytes = s.n_local * v4;
        NC(c, nccl().AllGather(base + size_t(s.rank) * bytes, base, bytes, ncclUint8_, s.comm, s.stream));
    }
    if (!c->rank_mode) NC(c, nccl().GroupEnd());
    return NBODY_OK;
}

// Where a position update of shard s goes: its own `which` buffer, plus every peer's in peer-store mode.
template <typename Real> PosOut<Real> make_out(nbody_ctx* c, Shard& s, int which)
{
    PosOut<Real> o;
    o.n = 0;
    o.dst[o.n++] = static_cast<Real*>(s.pos[which]);
    if (c->p2p)
        for (int r = 0; r < c->world; ++r)
            if (r != s.rank) o.dst[o.n++] = static_cast<Real*>(s.peer_pos[r][which]);
    return o;
}

// After a synchronisation: did an exchange barrier time out on a lost peer?
int check_exchange(nbody_ctx* c)
{
    if (!c->p2p) return NBODY_OK;
    for (Shard& s : c->shards) {
        unsigned e = 0;
        CU(c, cudaSetDevice(s.device));
        CU(c, cudaMemcpy(&e, s.xerr, sizeof e, cudaMemcpyDeviceToHost));
        if (e) return fail(c, NBODY_ECUDA, "position exchange timed out waiting for rank %u", e - 1);
    }
    return NBODY_OK;
}

template <typename Real> ForceParams<Real> make_params(nbody_ctx* c, Shard& s, int mode, double dt, double G, double eps)
{
    ForceParams<Real> p;
    p.pos_cur = static_cast<const Real*>(s.pos[c->cur]);
    p.out = make_out<Real>(c, s, c->cur ^ 1);
    p.vel = static_cast<Real*>(s.vel);
    p.acc = static_cast<Real*>(s.acc);
    p.partial = static_cast<Real*>(s.partial);
    p.temp = c->has_temp ? static_cast<Real*>(s.temp) : nullptr;
    p.near = reinterpret_cast<Real*>(s.near);  // written by the fp32 NEAR kernels only
    p.outliers = s.outliers;
    p.n_outliers = 0;
    p.bulk_mass = Real(0);
    p.tile_counter = s.counters;
    p.n_j = c->n_j;
    p.i_begin = s.i_begin;
    p.n_local = s.n_local;
    p.chunk_len = c->chunk_len;
    p.j_chunks = c->j_chunks;
    p.mode = mode;
    p.G = Real(G);
    p.eps2 = Real(eps) * Real(eps);  // epsSq = softening*softening, idx:269
    p.half_dt = Real(0.5) * Real(dt);  // halfDt = 0.5*dt, idx:248
    p.dt = Real(dt);
    p.heat_decay = Real(c->heat_decay);
    return p;
}

int launch_force(nbody_ctx* c, int mode, double dt, double G, double eps)
{
    // With radii set, the fp32 sweep also tracks each body's nearest neighbour for the
    // collision broad phase (contacts.cuh).  The result describes the positions the
    // sweep READ, so it stays valid only when the epilogue does not drift them.
    const bool track_near = c->has_radius && c->prec == NBODY_F32 && c->R == 4 && env_int("NBODY_NEAR", 1) != 0;
    c->near_valid = track_near && mode != kForceKick2Kick1;
    c->near_eps2 = double(float(eps) * float(eps));
    for (Shard& s : c->shards) {
        CU(c, cudaSetDevice(s.device));
        const unsigned grid = unsigned(s.n_local / c->i_tile) * unsigned(c->j_chunks);
        const bool time_it = c->timing && &s == &c->shards[0];
        if (time_it) {
            if (s.ev_used + 2 > s.ev.size()) {
                if (s.ev.size() < 2 * kEventPool) {
                    for (int k = 0; k < 2; ++k) {
                        cudaEvent_t e;
                        CU(c, cudaEventCreate(&e));
                        s.ev.push_back(e);
                    }
                } else {
                    int rc = fold_timing(c, s);
                    if (rc != NBODY_OK) return rc;
                }
            }
            CU(c, cudaEventRecord(s.ev[s.ev_used], s.stream));
        }
        if (c->prec == NBODY_F32) {
            ForceParams<float> p = make_params<float>(c, s, mode, dt, G, eps);
            const bool uniform = c->uniform;
            if (uniform && c->outliers.empty()) {
                p.G = float(G) * float(c->uniform_mass);  // equal masses: the common m joins G (force_f32.cuh, UNIFORM)
            } else if (uniform) {
                p.n_outliers = int(c->outliers.size());   // ... but for a few bodies, corrected per i-body
                p.bulk_mass = float(c->uniform_mass);
            }
            const size_t sm = force_f32_smem_bytes();
            const bool z = (eps == 0.0);
            const dim3 blk(kF32Threads + 32);
            if (track_near && uniform) {
                if (z) force_f32_near_kernel<4, true, true><<<grid, blk, sm, s.stream>>>(p);
                else force_f32_near_kernel<4, false, true><<<grid, blk, sm, s.stream>>>(p);
            } else if (track_near) {
                if (z) force_f32_near_kernel<4, true><<<grid, blk, sm, s.stream>>>(p);
                else force_f32_near_kernel<4, false><<<grid, blk, sm, s.stream>>>(p);
            } else if (uniform) {
                if (z) force_f32_kernel<4, true, kUniUnroll, true><<<grid, blk, sm, s.stream>>>(p);
                else force_f32_kernel<4, false, kUniUnroll, true><<<grid, blk, sm, s.stream>>>(p);
            } else if (c->R == 2) {
                if (z) force_f32_kernel<2, true><<<grid, blk, sm, s.stream>>>(p);
                else force_f32_kernel<2, false><<<grid, blk, sm, s.stream>>>(p);
            } else {
                if (z) force_f32_kernel<4, true><<<grid, blk, sm, s.stream>>>(p);
                else force_f32_kernel<4, false><<<grid, blk, sm, s.stream>>>(p);
            }
        } else {
            const ForceParams<double> p = make_params<double>(c, s, mode, dt, G, eps);
            const dim3 blk(kThreads + 32);
            const size_t sm = force_f64_smem_bytes();
            ForceParams<double> q = p;
            const bool uni = c->uniform && c->f64_math == 2;
            if (uni) q.G = G * c->uniform_mass;  // equal masses: the common m joins G (force_f64.cuh, UNIFORM)
            pick_f64_kernel(c->f64_math, eps == 0.0, uni, c->R, [&](auto kernel) { kernel<<<grid, blk, sm, s.stream>>>(q); });
        }
        c->launches++;
        CU(c, cudaGetLastError());
        if (time_it) {
            CU(c, cudaEventRecord(s.ev[s.ev_used + 1], s.stream));
            s.ev_used += 2;
        }
    }
    return NBODY_OK;
}

// The whole nsteps x integrate(dt) of a small system in one cluster launch (small.cuh).
template <typename Real> int launch_small_steps_t(nbody_ctx* c, Shard& s, double dt, double G, double eps, int nsteps)
{
    SmallParams<Real> p;
    p.pos = static_cast<Real*>(s.pos[c->cur]);
    p.vel = static_cast<Real*>(s.vel);
    p.acc = static_cast<Real*>(s.acc);
    p.temp = c->has_temp ? static_cast<Real*>(s.temp) : nullptr;
    p.n = int(c->n);
    p.nsteps = nsteps;
    p.G = Real(G);
    p.eps2 = Real(eps) * Real(eps);
    p.r2_floor = eps == 0.0 ? Real(1e-20) : Real(0);
    p.half_dt = Real(0.5) * Real(dt);
    p.dt = Real(dt);
    p.heat_decay = Real(c->heat_decay);
    small_step_kernel<Real><<<kSmallCluster, kSmallThreads, small_smem_bytes<Real>(), s.stream>>>(p);
    c->launches++;
    CU(c, cudaGetLastError());
    return NBODY_OK;
}

int launch_small_steps(nbody_ctx* c, double dt, double G, double eps, int nsteps)
{
    Shard& s = c->shards[0];
    CU(c, cudaSetDevice(s.device));
    c->near_valid = false;  // the bodies move and nobody tracked neighbours: a contact search scans every body
    return c->prec == NBODY_F32 ? launch_small_steps_t<float>(c, s, dt, G, eps, nsteps)
                                : launch_small_steps_t<double>(c, s, dt, G, eps, nsteps);
}

// idx:248-253 for every shard's own range, written to the OTHER buffer of every rank,
// then the exchange, then the buffers swap (same flow as a fused force step).
int launch_kick1_drift(nbody_ctx* c, double dt)
{
    const int nxt = c->cur ^ 1;
    for (Shard& s : c->shards) {
        CU(c, cudaSetDevice(s.device));
        const unsigned blocks = unsigned((s.n_local + 255) / 256);
        if (c->prec == NBODY_F32)
            kick1_drift_kernel<float><<<blocks, 256, 0, s.stream>>>((const float*)s.pos[c->cur], make_out<float>(c, s, nxt), (float*)s.vel,
                                                                   (const float*)s.acc, s.i_begin, s.n_local, 0.5f * float(dt), float(dt));
        else
            kick1_drift_kernel<double><<<blocks, 256, 0, s.stream>>>((const double*)s.pos[c->cur], make_out<double>(c, s, nxt), (double*)s.vel,
                                                                    (const double*)s.acc, s.i_begin, s.n_local, 0.5 * dt, dt);
        c->launches++;
        CU(c, cudaGetLastError());
    }
    c->near_valid = false;
    const int rc = exchange_positions(c, nxt);
    if (rc == NBODY_OK) c->cur = nxt;
    return rc;
}

int check_ready(nbody_ctx* c, const char* what)
{
    if (!c) return NBODY_EINVAL;
    if (c->n == 0) return fail(c, NBODY_EINVAL, "%s before upload", what);
    return NBODY_OK;
}

}  // namespace

extern "C" int nbody_step(nbody_ctx* c, double dt, double G, double softening, int nsteps)
{
    int rc = check_ready(c, "step");
    if (rc != NBODY_OK) return rc;
    if (nsteps < 0 || softening < 0.0) return fail(c, NBODY_EINVAL, "step: nsteps=%d softening=%g", nsteps, softening);
    if (nsteps == 0) return NBODY_OK;
    if (c->small_path && c->world == 1 && c->n <= size_t(kSmallMaxN) && !c->timing) return launch_small_steps(c, dt, G, softening, nsteps);
    // First loop of integrate() for the first step (idx:250-253); later steps
    // get theirs from the previous force kernel's epilogue, so positions cross
    // HBM once per step.
    rc = launch_kick1_drift(c, dt);
    if (rc != NBODY_OK) return rc;
    for (int s = 0; s < nsteps; ++s) {
        const bool more = s + 1 < nsteps;
        rc = launch_force(c, more ? kForceKick2Kick1 : kForceKick2, dt, G, softening);
        if (rc != NBODY_OK) return rc;
        if (more) {
            rc = exchange_positions(c, c->cur ^ 1);
            if (rc != NBODY_OK) return rc;
            c->cur ^= 1;
        }
    }
    return NBODY_OK;
}

extern "C" int nbody_kick1_drift(nbody_ctx* c, double dt)
{
    int rc = check_ready(c, "kick1_drift");
    if (rc != NBODY_OK) return rc;
    return launch_kick1_drift(c, dt);
}

extern "C" int nbody_force(nbody_ctx* c, double G, double softening)
{
    int rc = check_ready(c, "force");
    if (rc != NBODY_OK) return rc;
    if (softening < 0.0) return fail(c, NBODY_EINVAL, "force: softening=%g", softening);
    return launch_force(c, kForceOnly, 0.0, G, softening);
}

extern "C" int nbody_kick2(nbody_ctx* c, double dt, double heat_decay)
{
    int rc = check_ready(c, "kick2");
    if (rc != NBODY_OK) return rc;
    for (Shard& s : c->shards) {
        CU(c, cudaSetDevice(s.device));
        const unsigned blocks = unsigned((s.n_local + 255) / 256);
        if (c->prec == NBODY_F32)
            kick2_kernel<float><<<blocks, 256, 0, s.stream>>>((float*)s.vel, (const float*)s.acc, c->has_temp ? (float*)s.temp : nullptr,
                                                             s.n_local, 0.5f * float(dt), float(heat_decay));
        else
            kick2_kernel<double><<<blocks, 256, 0, s.stream>>>((double*)s.vel, (const double*)s.acc, c->has_temp ? (double*)s.temp : nullptr,
                                                              s.n_local, 0.5 * dt, heat_decay);
        c->launches++;
        CU(c, cudaGetLastError());
    }
    return NBODY_OK;
}

extern "C" int nbody_enable_timing(nbody_ctx* c, int enable)
{
    if (!c) return NBODY_EINVAL;
    int rc = fold_timing(c, c->shards[0]);
    c->timing = enable != 0;
    return rc;
}

extern "C" int nbody_read_timing(nbody_ctx* c, long long* launches, double* force_ms)
{
    if (!c) return NBODY_EINVAL;
    int rc = fold_timing(c, c->shards[0]);
    if (rc != NBODY_OK) return rc;
    if (launches) *launches = c->t_launches;
    if (force_ms) *force_ms = c->t_ms;
    c->t_launches = 0;
    c->t_ms = 0.0;
    return NBODY_OK;
}

// ============================================================================
// diagnostics
// ============================================================================

extern "C" int nbody_energy(nbody_ctx* c, double G, double softening, double* kinetic, double* potential, double momentum[3])
{
    int rc = check_ready(c, "energy");
    if (rc != NBODY_OK) return rc;
    double tot[5] = {0, 0, 0, 0, 0};
    std::vector<std::vector<double>> rows(c->shards.size());
    for (size_t k = 0; k < c->shards.size(); ++k) {
        Shard& s = c->shards[k];
        if (s.n_local_real == 0) continue;
        CU(c, cudaSetDevice(s.device));
        const unsigned blocks = unsigned((s.n_local_real + kEnergyThreads - 1) / kEnergyThreads);
        if (c->prec == NBODY_F32)
            energy_kernel<float><<<blocks, kEnergyThreads, 0, s.stream>>>((const float*)s.pos[c->cur], (const float*)s.vel, c->n, s.i_begin,
                                                                         s.n_local_real, softening * softening, s.energy_out);
        else
            energy_kernel<double><<<blocks, kEnergyThreads, 0, s.stream>>>((const double*)s.pos[c->cur], (const double*)s.vel, c->n,
                                                                          s.i_begin, s.n_local_real, softening * softening, s.energy_out);
        c->launches++;
        CU(c, cudaGetLastError());
        rows[k].resize(size_t(blocks) * 5);
        CU(c, cudaMemcpyAsync(rows[k].data(), s.energy_out, rows[k].size() * sizeof(double), cudaMemcpyDeviceToHost, s.stream));
    }
    for (size_t k = 0; k < c->shards.size(); ++k) {
        Shard& s = c->shards[k];
        CU(c, cudaSetDevice(s.device));
        CU(c, cudaStreamSynchronize(s.stream));
        for (size_t b = 0; b < rows[k].size() / 5; ++b)
            for (int q = 0; q < 5; ++q) tot[q] += rows[k][5 * b + q];
    }
    if (c->rank_mode && c->world > 1) {  // add the other processes' shares
        Shard& s = c->shards[0];
        CU(c, cudaMemcpyAsync(s.energy_out, tot, sizeof tot, cudaMemcpyHostToDevice, s.stream));
        NC(c, nccl().AllReduce(s.energy_out, s.energy_out, 5, ncclFloat64_, ncclSum_, s.comm, s.stream));
        CU(c, cudaMemcpyAsync(tot, s.energy_out, sizeof tot, cudaMemcpyDeviceToHost, s.stream));
        CU(c, cudaStreamSynchronize(s.stream));
    }
    if (kinetic) *kinetic = tot[0];
    if (potential) *potential = -G * tot[1];
    if (momentum) {
        momentum[0] = tot[2];
        momentum[1] = tot[3];
        momentum[2] = tot[4];
    }
    return NBODY_OK;
}

// ============================================================================
// collision broad phase (contacts.cuh)
// ============================================================================

extern "C" int nbody_set_radius(nbody_ctx* c, const void* radius)
{
    int rc = check_ready(c, "set_radius");
    if (rc != NBODY_OK) return rc;
    if (!radius) {
        c->has_radius = false;
        c->near_valid = false;
        return NBODY_OK;
    }
    std::vector<double> r(c->n);
    double rmax = 0.0;
    for (size_t i = 0; i < c->n; ++i) {
        r[i] = c->prec == NBODY_F32 ? double(static_cast<const float*>(radius)[i]) : static_cast<const double*>(radius)[i];
        if (!(r[i] >= 0.0)) return fail(c, NBODY_EINVAL, "set_radius: radius[%zu] = %g", i, r[i]);
        rmax = std::max(rmax, r[i]);
    }
    // A few dominant bodies (a galaxy's central mass, idx:750) would make every body a
    // candidate of the nearest-neighbour test; split them off when they are few.
    std::vector<double> sorted(r);
    std::nth_element(sorted.begin(), sorted.begin() + sorted.size() / 2, sorted.end());
    const double split = 4.0 * sorted[sorted.size() / 2];
    size_t n_big = 0;
    double small_max = 0.0;
    for (double v : r) {
        if (v > split) n_big++;
        else small_max = std::max(small_max, v);
    }
    const bool use_split = n_big > 0 && n_big <= std::max<size_t>(64, c->n / 256);
    c->r_split = use_split ? split : rmax;
    c->r_small_max = use_split ? small_max : rmax;
    for (Shard& s : c->shards) {
        CU(c, cudaSetDevice(s.device));
        CU(c, cudaMemsetAsync(s.radius, 0, c->n_pad * c->esz, s.stream));
        CU(c, cudaMemcpyAsync(s.radius, radius, c->n * c->esz, cudaMemcpyHostToDevice, s.stream));
        std::vector<unsigned> big;
        if (use_split)
            for (size_t i = s.i_begin; i < std::min(s.i_begin + s.n_local, c->n); ++i)
                if (r[i] > split) big.push_back(unsigned(i));
        cudaFree(s.big);
        s.big = nullptr;
        s.n_big = unsigned(big.size());
        if (s.n_big) {
            CU(c, cudaMalloc(&s.big, big.size() * sizeof(unsigned)));
            CU(c, cudaMemcpy(s.big, big.data(), big.size() * sizeof(unsigned), cudaMemcpyHostToDevice));
        }
    }
    for (Shard& s : c->shards) {
        CU(c, cudaSetDevice(s.device));
        CU(c, cudaStreamSynchronize(s.stream));
    }
    c->r_max = rmax;
    c->has_radius = true;
    c->near_valid = false;
    return NBODY_OK;
}

namespace {

template <typename Real, bool BOTH>
int launch_pairs(nbody_ctx* c, Shard& s, const unsigned* cand, unsigned n_cand, double margin)
{
    if (n_cand == 0) return NBODY_OK;
    const unsigned per_cta = kContactThreads * kContactPerThread;
    const unsigned gx = (n_cand + per_cta - 1) / per_cta;
    // enough CTAs to fill the chip: split the j range when there are few candidate blocks
    size_t jch = (size_t(c->sm_count) * 8 + gx - 1) / gx;
    const size_t max_jch = (c->n + kContactThreads - 1) / kContactThreads;
    jch = std::max<size_t>(1, std::min(jch, std::min<size_t>(max_jch, 65535)));
    const size_t chunk_len = round_up((c->n + jch - 1) / jch, size_t(kContactThreads));
    jch = (c->n + chunk_len - 1) / chunk_len;
    const dim3 grid(gx, unsigned(jch));
    contact_pairs_kernel<Real, BOTH><<<grid, kContactThreads, 0, s.stream>>>(static_cast<const Real*>(s.pos[c->cur]), static_cast<const Real*>(s.radius),
                                                                             cand, n_cand, s.i_begin, c->n, chunk_len, margin, s.pairs, s.pair_cap,
                                                                             s.pair_count);
    c->launches++;
    CU(c, cudaGetLastError());
    return NBODY_OK;
}

// Stage 1 on one shard: the nearest-neighbour values of the last force sweep -> candidate list
// (device, s.cand) and its length.
int shard_flag(nbody_ctx* c, Shard& s, double margin, unsigned* n_cand)
{
    *n_cand = 0;
    CU(c, cudaSetDevice(s.device));
    if (s.n_local_real == 0) return NBODY_OK;
    CU(c, cudaMemsetAsync(s.cand_count, 0, sizeof(unsigned), s.stream));
    const unsigned blocks = unsigned((s.n_local_real + 255) / 256);
    flag_near_kernel<<<blocks, 256, 0, s.stream>>>(s.near, static_cast<const float*>(s.radius), s.i_begin, s.n_local_real,
                                                   float(c->r_small_max + margin) * 1.000001f, float(c->r_split), float(c->near_eps2), s.cand,
                                                   s.cand_count);
    c->launches++;
    CU(c, cudaGetLastError());
    CU(c, cudaMemcpyAsync(n_cand, s.cand_count, sizeof *n_cand, cudaMemcpyDeviceToHost, s.stream));
    CU(c, cudaStreamSynchronize(s.stream));
    return NBODY_OK;
}

// Candidate x candidate (contacts.cuh): every device gets the candidates of all devices, concatenated in
// shard order (offsets[k] = where shard k's own begin), and dense (x, y, z, r) records of them.
int build_candidate_records(nbody_ctx* c, const std::vector<unsigned>& n_cand, std::vector<unsigned>& offsets, unsigned* total)
{
    offsets.assign(c->shards.size(), 0);
    unsigned F = 0;
    for (size_t k = 0; k < c->shards.size(); ++k) {
        offsets[k] = F;
        F += n_cand[k];
    }
    *total = F;
    std::vector<unsigned> all;
    if (c->shards.size() > 1) {
        all.resize(F);
        for (size_t k = 0; k < c->shards.size(); ++k) {
            Shard& s = c->shards[k];
            if (n_cand[k] == 0) continue;
            CU(c, cudaSetDevice(s.device));
            CU(c, cudaMemcpy(all.data() + offsets[k], s.cand, n_cand[k] * sizeof(unsigned), cudaMemcpyDeviceToHost));
        }
    }
    for (Shard& s : c->shards) {
        CU(c, cudaSetDevice(s.device));
        if (c->shards.size() > 1) {
            if (!s.cand_all) CU(c, cudaMalloc(&s.cand_all, c->cap_pad * sizeof(unsigned)));
            CU(c, cudaMemcpyAsync(s.cand_all, all.data(), F * sizeof(unsigned), cudaMemcpyHostToDevice, s.stream));
        }
        if (F > s.cand_rec_cap) {
            CU(c, cudaStreamSynchronize(s.stream));
            cudaFree(s.cand_rec);
            s.cand_rec = nullptr;
            s.cand_rec_cap = 0;
            const size_t want = std::min<size_t>(c->cap_pad, size_t(F) + F / 4 + 1024);
            CU(c, cudaMalloc(&s.cand_rec, want * sizeof(float4)));
            s.cand_rec_cap = want;
        }
        const unsigned* list = c->shards.size() > 1 ? s.cand_all : s.cand;
        gather_candidates_kernel<float><<<(F + 255) / 256, 256, 0, s.stream>>>(static_cast<const float*>(s.pos[c->cur]), static_cast<const float*>(s.radius),
                                                                               list, F, s.cand_rec);
        c->launches++;
        CU(c, cudaGetLastError());
    }
    if (c->shards.size() > 1)  // `all` is borrowed by the copies above
        for (Shard& s : c->shards) {
            CU(c, cudaSetDevice(s.device));
            CU(c, cudaStreamSynchronize(s.stream));
        }
    return NBODY_OK;
}

int launch_pairs_compact(nbody_ctx* c, Shard& s, unsigned i_off, unsigned n_i, unsigned n_all, double margin)
{
    if (n_i == 0) return NBODY_OK;
    const unsigned per_cta = kContactThreads * kContactPerThread;
    const unsigned gx = (n_i + per_cta - 1) / per_cta;
    size_t jch = (size_t(c->sm_count) * 8 + gx - 1) / gx;  // enough CTAs to fill the chip
    const size_t max_jch = (n_all + kContactThreads - 1) / kContactThreads;
    jch = std::max<size_t>(1, std::min(jch, std::min<size_t>(max_jch, 65535)));
    const unsigned chunk_len = unsigned(round_up((n_all + jch - 1) / jch, size_t(kContactThreads)));
    jch = (n_all + chunk_len - 1) / chunk_len;
    const unsigned* list = c->shards.size() > 1 ? s.cand_all : s.cand;
    contact_pairs_compact_kernel<float><<<dim3(gx, unsigned(jch)), kContactThreads, 0, s.stream>>>(s.cand_rec, list, i_off, n_i, n_all, chunk_len, margin,
                                                                                                    s.pairs, s.pair_cap, s.pair_count);
    c->launches++;
    CU(c, cudaGetLastError());
    return NBODY_OK;
}

// Stage 2 on one shard: `launch` queues its pair kernels; they run again with a larger buffer until every
// pair fits, then the pairs join `keys` as (i << 32 | j).
template <typename L> int shard_collect(nbody_ctx* c, Shard& s, L&& launch, std::vector<unsigned long long>& keys)
{
    CU(c, cudaSetDevice(s.device));
    if (s.pair_cap == 0) {
        s.pair_cap = 1 << 16;
        CU(c, cudaMalloc(&s.pairs, s.pair_cap * sizeof(uint2)));
    }
    for (;;) {
        CU(c, cudaMemsetAsync(s.pair_count, 0, sizeof(unsigned long long), s.stream));
        const int rc = launch();
        if (rc != NBODY_OK) return rc;
        unsigned long long found = 0;
        CU(c, cudaMemcpyAsync(&found, s.pair_count, sizeof found, cudaMemcpyDeviceToHost, s.stream));
        CU(c, cudaStreamSynchronize(s.stream));
        if (found > s.pair_cap) {  // grow and search again
            cudaFree(s.pairs);
            s.pairs = nullptr;
            s.pair_cap = size_t(found + found / 4 + 1024);
            CU(c, cudaMalloc(&s.pairs, s.pair_cap * sizeof(uint2)));
            continue;
        }
        if (found) {
            std::vector<uint2> host(found);
            CU(c, cudaMemcpyAsync(host.data(), s.pairs, found * sizeof(uint2), cudaMemcpyDeviceToHost, s.stream));
            CU(c, cudaStreamSynchronize(s.stream));
            for (const uint2& q : host) keys.push_back((unsigned long long)q.x << 32 | q.y);
        }
        return NBODY_OK;
    }
}

// Every shard's pairs: those whose lower index lies in the shard's own range, plus the lower-index
// partners of its big bodies.
int all_contacts(nbody_ctx* c, double margin, bool use_flags, std::vector<unsigned long long>& keys)
{
    const size_t ns = c->shards.size();
    std::vector<unsigned> n_cand(ns, 0), offsets(ns, 0);
    unsigned total = 0;
    bool compact = false;
    if (use_flags) {
        size_t flagged = 0;
        for (size_t k = 0; k < ns; ++k) {
            int rc = shard_flag(c, c->shards[k], margin, &n_cand[k]);
            if (rc != NBODY_OK) return rc;
            flagged += n_cand[k];
            c->cstat[0] += n_cand[k] + c->shards[k].n_big;
        }
        // F^2 tests among the candidates against F*N/2 of candidates x everybody (contacts.cuh)
        compact = flagged > 0 && 2 * flagged < c->n && env_int("NBODY_CONTACTS_COMPACT", 1) != 0;
        if (compact) {
            int rc = build_candidate_records(c, n_cand, offsets, &total);
            if (rc != NBODY_OK) return rc;
        }
    }
    if (compact) c->cstat[3] = 2;
    for (size_t k = 0; k < ns; ++k) {
        Shard& s = c->shards[k];
        if (s.n_local_real == 0) continue;
        const unsigned nc = use_flags ? n_cand[k] : unsigned(s.n_local_real);
        const unsigned n_big = use_flags ? s.n_big : 0;
        c->cstat[1] += nc + n_big;
        if (nc + n_big == 0) continue;
        int rc = shard_collect(c, s, [&]() -> int {
            int r;
            if (c->prec == NBODY_F32) {
                r = compact ? launch_pairs_compact(c, s, offsets[k], nc, total, margin)
                            : launch_pairs<float, false>(c, s, use_flags ? s.cand : nullptr, nc, margin);
                if (r == NBODY_OK) r = launch_pairs<float, true>(c, s, s.big, n_big, margin);
            } else {
                r = launch_pairs<double, false>(c, s, nullptr, nc, margin);  // fp64 contexts have no flags: every body is scanned
            }
            return r;
        }, keys);
        if (rc != NBODY_OK) return rc;
    }
    return NBODY_OK;
}

}  // namespace

extern "C" int nbody_contacts(nbody_ctx* c, double margin, unsigned* pairs, size_t cap, size_t* count)
{
    int rc = check_ready(c, "contacts");
    if (rc != NBODY_OK) return rc;
    if (!count) return fail(c, NBODY_EINVAL, "contacts: count is required");
    *count = 0;
    if (!c->has_radius) return fail(c, NBODY_EINVAL, "contacts before set_radius");
    if (!(margin >= 0.0)) return fail(c, NBODY_EINVAL, "contacts: margin=%g", margin);
    if (c->n > 0xffffffffull) return fail(c, NBODY_EINVAL, "contacts: n exceeds 32-bit pair indices");
    const bool use_flags = c->near_valid;
    c->cstat[0] = c->cstat[1] = c->cstat[2] = 0;
    c->cstat[3] = use_flags;
    std::vector<unsigned long long> keys;
    try {
        rc = all_contacts(c, margin, use_flags, keys);
        if (rc != NBODY_OK) return rc;
        std::sort(keys.begin(), keys.end());  // the reference's visiting order: ascending i, then j (idx:301-305)
        keys.erase(std::unique(keys.begin(), keys.end()), keys.end());  // a big body's scan repeats pairs its partner also found
    } catch (const std::bad_alloc&) {
        return fail(c, NBODY_ENOMEM, "contacts: host allocation failed");
    }
    c->cstat[2] = (long long)keys.size();
    *count = keys.size();
    if (keys.size() > cap) return fail(c, NBODY_ENOMEM, "contacts: %zu pairs found, room for %zu", keys.size(), cap);
    if (!pairs && !keys.empty()) return fail(c, NBODY_EINVAL, "contacts: pairs is NULL");
    for (size_t k = 0; k < keys.size(); ++k) {
        pairs[2 * k] = unsigned(keys[k] >> 32);
        pairs[2 * k + 1] = unsigned(keys[k] & 0xffffffffull);
    }
    return NBODY_OK;
}

extern "C" int nbody_contact_stats(nbody_ctx* c, long long out[4])
{
    if (!c || !out) return NBODY_EINVAL;
    for (int k = 0; k < 4; ++k) out[k] = c->cstat[k];
    return NBODY_OK;
}

// ============================================================================
// renderer feed: position history ring + ribbon trails (trails.cuh)
// ============================================================================

extern "C" int nbody_history_enable(nbody_ctx* c, int depth)
{
    if (!c) return NBODY_EINVAL;
    if (depth != 0 && (depth < 2 || depth > 256)) return fail(c, NBODY_EINVAL, "history_enable: depth=%d outside 2..256", depth);
    Shard& s = c->shards[0];
    CU(c, cudaSetDevice(s.device));
    CU(c, cudaStreamSynchronize(s.stream));
    cudaFree(s.hist);
    cudaFree(s.hist_alt);
    s.hist = s.hist_alt = nullptr;
    c->hist_depth = 0;
    c->hist_n = c->hist_prev_n = 0;
    c->hist_fresh_pending = false;
    c->hist_head = 0;
    if (depth == 0) return NBODY_OK;
    c->hist_stride = c->cap_pad;
    CU(c, cudaMalloc(&s.hist, size_t(depth) * c->hist_stride * 4 * c->esz));
    c->hist_depth = depth;
    const int smem = (depth + 1) * 512;  // trail_bodies() x sizeof(Real4) = 512 B per sample in either precision, one padding element per body
    if (c->prec == NBODY_F32) CU(c, cudaFuncSetAttribute(trail_ribbons_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    else CU(c, cudaFuncSetAttribute(trail_ribbons_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    return NBODY_OK;
}

extern "C" int nbody_history_push(nbody_ctx* c)
{
    int rc = check_ready(c, "history_push");
    if (rc != NBODY_OK) return rc;
    if (c->hist_depth == 0) return fail(c, NBODY_EINVAL, "history_push before history_enable");
    Shard& s = c->shards[0];
    CU(c, cudaSetDevice(s.device));
    const int fill = c->hist_n != c->n;
    const int* fresh = (!fill && c->hist_fresh_pending) ? s.hist_src : nullptr;  // bodies new to a remapped ring
    c->hist_fresh_pending = false;
    c->hist_prev_n = 0;
    if (fill) c->hist_head = 0;
    else c->hist_head = (c->hist_head + c->hist_depth - 1) % c->hist_depth;  // unshift: the newest sample moves one slot back
    const unsigned blocks = unsigned((c->n + 255) / 256);
    if (c->prec == NBODY_F32)
        history_push_kernel<float><<<blocks, 256, 0, s.stream>>>((const float*)s.pos[c->cur], (float4*)s.hist, c->n, c->hist_stride, c->hist_head,
                                                                 c->hist_depth, fill, fresh);
    else
        history_push_kernel<double><<<blocks, 256, 0, s.stream>>>((const double*)s.pos[c->cur], (double4*)s.hist, c->n, c->hist_stride, c->hist_head,
                                                                  c->hist_depth, fill, fresh);
    c->launches++;
    CU(c, cudaGetLastError());
    c->hist_n = c->n;
    return NBODY_OK;
}

extern "C" int nbody_history_remap(nbody_ctx* c, const int* source)
{
    int rc = check_ready(c, "history_remap");
    if (rc != NBODY_OK) return rc;
    if (!source) return fail(c, NBODY_EINVAL, "history_remap: source is NULL");
    const size_t old_n = c->hist_n ? c->hist_n : c->hist_prev_n;  // the list the ring was last pushed for
    if (c->hist_depth == 0 || old_n == 0) return fail(c, NBODY_EINVAL, "history_remap: no history to carry over (enable, then push)");
    for (size_t i = 0; i < c->n; ++i)
        if (source[i] >= 0 && size_t(source[i]) >= old_n)
            return fail(c, NBODY_EINVAL, "history_remap: source[%zu]=%d outside the previous list of %zu bodies", i, source[i], old_n);
    Shard& s = c->shards[0];
    CU(c, cudaSetDevice(s.device));
    const size_t ring_bytes = size_t(c->hist_depth) * c->hist_stride * 4 * c->esz;
    if (!s.hist_alt) CU(c, cudaMalloc(&s.hist_alt, ring_bytes));
    if (!s.hist_src) CU(c, cudaMalloc(&s.hist_src, c->cap_pad * sizeof(int)));
    CU(c, cudaMemcpyAsync(s.hist_src, source, c->n * sizeof(int), cudaMemcpyHostToDevice, s.stream));
    const unsigned blocks = unsigned((c->n + 255) / 256);
    if (c->prec == NBODY_F32)
        history_remap_kernel<float><<<blocks, 256, 0, s.stream>>>((const float4*)s.hist, (float4*)s.hist_alt, s.hist_src, c->n, c->hist_stride, c->hist_depth);
    else
        history_remap_kernel<double><<<blocks, 256, 0, s.stream>>>((const double4*)s.hist, (double4*)s.hist_alt, s.hist_src, c->n, c->hist_stride, c->hist_depth);
    c->launches++;
    CU(c, cudaGetLastError());
    CU(c, cudaStreamSynchronize(s.stream));  // `source` is borrowed for the call only
    std::swap(s.hist, s.hist_alt);
    c->hist_n = c->n;
    c->hist_prev_n = 0;
    c->hist_fresh_pending = true;
    return NBODY_OK;
}

namespace {
int history_ready(nbody_ctx* c, const char* what)
{
    int rc = check_ready(c, what);
    if (rc != NBODY_OK) return rc;
    if (c->hist_depth == 0 || c->hist_n != c->n || c->hist_fresh_pending)
        return fail(c, NBODY_EINVAL, "%s: no history for the current body list (enable, then push)", what);
    return NBODY_OK;
}
}  // namespace

extern "C" int nbody_history_download(nbody_ctx* c, void* out)
{
    int rc = history_ready(c, "history_download");
    if (rc != NBODY_OK) return rc;
    if (!out) return fail(c, NBODY_EINVAL, "history_download: out is NULL");
    Shard& s = c->shards[0];
    CU(c, cudaSetDevice(s.device));
    const size_t elems = c->n * size_t(c->hist_depth) * 3;
    const size_t bytes = elems * c->esz;
    if (bytes > s.ribbons_bytes) {
        CU(c, cudaStreamSynchronize(s.stream));
        cudaFree(s.ribbons);
        s.ribbons = nullptr;
        s.ribbons_bytes = 0;
        CU(c, cudaMalloc(&s.ribbons, bytes));
        s.ribbons_bytes = bytes;
    }
    const unsigned blocks = unsigned((c->n * size_t(c->hist_depth) + 255) / 256);
    if (c->prec == NBODY_F32)
        history_gather_kernel<float><<<blocks, 256, 0, s.stream>>>((const float4*)s.hist, (float*)s.ribbons, c->n, c->hist_stride, c->hist_head, c->hist_depth);
    else
        history_gather_kernel<double><<<blocks, 256, 0, s.stream>>>((const double4*)s.hist, (double*)s.ribbons, c->n, c->hist_stride, c->hist_head, c->hist_depth);
    c->launches++;
    CU(c, cudaGetLastError());
    CU(c, cudaMemcpyAsync(out, s.ribbons, bytes, cudaMemcpyDeviceToHost, s.stream));
    CU(c, cudaStreamSynchronize(s.stream));
    return NBODY_OK;
}

namespace {
// Builds the ribbons into Shard::ribbons (device); shared by the download call and the timing hook.
int build_ribbons(nbody_ctx* c, const double camera[3])
{
    Shard& s = c->shards[0];
    CU(c, cudaSetDevice(s.device));
    const size_t bytes = c->n * size_t(c->hist_depth) * 6 * sizeof(float);
    if (bytes > s.ribbons_bytes) {
        CU(c, cudaStreamSynchronize(s.stream));
        cudaFree(s.ribbons);
        s.ribbons = nullptr;
        s.ribbons_bytes = 0;
        CU(c, cudaMalloc(&s.ribbons, bytes));
        s.ribbons_bytes = bytes;
    }
    const size_t bpb = c->prec == NBODY_F32 ? trail_bodies<float>() : trail_bodies<double>();
    const unsigned blocks = unsigned((c->n + bpb - 1) / bpb);
    const size_t smem = size_t(c->hist_depth + 1) * 512;
    if (c->prec == NBODY_F32)
        trail_ribbons_kernel<float><<<blocks, kTrailThreads, smem, s.stream>>>((const float4*)s.hist, (const float*)s.radius, s.ribbons, c->n, c->hist_stride,
                                                                               c->hist_head, c->hist_depth, c->hist_depth, float(camera[0]),
                                                                               float(camera[1]), float(camera[2]));
    else
        trail_ribbons_kernel<double><<<blocks, kTrailThreads, smem, s.stream>>>((const double4*)s.hist, (const double*)s.radius, s.ribbons, c->n,
                                                                                c->hist_stride, c->hist_head, c->hist_depth, c->hist_depth, camera[0],
                                                                                camera[1], camera[2]);
    c->launches++;
    CU(c, cudaGetLastError());
    return NBODY_OK;
}
}  // namespace

extern "C" int nbody_trail_ribbons(nbody_ctx* c, const double camera[3], float* out)
{
    int rc = history_ready(c, "trail_ribbons");
    if (rc != NBODY_OK) return rc;
    if (!camera) return fail(c, NBODY_EINVAL, "trail_ribbons: camera is NULL");
    if (!c->has_radius) return fail(c, NBODY_EINVAL, "trail_ribbons before set_radius");
    rc = build_ribbons(c, camera);
    if (rc != NBODY_OK) return rc;
    Shard& s = c->shards[0];
    if (out) {  // out == NULL: build on the device only (timing)
        CU(c, cudaMemcpyAsync(out, s.ribbons, c->n * size_t(c->hist_depth) * 6 * sizeof(float), cudaMemcpyDeviceToHost, s.stream));
        CU(c, cudaStreamSynchronize(s.stream));
    }
    return NBODY_OK;
}

// ============================================================================
// snapshots
// ============================================================================

namespace {
struct SnapHeader {
    char magic[8];
    unsigned long long n, step;
    double dt, G, softening;
    unsigned prec, flags;
    unsigned long long reserved;
};
static_assert(sizeof(SnapHeader) == 64, "snapshot header is 64 bytes");
const char kSnapMagic[8] = {'N', 'B', 'O', 'D', 'Y', 'S', 'N', '1'};
}  // namespace

extern "C" int nbody_snapshot_save(nbody_ctx* c, const char* path, unsigned long long step, double dt, double G, double softening)
{
    int rc = check_ready(c, "snapshot_save");
    if (rc != NBODY_OK) return rc;
    if (!path) return fail(c, NBODY_EINVAL, "snapshot_save: path is NULL");
    if (c->rank_mode && c->world > 1) return fail(c, NBODY_EINVAL, "snapshot_save: a rank holds only its own vel/acc; gather on the host instead");
    const size_t n = c->n, esz = c->esz;
    std::vector<char> pm, vel, acc, temp, rad;
    try {
        pm.resize(4 * n * esz);
        vel.resize(4 * n * esz);
        acc.resize(4 * n * esz);
        if (c->has_temp) temp.resize(n * esz);
        if (c->has_radius) rad.resize(n * esz);
    } catch (const std::bad_alloc&) {
        return fail(c, NBODY_ENOMEM, "snapshot_save: host allocation failed");
    }
    rc = nbody_download(c, pm.data(), vel.data(), acc.data(), c->has_temp ? temp.data() : nullptr);
    if (rc != NBODY_OK) return rc;
    if (c->has_radius) {
        Shard& s = c->shards[0];
        CU(c, cudaSetDevice(s.device));
        CU(c, cudaMemcpy(rad.data(), s.radius, n * esz, cudaMemcpyDeviceToHost));
    }
    SnapHeader h;
    memset(&h, 0, sizeof h);
    memcpy(h.magic, kSnapMagic, 8);
    h.n = n;
    h.step = step;
    h.dt = dt;
    h.G = G;
    h.softening = softening;
    h.prec = unsigned(c->prec);
    h.flags = (c->has_temp ? 1u : 0u) | (c->has_radius ? 2u : 0u);
    FILE* f = fopen(path, "wb");
    if (!f) return fail(c, NBODY_EIO, "snapshot_save: cannot open %s", path);
    bool ok = fwrite(&h, sizeof h, 1, f) == 1;
    for (const std::vector<char>* a : {&pm, &vel, &acc, &temp, &rad})
        if (ok && !a->empty()) ok = fwrite(a->data(), 1, a->size(), f) == a->size();
    ok = (fclose(f) == 0) && ok;
    return ok ? NBODY_OK : fail(c, NBODY_EIO, "snapshot_save: short write to %s", path);
}

extern "C" int nbody_snapshot_load(nbody_ctx* c, const char* path, unsigned long long* step, double* dt, double* G, double* softening)
{
    if (!c) return NBODY_EINVAL;
    if (!path) return fail(c, NBODY_EINVAL, "snapshot_load: path is NULL");
    FILE* f = fopen(path, "rb");
    if (!f) return fail(c, NBODY_EIO, "snapshot_load: cannot open %s", path);
    SnapHeader h;
    if (fread(&h, sizeof h, 1, f) != 1 || memcmp(h.magic, kSnapMagic, 8) != 0) {
        fclose(f);
        return fail(c, NBODY_EIO, "snapshot_load: %s is not an NBODYSN1 snapshot", path);
    }
    if (h.prec != unsigned(c->prec) || h.n == 0 || h.n > c->capacity || (h.flags & ~3u)) {
        fclose(f);
        return fail(c, NBODY_EINVAL, "snapshot_load: n=%llu precision=%u flags=%u do not fit this context (capacity %zu, precision %d)", h.n, h.prec,
                    h.flags, c->capacity, int(c->prec));
    }
    const size_t n = size_t(h.n), esz = c->esz;
    std::vector<char> pm, vel, acc, temp, rad;
    try {
        pm.resize(4 * n * esz);
        vel.resize(4 * n * esz);
        acc.resize(4 * n * esz);
        if (h.flags & 1u) temp.resize(n * esz);
        if (h.flags & 2u) rad.resize(n * esz);
    } catch (const std::bad_alloc&) {
        fclose(f);
        return fail(c, NBODY_ENOMEM, "snapshot_load: host allocation failed");
    }
    bool ok = true;
    for (std::vector<char>* a : {&pm, &vel, &acc, &temp, &rad})
        if (ok && !a->empty()) ok = fread(a->data(), 1, a->size(), f) == a->size();
    char extra;
    ok = ok && fread(&extra, 1, 1, f) == 0;  // nothing may follow
    fclose(f);
    if (!ok) return fail(c, NBODY_EIO, "snapshot_load: %s is truncated or has trailing bytes", path);
    int rc = nbody_upload(c, n, pm.data(), vel.data(), acc.data(), (h.flags & 1u) ? temp.data() : nullptr);
    if (rc != NBODY_OK) return rc;
    if (h.flags & 2u) {
        rc = nbody_set_radius(c, rad.data());
        if (rc != NBODY_OK) return rc;
    }
    if (step) *step = h.step;
    if (dt) *dt = h.dt;
    if (G) *G = h.G;
    if (softening) *softening = h.softening;
    return NBODY_OK;
}

// ============================================================================
// headless body renderer (render.cu)
// ============================================================================

extern "C" int nbody_set_materials(nbody_ctx* c, const int* type, const float* seed)
{
    int rc = check_ready(c, "set_materials");
    if (rc != NBODY_OK) return rc;
    Shard& s = c->shards[0];
    CU(c, cudaSetDevice(s.device));
    if (type)
        for (size_t i = 0; i < c->n; ++i)
            if (type[i] < 0 || type[i] > 2) return fail(c, NBODY_EINVAL, "set_materials: type[%zu] = %d (0 rock, 1 metal, 2 ice)", i, type[i]);
    if (!s.mat_type) {
        CU(c, cudaMalloc(&s.mat_type, c->cap_pad * sizeof(int)));
        CU(c, cudaMalloc(&s.mat_seed, c->cap_pad * sizeof(float)));
    }
    CU(c, cudaMemsetAsync(s.mat_type, 0, c->n * sizeof(int), s.stream));
    CU(c, cudaMemsetAsync(s.mat_seed, 0, c->n * sizeof(float), s.stream));
    if (type) CU(c, cudaMemcpyAsync(s.mat_type, type, c->n * sizeof(int), cudaMemcpyHostToDevice, s.stream));
    if (seed) CU(c, cudaMemcpyAsync(s.mat_seed, seed, c->n * sizeof(float), cudaMemcpyHostToDevice, s.stream));
    CU(c, cudaStreamSynchronize(s.stream));
    c->has_materials = true;
    return NBODY_OK;
}

extern "C" int nbody_render(nbody_ctx* c, int width, int height, const double cam_position[3], const double cam_target[3],
                            double fov_deg, float* rgba, int* ids)
{
    int rc = check_ready(c, "render");
    if (rc != NBODY_OK) return rc;
    if (width < 1 || height < 1 || size_t(width) * size_t(height) > (size_t(1) << 28))
        return fail(c, NBODY_EINVAL, "render: %d x %d", width, height);
    if (!cam_position || !cam_target || !(fov_deg > 0.0 && fov_deg < 180.0)) return fail(c, NBODY_EINVAL, "render: camera");
    if (!c->has_radius) return fail(c, NBODY_EINVAL, "render before set_radius");
    if (c->rank_mode && c->world > 1) return fail(c, NBODY_EINVAL, "render: a rank holds only its own temperatures; render from a single-process context");
    if (c->n > 0xffffffffull) return fail(c, NBODY_EINVAL, "render: n exceeds the 32-bit body id of the depth buffer");
    // Matrix4.lookAt(position, target, up = +y) in double, stored as float (render.cuh)
    RenderCamera cam;
    double z[3] = {cam_position[0] - cam_target[0], cam_position[1] - cam_target[1], cam_position[2] - cam_target[2]};
    double l = sqrt(z[0] * z[0] + z[1] * z[1] + z[2] * z[2]);
    if (!(l > 0.0)) return fail(c, NBODY_EINVAL, "render: camera position equals its target");
    for (double& v : z) v /= l;
    double x[3] = {1.0 * z[2] - 0.0 * z[1], 0.0 * z[0] - 0.0 * z[2], 0.0 * z[1] - 1.0 * z[0]};
    l = sqrt(x[0] * x[0] + x[1] * x[1] + x[2] * x[2]);
    if (!(l > 0.0)) return fail(c, NBODY_EINVAL, "render: the camera looks along its up axis");
    for (double& v : x) v /= l;
    const double y[3] = {z[1] * x[2] - z[2] * x[1], z[2] * x[0] - z[0] * x[2], z[0] * x[1] - z[1] * x[0]};
    for (int k = 0; k < 3; ++k) {
        cam.pos[k] = float(cam_position[k]);
        cam.xa[k] = float(x[k]);
        cam.ya[k] = float(y[k]);
        cam.za[k] = float(z[k]);
    }
    cam.tan_half = float(tan(fov_deg * 3.14159265358979323846 / 360.0));
    cam.aspect = float(width) / float(height);
    cam.near = 0.1f;  // PerspectiveCamera(45, aspect, 0.1, 3000), idx:704

    Shard& s = c->shards[0];
    CU(c, cudaSetDevice(s.device));
    const size_t pixels = size_t(width) * size_t(height);
    if (pixels > s.image_pixels) {
        CU(c, cudaStreamSynchronize(s.stream));
        render_scratch_free(s.render_scratch);
        cudaFree(s.image);
        cudaFree(s.image_ids);
        s.image = nullptr;
        s.image_ids = nullptr;
        s.image_pixels = 0;
        CU(c, cudaMalloc(&s.image, pixels * sizeof(float4)));
        CU(c, cudaMalloc(&s.image_ids, pixels * sizeof(int)));
        s.image_pixels = pixels;
    }
    // temperatures: the renderer's `temperature` uniform (idx:553); sharded state is gathered on this device
    const void* temp = c->has_temp ? s.temp : nullptr;
    if (c->has_temp && c->shards.size() > 1) {
        if (!s.temp_full) CU(c, cudaMalloc(&s.temp_full, c->cap_pad * c->esz));
        for (Shard& o : c->shards) {
            if (o.n_local_real == 0) continue;
            CU(c, cudaSetDevice(o.device));
            CU(c, cudaStreamSynchronize(o.stream));
            CU(c, cudaMemcpy(static_cast<char*>(s.temp_full) + std::min(o.i_begin, c->n) * c->esz, o.temp, o.n_local_real * c->esz, cudaMemcpyDefault));
        }
        CU(c, cudaSetDevice(s.device));
        temp = s.temp_full;
    }
    const int* type = c->has_materials ? s.mat_type : nullptr;
    const float* seed = c->has_materials ? s.mat_seed : nullptr;
    cudaError_t e;
    if (c->prec == NBODY_F32)
        e = render_launch<float>((const float*)s.pos[c->cur], (const float*)s.radius, (const float*)temp, type, seed, c->n, cam, width, height, s.render_scratch,
                                 s.image, ids ? s.image_ids : nullptr, s.stream);
    else
        e = render_launch<double>((const double*)s.pos[c->cur], (const double*)s.radius, (const double*)temp, type, seed, c->n, cam, width, height,
                                  s.render_scratch, s.image, ids ? s.image_ids : nullptr, s.stream);
    c->launches += 4;  // bin count, scan, bin fill, tile render
    if (e != cudaSuccess) return fail(c, NBODY_ECUDA, "render: %s", cudaGetErrorString(e));
    if (rgba) CU(c, cudaMemcpyAsync(rgba, s.image, pixels * sizeof(float4), cudaMemcpyDeviceToHost, s.stream));
    if (ids) CU(c, cudaMemcpyAsync(ids, s.image_ids, pixels * sizeof(int), cudaMemcpyDeviceToHost, s.stream));
    if (rgba || ids) CU(c, cudaStreamSynchronize(s.stream));  // rgba == NULL: the frame stays on the device (timing)
    return NBODY_OK;
}
