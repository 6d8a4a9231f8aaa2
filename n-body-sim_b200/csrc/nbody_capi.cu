// libnbody.so -- implementation of include/nbody.h.
// One nbody_ctx owns one Shard per device it drives.  In "rank mode"
// (nbody_create_rank) the process holds a single Shard that is rank r of a
// world of processes; otherwise it holds n_gpus Shards (ranks 0..n_gpus-1 of a
// world inside this process).  Both run the same per-step schedule; only the
// NCCL communicator construction differs.
#include "capi_ctx.cuh"

#include "exchange.cuh"
#include "force_f32.cuh"
#include "force_f32_sym.cuh"
#include "force_f32_symw.cuh"
#include "sym_coop.cuh"
#include "force_f64.cuh"
#include "force_f64_sym.cuh"
#include "integrate.cuh"
#include "small.cuh"
#include "medium.cuh"


#ifndef NB_UNIFORM_UNROLL
#define NB_UNIFORM_UNROLL 8
#endif
constexpr int kUniUnroll = NB_UNIFORM_UNROLL;  // inner-loop unroll of the equal-mass force kernel
// i-bodies per thread of the pairs-once fp32 pass: 8 (168 registers, 2 CTAs/SM) beats 4 (128 registers, 3 CTAs/SM) for both
// mass variants once each runs at its best unroll (force_f32_sym.cuh; N=1M: 4,240 against 3,920 and 3,608 against 3,526).
constexpr int kSym32EqualMassR = 8, kSym32GeneralR = 8;
constexpr size_t kSym32SmallN = 49152;              // below: 4 bodies per thread
constexpr int kSym32DefaultMinN = 32768;            // smallest system that takes the CTA-wide form (below: too few units for the SMs)
constexpr int kSymwDefaultMinN = 5121;              // the warp-unit form (force_f32_symw.cuh) takes [min, max) bodies whatever the masses,
constexpr int kSymwDefaultMaxN = 393216;            // and general masses at every size from min up
constexpr int kSymCoopDefaultMaxN = 14336;          // sym_coop.cuh: all steps of a call in one cooperative launch up to here
constexpr size_t kSymwTinyN = 7168;                 // below: 2 bodies per thread (N=6,144: 24.3 us/step against 28.0 with 4; 8,192: 35.4 against 32.6)
constexpr size_t kSymwSmallN = 28672;               // below: 4 bodies per thread, from here 8

namespace {

int check_exchange(nbody_ctx* c);  // defined with the exchange code below
int resolve_mass_detection(nbody_ctx* c);  // defined with the upload code below

// fp64 force kernel: two i-bodies per thread pay from ~32k bodies up (N=65,536: 961 -> 974 G interactions/s with
// equal masses, 906 -> 948 without) and cost below (N=16,384: 304 -> 326 us/step): profiles/r1e_f64_r_uniform.jsonl
constexpr size_t kF64TwoBodyMinN = 32768;
// fp32: 128-body j tiles up to this many bodies.  Measured us/step with 512- / 128-body tiles (scripts/gpu_25.sh,
// profiles/r1f_f32_tile_sweep.jsonl): N=600 24.5 / 12.3, 1,024 24.5 / 13.1, 2,048 24.6 / 14.3, 4,096 24.6 / 20.4,
// 8,192 41.0 / 42.9, 16,384 116 / 128.  (Two i-bodies per thread change nothing there: the 512 x 512 CTA is a
// dependent chain, not an issue-bound one; profiles/r1f_f32_midsize_r.jsonl.)
constexpr size_t kF32SmallTileMaxN = 4096;
// The pairs-once pass (force_f64_sym.cuh) serves one-GPU MATH 2 contexts from 3,072 bodies up, while its partial
// sums fit NBODY_F64_SYM_MAX_MB (default 4 GiB: N ~ 300k).  Below, its two launches and few CTAs lose to the
// one-sided kernel (us/step one-sided vs pairs-once: N=2,048 16.8 / 20.5, N=4,096 38.1 / 30.9, N=16,384 291 / 190;
// profiles/r1f_f64_sym_r.jsonl).
constexpr size_t kSymMinN = 3072;
size_t sym_partial_bytes(size_t n_pad, int r, int k, size_t* bytes_i, size_t* bytes_j)
{
    const size_t nib = n_pad / (size_t(kSymThreads) * r);
    const size_t bi = size_t(sym_chunks_of_row(int(nib), r, k, 0)) * 3 * n_pad * sizeof(double);
    const size_t bj = nib * 3 * n_pad * sizeof(double);
    if (bytes_i) *bytes_i = bi;
    if (bytes_j) *bytes_j = bj;
    return bi + bj;
}
// j-blocks per CTA: as many as still leave ~48 CTAs per SM (NBODY_F64_SYM_K overrides); 0 = not this pass.
// Measured (profiles/r1f_f64_sym.jsonl): N=65,536 runs at 1,599 / 1,617 / 1,622 / 1,504 G interactions/s with
// k = 1 / 2 / 4 / 19, N=16,384 at 192 / 204 / 262 us per step with k = 1 / 2 / 4.
long long sym_cta_count(size_t n_pad, int r, int k)
{
    const int nib = int(n_pad / (size_t(kSymThreads) * r));
    long long ctas = 0;
    for (int I = 0; I < nib; ++I) ctas += sym_chunks_of_row(nib, r, k, I);
    return ctas;
}
// Shape of the pass for n bodies: i-bodies per thread (*r_out) and j-blocks per CTA (returned; 0 = not this pass).
// Small systems take fewer bodies per thread so that every SM still gets several CTAs (NBODY_F64_SYM_R overrides;
// measured best: R = 1 at N=4,096, 2 at 8,192, 4 from 12,288 up).
int sym_pick_k(const nbody_ctx* c, size_t n, int* r_out)
{
    *r_out = kSymMaxR;
    if (!c->f64_sym_allowed || n < kSymMinN) return 0;
    const size_t n_pad = round_up(n, size_t(kSymIBlock));
    const long long sms = std::max(c->sm_count, 1);
    int r = env_int("NBODY_F64_SYM_R", 0);
    if (r != 1 && r != 2 && r != 4) {
        r = 4;
        while (r > 1 && sym_cta_count(n_pad, r, 1) < 6 * sms) r /= 2;
    }
    *r_out = r;
    const int forced = env_int("NBODY_F64_SYM_K", 0);
    int k = 1;
    if (forced > 0) {
        k = forced;
    } else {
        for (int t = 2; t <= 64; ++t) {
            if (sym_cta_count(n_pad, r, t) < 48 * sms) break;
            k = t;
        }
    }
    return k;  // (the array form of the pass, softening == 0 on one GPU, checks its memory budget at launch)
}

int f64_bodies_per_thread(const nbody_ctx* c, size_t n)
{
    if (c->f64_math != 2) return 1;
    if (c->f64_sym_allowed && n >= kSymMinN) return kSymIBlock / kThreads;  // its 512-body i-blocks set the padding
    if (c->f64_R_forced == 1 || c->f64_R_forced == 2) return c->f64_R_forced;
    return n >= kF64TwoBodyMinN ? 2 : 1;
}

size_t pad_unit(const nbody_ctx* c) { return pad_unit_of(c); }

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
    // everything a peer maps lives in ONE allocation (one CUDA IPC handle in rank mode): the two position buffers, the
    // fixed-point accumulators of the fp32 pairs-once pass (force_f32_sym.cuh), the exchange flags and that pass's control words
    {
        const size_t pos_bytes = round_up(c->cap_pad * v4, 256);
        // fp32: [3][cap] sums + poison; fp64: hi [3][cap] + poison, then lo [3][cap] (force_f64_sym.cuh, FIXED)
        const size_t acc_bytes = round_up(((c->prec == NBODY_F32 ? 3 : 6) * c->cap_pad + 2) * sizeof(unsigned long long), 256);
        const size_t flag_bytes = 256;  // kMaxPeers epoch flags, then 2 control words
        s.arena_bytes = 2 * pos_bytes + acc_bytes + flag_bytes;
        CU(c, cudaMalloc(&s.arena, s.arena_bytes));
        CU(c, cudaMemset(s.arena, 0, s.arena_bytes));
        s.off_pos1 = pos_bytes;
        s.off_acc = 2 * pos_bytes;
        s.off_flags = 2 * pos_bytes + acc_bytes;
        char* base = static_cast<char*>(s.arena);
        s.pos[0] = base;
        s.pos[1] = base + s.off_pos1;
        s.accfix = acc_bytes ? reinterpret_cast<unsigned long long*>(base + s.off_acc) : nullptr;
        s.flags = reinterpret_cast<unsigned long long*>(base + s.off_flags);
        s.sym_ctl = reinterpret_cast<unsigned*>(s.flags + kMaxPeers);
    }
    CU(c, cudaMalloc(&s.stage, c->cap_pad * v4));
    CU(c, cudaMalloc(&s.vel, local_cap * v4));
    CU(c, cudaMalloc(&s.acc, local_cap * v4));
    CU(c, cudaMalloc(&s.temp, local_cap * c->esz));
    // (an fp64 upload re-picks its i-tile per body count: size for the smallest one, kThreads)
    const size_t ntiles = local_cap / size_t(c->prec == NBODY_F64 ? kThreads : c->i_tile) + 1;
    CU(c, cudaMalloc(&s.counters, ntiles * sizeof(unsigned)));
    CU(c, cudaMemset(s.counters, 0, ntiles * sizeof(unsigned)));
    CU(c, cudaMalloc(&s.medium_arrivals, sizeof(unsigned)));
    CU(c, cudaMalloc(&s.xerr, sizeof(unsigned)));
    CU(c, cudaMemset(s.xerr, 0, sizeof(unsigned)));
    s.energy_rows = (local_cap + kEnergyThreads - 1) / kEnergyThreads + 1;
    CU(c, cudaMalloc(&s.energy_out, s.energy_rows * 5 * sizeof(double)));
    CU(c, cudaMalloc(&s.radius, c->cap_pad * c->esz));
    CU(c, cudaMalloc(&s.mass_sums, (c->cap_pad / 256 + 1) * sizeof(double)));
    CU(c, cudaMalloc(&s.outliers, kMaxOutliers * sizeof(MassOutlier)));
    CU(c, cudaMalloc(&s.outlier_count, sizeof(unsigned)));
    CU(c, cudaMalloc(&s.near, local_cap * sizeof(float)));
    CU(c, cudaMalloc(&s.cand, local_cap * sizeof(unsigned)));
    CU(c, cudaMalloc(&s.cand_count, sizeof(unsigned)));
    CU(c, cudaMalloc(&s.pair_count, sizeof(unsigned long long)));
    if (c->world == 1 && env_int("NBODY_SMALL_IO", 1) != 0) {  // one device: small systems travel as one record (integrate.cuh)
        const size_t bytes = 13 * std::min(c->cap_pad, kSmallIoMaxN) * c->esz + 16;
        CU(c, cudaMalloc(&s.io_dev, bytes));
        CU(c, cudaHostAlloc(&s.io_host, bytes, cudaHostAllocMapped | cudaHostAllocPortable));
        CU(c, cudaEventCreateWithFlags(&s.io_done, cudaEventDisableTiming));
    }
    return NBODY_OK;
}

// The fp32 force kernel's variants (force_f32.cuh): nearest-neighbour tracking, equal-mass sweep, softening == 0, and the
// ring-stage size `tp` (j-pairs: 512- or 128-body tiles).  Two i-bodies per thread (NBODY_R=2) exist for the plain general
// sweep on 512-body tiles only; every other request with R = 2 gets the R = 4 kernel's nearest relative and must not
// happen (create/upload keep R = 4 whenever tracking, equal masses or small tiles are in play).
template <typename F> void pick_f32_kernel(bool near, bool uniform, bool z, int R, int tp, F&& f)
{
    auto by_tile = [&](auto nr, auto un, auto zs) {
        constexpr bool N = decltype(nr)::value, U = decltype(un)::value, Z = decltype(zs)::value;
        if (tp == kF32SmallTilePairs) {
            if (N) return f(force_f32_near_kernel<4, Z, U, NB_NEAR_UNROLL, kF32SmallTilePairs>);
            return f(force_f32_kernel<4, Z, U ? kUniUnroll : kF32Unroll, U, kF32SmallTilePairs>);
        }
        if (N) return f(force_f32_near_kernel<4, Z, U>);
        if (!U && R == 2) return f(force_f32_kernel<2, Z>);
        return f(force_f32_kernel<4, Z, U ? kUniUnroll : kF32Unroll, U>);
    };
    auto by_z = [&](auto nr, auto un) { return z ? by_tile(nr, un, std::true_type{}) : by_tile(nr, un, std::false_type{}); };
    auto by_u = [&](auto nr) { return uniform ? by_z(nr, std::true_type{}) : by_z(nr, std::false_type{}); };
    return near ? by_u(std::true_type{}) : by_u(std::false_type{});
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

template <typename F> void pick_f64_sym_kernel(bool zero_soft, bool uniform, int r, F&& f)
{
    auto by_r = [&](auto zs, auto un) {
        constexpr bool Z = decltype(zs)::value, U = decltype(un)::value;
        if (r == 1) return f(force_f64_sym_kernel<Z, U, 1>);
        if (r == 2) return f(force_f64_sym_kernel<Z, U, 2>);
        return f(force_f64_sym_kernel<Z, U, 4>);
    };
    if (zero_soft) return uniform ? by_r(std::true_type{}, std::true_type{}) : by_r(std::true_type{}, std::false_type{});
    return uniform ? by_r(std::false_type{}, std::true_type{}) : by_r(std::false_type{}, std::false_type{});
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

template <typename F> void pick_f32_sym_kernel(int R, bool uniform, F&& f)
{
    if (R == 8) return uniform ? f(force_f32_sym_kernel<8, true>) : f(force_f32_sym_kernel<8, false>);
    return uniform ? f(force_f32_sym_kernel<4, true>) : f(force_f32_sym_kernel<4, false>);
}

template <typename F> void pick_f32_symw_kernel(int R, bool uniform, F&& f)
{
    if (R == 8) return uniform ? f(force_f32_symw_kernel<8, true>) : f(force_f32_symw_kernel<8, false>);
    if (R == 2) return uniform ? f(force_f32_symw_kernel<2, true>) : f(force_f32_symw_kernel<2, false>);
    return uniform ? f(force_f32_symw_kernel<4, true>) : f(force_f32_symw_kernel<4, false>);
}

int set_kernel_attrs(nbody_ctx* c)
{
    const int sm32 = int(force_f32_smem_bytes());
    const int sm64 = int(force_f64_smem_bytes());
    for (int near = 0; near < 2; ++near)
        for (int uni = 0; uni < 2; ++uni)
            for (int z = 0; z < 2; ++z)
                for (int R : {2, 4}) {  // (the 2 KiB-tile variants stay below the 48 KiB default)
                    cudaError_t e = cudaSuccess;
                    pick_f32_kernel(near != 0, uni != 0, z != 0, R, kF32TilePairs,
                                    [&](auto kernel) { e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, sm32); });
                    if (e != cudaSuccess) return fail(c, NBODY_ECUDA, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
                }
    for (int R : {4, 8})
        for (int uni = 0; uni < 2; ++uni) {
            cudaError_t e = cudaSuccess;
            pick_f32_sym_kernel(R, uni != 0, [&](auto kernel) {
                e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(force_f32_sym_smem_bytes()));
            });
            if (e != cudaSuccess) return fail(c, NBODY_ECUDA, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
        }
    CU(c, cudaFuncSetAttribute(small_step_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(small_smem_bytes<float>())));
    CU(c, cudaFuncSetAttribute(small_step_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(small_smem_bytes<double>())));
    CU(c, cudaFuncSetAttribute(medium_step_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(medium_smem_bytes<float>())));
    CU(c, cudaFuncSetAttribute(medium_step_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(medium_smem_bytes<double>())));
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
        // pairs-once pass (force_f32_sym.cuh): the default for systems of sym32_min_n bodies and more
        c->sym32_allowed = env_int("NBODY_F32_SYM", 1) != 0 && c->R == 4;
        c->sym32_R_forced = env_int("NBODY_F32_SYM_R", 0);  // 4 or 8; otherwise by kernel variant (build_sym32_table)
        c->sym32_min_n = size_t(std::max(env_int("NBODY_F32_SYM_MIN", kSym32DefaultMinN), 1));
        // ... and its warp-unit form (force_f32_symw.cuh)
        c->symw_min_n = size_t(std::max(env_int("NBODY_F32_SYMW_MIN", kSymwDefaultMinN), 1));
        c->symw_max_n = env_int("NBODY_F32_SYMW", 1) != 0 ? size_t(std::max(env_int("NBODY_F32_SYMW_MAX", kSymwDefaultMaxN), 0)) : 0;
        c->symw_R_forced = env_int("NBODY_F32_SYMW_R", 0);
        c->symw_general_large = env_int("NBODY_F32_SYMW_GENERAL", 1) != 0;
        // one cooperative launch per call up to 14,336 bodies (us per step, launch per step -> cooperative: N=6,144 25.5 -> 21.1, 8,192
        // 33.5 -> 28.6, 12,288 57.1 -> 54.4, 16,384 91.3 -> 89.0 with equal and 95.5 -> 96.1 with general masses, 32,768 312.7 -> 319.8:
        // profiles/r2v_fine_lab.jsonl); the kernel itself takes up to kSymCoopMaxN
        c->sym_coop_max_n = size_t(std::min<long long>(std::max(env_int("NBODY_SYM_COOP_MAX", kSymCoopDefaultMaxN), 0), (long long)kSymCoopMaxN));
    } else {
        c->f64_math = env_int("NBODY_F64_MATH", 2);
        c->f64_R_forced = env_int("NBODY_F64_R", 0);
        // (forcing the one-sided kernel's shape, or another arithmetic, selects the one-sided kernel)
        c->f64_sym_allowed = env_int("NBODY_F64_SYM", 1) != 0 && c->f64_math == 2 && c->f64_R_forced == 0;
        c->R = f64_bodies_per_thread(c, capacity);  // for cap_pad: the largest i-tile an upload may pick
        c->i_tile = kThreads * c->R;
        c->j_unit = kJUnitF64;
    }
    // one cluster up to here, all SMs above (measured crossover, profiles/r1f_small_vs_medium_crossover.jsonl: fp64 2.33 / 3.06 us
    // per step at 151 bodies and 4.26 / 3.49 at 300; fp32 2.17 / 2.57 at 300 and 3.64 / 2.66 at 512)
    const int small_default = prec == NBODY_F64 ? kSmallDefaultMaxN64 : kSmallDefaultMaxN32;
    c->small_max_n = size_t(std::min(std::max(env_int("NBODY_SMALL_MAX", small_default), 0), kSmallMaxN));
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
    DeviceGuard device_guard_;
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
    DeviceGuard device_guard_;
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

void world_free(nbody_ctx* c);  // page_world.cu

extern "C" void nbody_destroy(nbody_ctx* c)
{
    DeviceGuard device_guard_;
    if (!c) return;
    if (!c->shards.empty()) {
        cudaSetDevice(c->shards[0].device);
        world_free(c);
    }
    for (Shard& s : c->shards) {
        cudaSetDevice(s.device);
        if (s.own_stream) cudaStreamSynchronize(s.own_stream);
        if (s.comm && nccl().ok()) nccl().CommDestroy(s.comm);
        for (cudaEvent_t e : s.ev) cudaEventDestroy(e);
        for (void* m : s.ipc_opened) cudaIpcCloseMemHandle(m);
        cudaFree(s.arena);  // pos[0], pos[1], accfix, flags, sym_ctl
        cudaFree(s.xerr);
        cudaFree(s.medium_arrivals);
        cudaFree(s.stage);
        cudaFree(s.io_dev);
        if (s.io_host) cudaFreeHost(s.io_host);
        if (s.io_done) cudaEventDestroy(s.io_done);
        cudaFree(s.vel);
        cudaFree(s.acc);
        cudaFree(s.temp);
        cudaFree(s.partial);
        cudaFree(s.sym_part_i);
        cudaFree(s.sym_part_j);
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
        cudaFree(s.sym_units);
        cudaFree(s.mass_sums);
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
                a.peer_acc[b.rank] = b.accfix;
                a.peer_ctl[b.rank] = b.sym_ctl;
            }
        }
        c->p2p = true;
        return NBODY_OK;
    }
    struct Blob {
        cudaIpcMemHandle_t arena;
        int ok;
    };
    Shard& s = c->shards[0];
    CU(c, cudaSetDevice(s.device));
    std::vector<Blob> all(c->world);
    Blob mine;
    memset(&mine, 0, sizeof mine);
    mine.ok = cudaIpcGetMemHandle(&mine.arena, s.arena) == cudaSuccess;
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
            char* base = static_cast<char*>(s.arena);
            if (r != c->my_rank) {
                void* ptr = nullptr;
                if (cudaIpcOpenMemHandle(&ptr, all[r].arena, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
                    cudaGetLastError();
                    ok = false;
                    break;
                }
                s.ipc_opened.push_back(ptr);
                base = static_cast<char*>(ptr);
            }
            // (every rank's arena has the same layout: same precision and capacity)
            s.peer_pos[r][0] = base;
            s.peer_pos[r][1] = base + s.off_pos1;
            s.peer_acc[r] = s.accfix ? base + s.off_acc : nullptr;
            s.peer_flags.flags[r] = reinterpret_cast<unsigned long long*>(base + s.off_flags);
            s.peer_ctl[r] = s.peer_flags.flags[r] + kMaxPeers;
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
    if (c->world == 1 && c->small_path && env_int("NBODY_MEDIUM", 1) != 0) {
        // one CTA per SM, all co-resident (the kernel's grid barrier needs that: cooperative launch)
        int coop = 0, per_sm = 0;
        CU(c, cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, c->shards[0].device));
        if (c->prec == NBODY_F32)
            CU(c, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, medium_step_kernel<float>, kMediumThreads, medium_smem_bytes<float>()));
        else
            CU(c, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, medium_step_kernel<double>, kMediumThreads, medium_smem_bytes<double>()));
        c->medium_max_n = size_t(c->prec == NBODY_F32 ? medium_max_n<float>() : medium_max_n<double>());
        c->medium_max_n = std::min(c->medium_max_n, size_t(std::max(env_int("NBODY_MEDIUM_MAX", int(c->medium_max_n)), 0)));
        if (coop && per_sm >= 1 && size_t(c->sm_count) * kMediumThreads >= c->medium_max_n) c->medium_ctas = c->sm_count;
    }
    if (c->world == 1 && c->prec == NBODY_F32 && env_int("NBODY_SYM_COOP", 1) != 0) {  // sym_coop.cuh needs cooperative launches too
        int coop = 0;
        CU(c, cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, c->shards[0].device));
        c->sym_coop = coop != 0;
    }
    if (c->medium_ctas == 0 && !getenv("NBODY_SMALL_MAX")) c->small_max_n = kSmallMaxN;  // nobody else runs all steps in one launch
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
    DeviceGuard device_guard_;
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
    DeviceGuard device_guard_;
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
    DeviceGuard device_guard_;
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
    DeviceGuard device_guard_;
    if (!c) return NBODY_EINVAL;
    c->heat_decay = heat_decay;
    return NBODY_OK;
}

extern "C" int nbody_set_stream(nbody_ctx* c, void* cuda_stream)
{
    DeviceGuard device_guard_;
    if (!c) return NBODY_EINVAL;
    Shard& s = c->shards[0];
    CU(c, cudaSetDevice(s.device));
    CU(c, cudaStreamSynchronize(s.stream));
    s.stream = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : s.own_stream;
    return NBODY_OK;
}

extern "C" int nbody_sync(nbody_ctx* c)
{
    DeviceGuard device_guard_;
    if (!c) return NBODY_EINVAL;
    for (Shard& s : c->shards) {
        CU(c, cudaSetDevice(s.device));
        CU(c, cudaStreamSynchronize(s.stream));
    }
    return check_exchange(c);
}

extern "C" int nbody_get_info(nbody_ctx* c, nbody_info* o)
{
    DeviceGuard device_guard_;
    if (!c || !o) return NBODY_EINVAL;
    memset(o, 0, sizeof *o);
    if (c->detect_pending) {
        const int rcd = resolve_mass_detection(c);
        if (rcd != NBODY_OK) return rcd;
    }
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
    o->f64_symmetric = c->f64_sym_k;
    o->f32_symmetric = c->sym32_run;
    o->f32_sym_bodies_per_thread = c->sym32_run > 0 ? c->sym32_R : 0;
    o->f32_sym_last_pass = c->sym32_used ? 1 : 0;
    o->upload_h2d_bytes = c->upload_h2d_bytes;
    o->f32_sym_warp_units = c->sym32_run > 0 && c->sym32_fine ? 1 : 0;
    return NBODY_OK;
}

// ============================================================================
// state
// ============================================================================

namespace {

int queue_exchange_barrier(nbody_ctx* c);  // defined with the exchange code below

template <typename Real> int upload_t(nbody_ctx* c, size_t n, const Real* pm, const Real* vel, const Real* acc, const Real* temp)
{
    const size_t v4 = 4 * sizeof(Real);
    bool small_io = false;
    // One process per GPU with peer stores: a peer that is a step ahead (or behind) reads and writes THIS rank's position
    // buffers and accumulators.  The upload is a collective there: an epoch barrier before the buffers are rewritten (every
    // rank has finished the work it had queued) and one after (every rank's buffers are packed before any peer store).
    const bool handshake = c->rank_mode && c->world > 1 && c->p2p;
    if (handshake) {
        const int rcb = queue_exchange_barrier(c);
        if (rcb != NBODY_OK) return rcb;
    }
    const bool want_mass = sizeof(Real) == 4 ? (c->sym32_allowed && (n >= c->sym32_min_n || (n >= c->symw_min_n && n < c->symw_max_n))) : c->f64_sym_k > 0;
    double small_mass = 0.0;
    // several devices with peer mappings: each shard uploads and packs its own slice only (pack_pos_shard_kernel)
    const bool sharded = c->world > 1 && c->p2p && env_int("NBODY_SHARDED_UPLOAD", 1) != 0;
    c->upload_sharded = sharded;
    c->upload_h2d_bytes = 0;
    if (sharded && !c->rank_mode)  // (one process: a shard's pack stores into the other shards' buffers -- let them go idle first)
        for (Shard& s : c->shards) {
            CU(c, cudaSetDevice(s.device));
            CU(c, cudaStreamSynchronize(s.stream));
        }
    for (Shard& s : c->shards) {
        CU(c, cudaSetDevice(s.device));
        s.n_local = c->n_pad / c->world;
        s.i_begin = size_t(s.rank) * s.n_local;
        const size_t lo = std::min(s.i_begin, n), hi = std::min(s.i_begin + s.n_local, n);
        s.n_local_real = hi - lo;
        if (s.io_host && n <= kSmallIoMaxN) {
            // small system on one device: one pinned record and one kernel (pack_state_kernel), no synchronise.  Up to
            // kSmallIoZeroCopyMaxN bodies the kernel reads the record where it lies (mapped pinned memory, a few us over
            // PCIe) and no copy operation enters the stream at all; above, one copy brings it to the device first.
            Real* rec = static_cast<Real*>(s.io_host);
            const bool zero_copy = n <= kSmallIoZeroCopyMaxN;
            CU(c, cudaEventSynchronize(s.io_done));  // the previous upload's last read of the record (long done, as a rule)
            memcpy(rec, pm, n * v4);
            memcpy(rec + 4 * n, vel, n * v4);
            if (acc) memcpy(rec + 8 * n, acc, n * v4); else memset(rec + 8 * n, 0, n * v4);
            if (temp) memcpy(rec + 12 * n, temp, n * sizeof(Real)); else memset(rec + 12 * n, 0, n * sizeof(Real));
            if (!zero_copy) CU(c, cudaMemcpyAsync(s.io_dev, rec, 13 * n * sizeof(Real), cudaMemcpyHostToDevice, s.stream));
            if (c->detect_masses) CU(c, cudaMemsetAsync(s.outlier_count, 0, sizeof(unsigned), s.stream));
            pack_state_kernel<Real><<<unsigned((c->n_pad + 255) / 256), 256, 0, s.stream>>>(
                zero_copy ? rec : static_cast<const Real*>(s.io_dev), static_cast<Real*>(s.pos[0]), static_cast<Real*>(s.pos[1]), static_cast<Real*>(s.vel),
                static_cast<Real*>(s.acc), static_cast<Real*>(s.temp), n, c->n_pad, Real(sizeof(Real) == 8 ? 1e150 : 1e18), Real(c->uniform_mass),
                c->detect_masses ? s.outlier_count : nullptr, s.outliers);
            c->launches++;
            CU(c, cudaGetLastError());
            CU(c, cudaEventRecord(s.io_done, s.stream));
            small_io = true;
            if (want_mass) {  // (a few thousand bodies: the host adds them, in index order)
                if (s.accfix) {  // the pairs-once pass may take this upload: its accumulators, poison words and counter start at zero
                    CU(c, cudaMemsetAsync(s.accfix, 0, ((sizeof(Real) == 4 ? 3 : 6) * c->cap_pad + 2) * sizeof(unsigned long long), s.stream));
                    CU(c, cudaMemsetAsync(s.sym_ctl, 0, sizeof(unsigned), s.stream));
                    CU(c, cudaMemsetAsync(s.sym_ctl + 2, 0, sizeof(unsigned), s.stream));
                }
                double t = 0.0;
                for (size_t i = 0; i < n; ++i) t += std::fabs(double(pm[4 * i + 3]));
                small_mass = t;
            }
        } else if (sharded) {
        // positions + masses: the OWN slice of the host AoS -> staging -> every rank's device layout (both buffers, peer stores)
        if (s.n_local_real) CU(c, cudaMemcpyAsync(s.stage, pm + 4 * lo, s.n_local_real * v4, cudaMemcpyHostToDevice, s.stream));
        c->upload_h2d_bytes += (long long)(s.n_local_real * v4);
        if (c->detect_masses) CU(c, cudaMemsetAsync(s.outlier_count, 0, sizeof(unsigned), s.stream));
        PackDst<Real> dst;
        dst.n = 0;
        for (int r = 0; r < c->world; ++r) {
            dst.p[dst.n++] = static_cast<Real*>(s.peer_pos[r][0]);
            dst.p[dst.n++] = static_cast<Real*>(s.peer_pos[r][1]);
        }
        pack_pos_shard_kernel<Real><<<unsigned((s.n_local + 255) / 256), 256, 0, s.stream>>>(
            static_cast<const Real*>(s.stage), dst, s.i_begin, s.n_local, n, Real(sizeof(Real) == 8 ? 1e150 : 1e18), Real(c->uniform_mass),
            c->detect_masses ? s.outlier_count : nullptr, s.outliers, want_mass ? s.mass_sums + s.i_begin / 256 : nullptr);
        CUK(c, "pack_pos_shard_kernel", s.device);
        } else {
        // positions + masses: host AoS -> staging -> device layout (both buffers)
        CU(c, cudaMemcpyAsync(s.stage, pm, n * v4, cudaMemcpyHostToDevice, s.stream));
        c->upload_h2d_bytes += (long long)(n * v4);
        const unsigned blocks = unsigned((c->n_pad + 255) / 256);
        const bool detect = c->detect_masses && &s == &c->shards[0];  // every shard packs the same array: one looks
        if (detect) CU(c, cudaMemsetAsync(s.outlier_count, 0, sizeof(unsigned), s.stream));
        pack_pos_kernel<Real><<<blocks, 256, 0, s.stream>>>(static_cast<const Real*>(s.stage), static_cast<Real*>(s.pos[0]),
                                                           static_cast<Real*>(s.pos[1]), n, c->n_pad, Real(sizeof(Real) == 8 ? 1e150 : 1e18), Real(c->uniform_mass),
                                                           detect ? s.outlier_count : nullptr, s.outliers,
                                                           want_mass && &s == &c->shards[0] ? s.mass_sums : nullptr);
        CUK(c, "pack_pos_kernel", s.device);
        }
        if (!small_io) {
        if (s.accfix) {  // (zero unless a pass was abandoned)
            CU(c, cudaMemsetAsync(s.accfix, 0, ((sizeof(Real) == 4 ? 3 : 6) * c->cap_pad + 2) * sizeof(unsigned long long), s.stream));
            CU(c, cudaMemsetAsync(s.sym_ctl, 0, sizeof(unsigned), s.stream));
            CU(c, cudaMemsetAsync(s.sym_ctl + 2, 0, sizeof(unsigned), s.stream));
        }
        c->launches++;
        CU(c, cudaGetLastError());
        CU(c, cudaMemsetAsync(s.vel, 0, s.n_local * v4, s.stream));
        CU(c, cudaMemsetAsync(s.acc, 0, s.n_local * v4, s.stream));
        CU(c, cudaMemsetAsync(s.temp, 0, s.n_local * sizeof(Real), s.stream));
        if (s.n_local_real) {
            CU(c, cudaMemcpyAsync(s.vel, vel + 4 * lo, s.n_local_real * v4, cudaMemcpyHostToDevice, s.stream));
            if (acc) CU(c, cudaMemcpyAsync(s.acc, acc + 4 * lo, s.n_local_real * v4, cudaMemcpyHostToDevice, s.stream));
            if (temp) CU(c, cudaMemcpyAsync(s.temp, temp + lo, s.n_local_real * sizeof(Real), cudaMemcpyHostToDevice, s.stream));
            c->upload_h2d_bytes += (long long)(s.n_local_real * (v4 * (acc ? 2 : 1) + (temp ? sizeof(Real) : 0)));
        }
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
    // host buffers are borrowed for the duration of the call only (the small-system path has copied them already)
    c->upload_deferred = small_io;
    c->sum_abs_mass = small_mass;
    if (small_io) return NBODY_OK;
    if (handshake) {
        const int rcb = queue_exchange_barrier(c);
        if (rcb != NBODY_OK) return rcb;
    }
    c->sum_abs_mass = 0.0;
    if (want_mass) {  // sum |m|, block sums added in index order: the same number on every rank and in every run
        std::vector<double> sums((c->n_pad + 255) / 256);
        if (!sharded) {
            Shard& s0 = c->shards[0];
            CU(c, cudaSetDevice(s0.device));
            CU(c, cudaMemcpyAsync(sums.data(), s0.mass_sums, sums.size() * sizeof(double), cudaMemcpyDeviceToHost, s0.stream));
            CU(c, cudaStreamSynchronize(s0.stream));
        } else if (!c->rank_mode) {  // every shard holds the sums of its own blocks
            for (Shard& s : c->shards) {
                CU(c, cudaSetDevice(s.device));
                const size_t b0 = s.i_begin / 256, nb = s.n_local / 256;
                CU(c, cudaMemcpyAsync(sums.data() + b0, s.mass_sums + b0, nb * sizeof(double), cudaMemcpyDeviceToHost, s.stream));
                CU(c, cudaStreamSynchronize(s.stream));
            }
        } else {                     // ... of which the other ranks' arrive by all-gather
            Shard& s = c->shards[0];
            CU(c, cudaSetDevice(s.device));
            const size_t nb = s.n_local / 256;
            NC(c, nccl().AllGather(s.mass_sums + size_t(s.rank) * nb, s.mass_sums, nb * sizeof(double), ncclUint8_, s.comm, s.stream));
            CU(c, cudaMemcpyAsync(sums.data(), s.mass_sums, sums.size() * sizeof(double), cudaMemcpyDeviceToHost, s.stream));
            CU(c, cudaStreamSynchronize(s.stream));
        }
        for (double v : sums) c->sum_abs_mass += v;
    }
    for (Shard& s : c->shards) {
        CU(c, cudaSetDevice(s.device));
        CU(c, cudaStreamSynchronize(s.stream));
    }
    return NBODY_OK;
}

template <typename Real> int download_t(nbody_ctx* c, Real* pm, Real* vel, Real* acc, Real* temp, bool own_only = false)
{
    const size_t v4 = 4 * sizeof(Real);
    if (own_only && pm && c->n) {  // positions of the own range(s) only, like vel / acc / temp below
        for (Shard& s : c->shards) {
            if (s.n_local_real == 0) continue;
            CU(c, cudaSetDevice(s.device));
            const size_t lo = std::min(s.i_begin, c->n);
            unpack_pos_range_kernel<Real><<<unsigned((s.n_local_real + 255) / 256), 256, 0, s.stream>>>(static_cast<const Real*>(s.pos[c->cur]),
                                                                                                   static_cast<Real*>(s.stage), lo, s.n_local_real);
            c->launches++;
            CU(c, cudaGetLastError());
            CU(c, cudaMemcpyAsync(pm + 4 * lo, s.stage, s.n_local_real * v4, cudaMemcpyDeviceToHost, s.stream));
        }
        pm = nullptr;
    }
    if (c->shards[0].io_host && c->n <= kSmallIoMaxN) {
        // small system on one device: one kernel gathers the state into a record, one copy, one synchronise (integrate.cuh)
        Shard& s = c->shards[0];
        const size_t n = c->n;
        const bool zero_copy = n <= kSmallIoZeroCopyMaxN;  // the kernel stores straight into the pinned record
        CU(c, cudaSetDevice(s.device));
        unpack_state_kernel<Real><<<unsigned((n + 255) / 256), 256, 0, s.stream>>>(static_cast<const Real*>(s.pos[c->cur]), static_cast<const Real*>(s.vel),
                                                                               static_cast<const Real*>(s.acc), static_cast<const Real*>(s.temp),
                                                                               static_cast<Real*>(zero_copy ? s.io_host : s.io_dev), n);
        c->launches++;
        CU(c, cudaGetLastError());
        const Real* rec = static_cast<const Real*>(s.io_host);
        if (!zero_copy) CU(c, cudaMemcpyAsync(s.io_host, s.io_dev, 13 * n * sizeof(Real), cudaMemcpyDeviceToHost, s.stream));
        CU(c, cudaStreamSynchronize(s.stream));
        if (pm) memcpy(pm, rec, n * v4);
        if (vel) memcpy(vel, rec + 4 * n, n * v4);
        if (acc) memcpy(acc, rec + 8 * n, n * v4);
        if (temp) memcpy(temp, rec + 12 * n, n * sizeof(Real));
        return check_exchange(c);
    }
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

// Partial-sum arrays of the fp64 pairs-once pass (force_f64_sym.cuh) for the current n_pad; they only grow.
int ensure_sym_buffers(nbody_ctx* c)
{
    Shard& s = c->shards[0];
    CU(c, cudaSetDevice(s.device));
    size_t bi = 0, bj = 0;
    const size_t budget = size_t(std::max(1, env_int("NBODY_F64_SYM_MAX_MB", 4096))) << 20;
    if (sym_partial_bytes(c->n_pad, c->f64_sym_r, c->f64_sym_k, &bi, &bj) > budget) return 1;  // the one-sided kernel then
    if (bi > s.sym_bytes_i) {
        cudaFree(s.sym_part_i);
        s.sym_part_i = nullptr;
        s.sym_bytes_i = 0;
        CU(c, cudaMalloc(&s.sym_part_i, bi));
        s.sym_bytes_i = bi;
    }
    if (bj > s.sym_bytes_j) {
        cudaFree(s.sym_part_j);
        s.sym_part_j = nullptr;
        s.sym_bytes_j = 0;
        CU(c, cudaMalloc(&s.sym_part_j, bj));
        s.sym_bytes_j = bj;
    }
    return NBODY_OK;
}

// The unit table of the fp32 pairs-once pass (force_f32_sym.cuh) for the current body count: row I = i-tile I against its own
// position tiles (one diagonal unit), then against the tiles above it in runs of `run`.  A pure function of n (and of the
// kernel shape), never of the GPU count: every partial sum a unit produces is the same number wherever the unit runs, and
// the GPUs share the table by cost (position-tile visits), in order.
void make_sym32_units(size_t n, int R, int run, int world, std::vector<SymUnit>& units, std::vector<unsigned>& split)
{
    const size_t itile = size_t(kF32Threads) * R, spt = itile / kPosTile;
    const size_t NS = (n + kPosTile - 1) / kPosTile;  // position tiles that hold bodies
    const size_t T = (n + itile - 1) / itile;
    std::vector<unsigned long long> cost;  // position-tile visits up to and including unit u
    unsigned long long total = 0;
    units.clear();
    for (size_t I = 0; I < T; ++I) {
        const size_t s0 = I * spt, s1 = std::min((I + 1) * spt, NS);
        units.push_back(SymUnit{unsigned(I), unsigned(s0) | unsigned(s1 - s0) << 24 | kSymUnitDiag});
        cost.push_back(total += s1 - s0);
        for (size_t st = s1; st < NS; st += size_t(run)) {
            const size_t cnt = std::min(size_t(run), NS - st);
            units.push_back(SymUnit{unsigned(I), unsigned(st) | unsigned(cnt) << 24});
            cost.push_back(total += cnt);
        }
    }
    split.assign(size_t(world) + 1, 0);
    size_t u = 0;
    for (int r = 1; r < world; ++r) {
        const unsigned long long target = total / world * r;
        while (u < units.size() && cost[u] < target) ++u;
        split[r] = unsigned(u);
    }
    split[world] = unsigned(units.size());
}

// The warp-unit form's table (force_f32_symw.cuh): i-blocks of R/2 groups of 64 bodies, each against its own groups (the
// diagonal unit, one-sided) and against the groups above it in runs of `run`.  The same rules: a function of n and of the
// kernel shape only; the GPUs share it by cost (group visits), in order.
void make_symw_units(size_t n, int R, int run, int world, std::vector<SymUnit>& units, std::vector<unsigned>& split)
{
    const size_t h = size_t(R) / 2;
    const size_t NG = (n + kSymwGroup - 1) / kSymwGroup;  // groups that hold bodies
    const size_t NB = (NG + h - 1) / h;
    std::vector<unsigned long long> cost;
    unsigned long long total = 0;
    units.clear();
    for (size_t b = 0; b < NB; ++b) {
        const size_t s0 = b * h, s1 = std::min((b + 1) * h, NG);
        units.push_back(SymUnit{unsigned(b), unsigned(s0) | unsigned(s1 - s0) << 24 | kSymUnitDiag});
        cost.push_back(total += s1 - s0);
        for (size_t st = s1; st < NG; st += size_t(run)) {
            const size_t cnt = std::min(size_t(run), NG - st);
            units.push_back(SymUnit{unsigned(b), unsigned(st) | unsigned(cnt) << 24});
            cost.push_back(total += cnt);
        }
    }
    split.assign(size_t(world) + 1, 0);
    size_t u = 0;
    for (int r = 1; r < world; ++r) {
        const unsigned long long target = total / world * r;
        while (u < units.size() && cost[u] < target) ++u;
        split[r] = unsigned(u);
    }
    split[world] = unsigned(units.size());
}

int build_sym32_table(nbody_ctx* c)
{
    c->sym32_run = 0;
    const bool mid = c->n >= c->symw_min_n && c->n < c->symw_max_n;
    if (c->prec != NBODY_F32 || !c->sym32_allowed || (!mid && c->n < c->sym32_min_n)) return NBODY_OK;
    if (c->detect_pending) {  // (a small upload returned before its mass detection was read)
        const int rcd = resolve_mass_detection(c);
        if (rcd != NBODY_OK) return rcd;
    }
    // Which form (profiles/r2r_fine_lab.jsonl, r2r_big.jsonl, r2o_*; G interactions/s, before -> warp units): below symw_max_n the warp
    // units always (N=8,192: 1,722 -> 2,016; 16,384: 2,428 -> 2,948; 32,768: 3,163 -> 3,637; 49,152: 3,663 -> 3,894; 65,536 and
    // 262,144 bodies with equal masses: 3,869 -> 3,881 and 4,167 -> 4,178, a tie).  Above it, general masses keep them -- N=1M:
    // 3,588 -> 3,834 -- and equal masses take the CTA-wide kernel, whose TMA ring is worth 0.5 % there (4,229 against 4,209).
    const bool fine = mid || (c->symw_max_n > 0 && c->n >= c->symw_min_n && !c->uniform && c->symw_general_large);
    int R, run;
    if (fine) {
        // 4 bodies per thread (12 warps per SM) up to ~28k bodies, 8 (8 warps) above: N=24,576 3,313 against 3,055, N=49,152 3,590 against 3,720
        R = (c->symw_R_forced == 2 || c->symw_R_forced == 4 || c->symw_R_forced == 8) ? c->symw_R_forced : (c->n < kSymwTinyN ? 2 : c->n < kSymwSmallN ? 4 : 8);
        // groups per unit: about eight units per resident warp of ONE device (a function of n only: the table, and with it every
        // bit of the result, must not depend on the GPU count), at least one group (~2 us of a warp at R = 4), and few enough
        // units that the table stays a few tens of MB at several million bodies
        const size_t NG = (c->n + kSymwGroup - 1) / kSymwGroup;
        const size_t rounds = NG * NG / size_t(R);
        const size_t warps = size_t(c->sm_count) * (R == 8 ? 8 : R >= 4 ? 12 : 16);
        run = env_int("NBODY_F32_SYMW_RUN", 0);
        if (run <= 0) {
            run = int(std::min<size_t>(16, std::max<size_t>(1, rounds / (warps * 8))));
            const size_t cells = NG * ((NG + size_t(R) / 2 - 1) / (size_t(R) / 2)) / 2;  // (i-block, group) cells above the diagonal
            while (run < kSymUnitMaxRun && cells / size_t(run) > (size_t(4) << 20)) ++run;
        }
        if (NG >= (1u << 24)) return NBODY_OK;
    } else {
        // (below ~49k bodies the 512-body i-tiles of R = 4 give the 444 resident CTAs twice as many units: N=32,768 2,961 against
        // 2,778 G interactions/s; from 65,536 up R = 8 wins, 3,767 against 3,534: profiles/r2a_sym_lab.jsonl)
        R = (c->sym32_R_forced == 4 || c->sym32_R_forced == 8) ? c->sym32_R_forced
            : (c->n < kSym32SmallN ? 4 : (c->uniform ? kSym32EqualMassR : kSym32GeneralR));
        const size_t NS = (c->n + kPosTile - 1) / kPosTile;  // position tiles that hold bodies
        if (NS >= (1u << 24)) return NBODY_OK;
        run = env_int("NBODY_F32_SYM_RUN", 0);
        // N=1M: 2 tiles (1,024 bodies) per unit = 0.13 ms of one CTA: the tail of a device's share stays ~0.2 % even when 8 GPUs
        // split the step (units of 1..16 tiles run within 0.5 % of each other on one GPU, profiles/r2a_sym_lab.jsonl)
        if (run <= 0) run = int(std::min<size_t>(8, std::max<size_t>(1, NS / 1024)));
    }
    c->sym32_R = R;
    c->sym32_fine = fine;
    run = std::min(run, kSymUnitMaxRun);
    if (c->sym32_table_n != c->n || c->sym32_table_R != R || c->sym32_table_run != run || c->sym32_table_fine != fine) {
        std::vector<SymUnit> units;
        if (fine) make_symw_units(c->n, R, run, c->world, units, c->sym32_split);
        else make_sym32_units(c->n, R, run, c->world, units, c->sym32_split);
        for (Shard& s : c->shards) {
            CU(c, cudaSetDevice(s.device));
            if (units.size() > s.sym_units_cap) {
                CU(c, cudaStreamSynchronize(s.stream));
                cudaFree(s.sym_units);
                s.sym_units = nullptr;
                s.sym_units_cap = 0;
                CU(c, cudaMalloc(&s.sym_units, units.size() * sizeof(SymUnit)));
                s.sym_units_cap = units.size();
            }
            CU(c, cudaMemcpyAsync(s.sym_units, units.data(), units.size() * sizeof(SymUnit), cudaMemcpyHostToDevice, s.stream));
            CU(c, cudaStreamSynchronize(s.stream));  // (`units` is pageable and goes out of scope)
        }
        c->sym32_table_n = c->n;
        c->sym32_table_R = R;
        c->sym32_table_run = run;
        c->sym32_table_fine = fine;
    }
    c->sym32_run = run;
    return NBODY_OK;
}

// After upload_t's final synchronise: what pack_pos_kernel found.  Up to kMaxOutliers exceptions keep
// the equal-mass sweep (they are corrected per i-body); the list is sorted so that the order of the
// corrections, and with it every bit of the result, does not depend on the order the atomics landed in.
int finish_mass_detection(nbody_ctx* c)
{
    // (exceptions are corrected in the fp32 epilogue only: an fp64 context takes the sweep for exactly equal masses)
    const size_t max_out = c->prec == NBODY_F64 ? 0 : std::min<size_t>(size_t(std::max(0, env_int("NBODY_OUTLIERS", kMaxOutliers))), kMaxOutliers);
    struct Found {
        unsigned count;
        MassOutlier list[kMaxOutliers];
    };
    std::vector<Found> found;
    auto read_shard = [&](Shard& s, Found& f) -> int {
        CU(c, cudaSetDevice(s.device));
        CU(c, cudaMemcpy(&f.count, s.outlier_count, sizeof(unsigned), cudaMemcpyDeviceToHost));
        const unsigned k = std::min<unsigned>(f.count, kMaxOutliers);
        if (k) CU(c, cudaMemcpy(f.list, s.outliers, k * sizeof(MassOutlier), cudaMemcpyDeviceToHost));
        return NBODY_OK;
    };
    if (!c->upload_sharded) {           // one shard looked at the whole array
        found.resize(1);
        const int rc = read_shard(c->shards[0], found[0]);
        if (rc != NBODY_OK) return rc;
    } else if (!c->rank_mode) {         // every shard looked at its own slice
        found.resize(c->shards.size());
        for (size_t k = 0; k < c->shards.size(); ++k) {
            const int rc = read_shard(c->shards[k], found[k]);
            if (rc != NBODY_OK) return rc;
        }
    } else {                            // ... and the other ranks' findings arrive by all-gather
        found.resize(c->world);
        Shard& s = c->shards[0];
        Found mine;
        memset(&mine, 0, sizeof mine);
        int rc = read_shard(s, mine);
        if (rc != NBODY_OK) return rc;
        Found* dev = nullptr;
        CU(c, cudaMalloc(&dev, sizeof(Found) * c->world));
        auto gather = [&]() -> int {
            CU(c, cudaMemcpyAsync(dev + c->my_rank, &mine, sizeof(Found), cudaMemcpyHostToDevice, s.stream));
            NC(c, nccl().AllGather(dev + c->my_rank, dev, sizeof(Found), ncclUint8_, s.comm, s.stream));
            CU(c, cudaMemcpyAsync(found.data(), dev, sizeof(Found) * c->world, cudaMemcpyDeviceToHost, s.stream));
            CU(c, cudaStreamSynchronize(s.stream));
            return NBODY_OK;
        };
        rc = gather();
        cudaFree(dev);
        if (rc != NBODY_OK) return rc;
    }
    size_t count = 0;
    for (const Found& f : found) count += f.count;
    if (count > max_out) return NBODY_OK;  // too many different masses: the general kernel
    c->uniform = true;
    if (count == 0) return NBODY_OK;
    for (const Found& f : found) c->outliers.insert(c->outliers.end(), f.list, f.list + f.count);
    std::sort(c->outliers.begin(), c->outliers.end(), [](const MassOutlier& a, const MassOutlier& b) { return a.index < b.index; });
    for (Shard& s : c->shards) {
        CU(c, cudaSetDevice(s.device));
        CU(c, cudaMemcpy(s.outliers, c->outliers.data(), count * sizeof(MassOutlier), cudaMemcpyHostToDevice));
    }
    return NBODY_OK;
}

// The small-system upload leaves the mass-detection count in flight; whoever needs `uniform` first waits for it.
int resolve_mass_detection(nbody_ctx* c)
{
    if (!c->detect_pending) return NBODY_OK;
    c->detect_pending = false;
    Shard& s0 = c->shards[0];
    CU(c, cudaSetDevice(s0.device));
    CU(c, cudaStreamSynchronize(s0.stream));
    return finish_mass_detection(c);
}

}  // namespace

extern "C" int nbody_upload(nbody_ctx* c, size_t n, const void* pos_mass, const void* vel, const void* acc, const void* temp)
{
    DeviceGuard device_guard_;
    if (!c) return NBODY_EINVAL;
    if (n == 0 || n > c->capacity) return fail(c, NBODY_EINVAL, "upload: n=%zu outside 1..capacity=%zu", n, c->capacity);
    if (!pos_mass || !vel) return fail(c, NBODY_EINVAL, "upload: pos_mass and vel are required");
    cudaGetLastError();  // (see check_ready)
    c->n = n;
    c->f64_sym_k = 0;
    if (c->prec == NBODY_F32) {
        // small systems: 128-body j tiles, four times as many and four times shorter CTAs (force_f32.cuh)
        const int forced = env_int("NBODY_F32_TILE", 0);  // bodies per tile: 128 or 512; 0 = by body count
        const bool small_tile = c->R == 4 && (forced == 2 * kF32SmallTilePairs || (forced != 2 * kF32TilePairs && n <= kF32SmallTileMaxN));
        c->j_unit = small_tile ? 2 * kF32SmallTilePairs : kJUnit;
    }
    if (c->prec == NBODY_F64) {
        c->R = f64_bodies_per_thread(c, n);
        c->i_tile = kThreads * c->R;
        c->f64_sym_k = sym_pick_k(c, n, &c->f64_sym_r);
    }
    c->n_pad = round_up(n, pad_unit(c));
    c->n_j = round_up(n, size_t(c->j_unit));
    c->cur = 0;
    c->has_temp = temp != nullptr;
    c->uniform = false;
    c->outliers.clear();
    c->detect_masses = false;
    c->detect_pending = false;
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
    c->detect_pending = rc == NBODY_OK && c->detect_masses && c->upload_deferred;
    if (rc == NBODY_OK && c->detect_masses && !c->detect_pending) rc = finish_mass_detection(c);
    c->detect_masses = false;
    if (rc == NBODY_OK) rc = build_sym32_table(c);
    return rc;
}

// page_world.cu: a world step changed the body count on the device.  Re-derive what nbody_upload derives from n (fp64, one GPU);
// the kernel has already rewritten the rows beyond n as pads, and the radii stay valid (it maintains them).
int world_adopt_count(nbody_ctx* c, size_t n)
{
    if (n == 0 || n > c->capacity) return fail(c, NBODY_EINVAL, "world: the device holds %zu bodies, capacity %zu", n, c->capacity);
    Shard& s = c->shards[0];
    CU(c, cudaSetDevice(s.device));
    const bool changed = n != c->n;
    c->n = n;
    c->R = f64_bodies_per_thread(c, n);
    c->i_tile = kThreads * c->R;
    c->f64_sym_k = sym_pick_k(c, n, &c->f64_sym_r);
    c->n_pad = round_up(n, pad_unit(c));
    c->n_j = round_up(n, size_t(c->j_unit));
    c->has_temp = true;
    c->uniform = false;  // (merged masses: the general kernels)
    c->outliers.clear();
    c->detect_pending = false;
    c->near_valid = false;
    if (changed) {
        if (c->hist_n) c->hist_prev_n = c->hist_n;
        c->hist_n = 0;
        c->has_materials = false;
    }
    choose_chunks(c);
    s.n_local = c->n_pad;
    s.i_begin = 0;
    s.n_local_real = n;
    const size_t need = c->j_chunks > 1 ? size_t(c->j_chunks) * s.n_local * 4 * c->esz : 0;
    if (need > s.partial_bytes) {
        cudaFree(s.partial);
        s.partial = nullptr;
        s.partial_bytes = 0;
        CU(c, cudaMalloc(&s.partial, need));
        s.partial_bytes = need;
    }
    return NBODY_OK;
}

extern "C" int nbody_download_own(nbody_ctx* c, void* pos_mass, void* vel, void* acc, void* temp)
{
    DeviceGuard device_guard_;
    if (!c) return NBODY_EINVAL;
    if (c->n == 0) return fail(c, NBODY_EINVAL, "download before upload");
    return c->prec == NBODY_F32 ? download_t<float>(c, (float*)pos_mass, (float*)vel, (float*)acc, (float*)temp, true)
                                : download_t<double>(c, (double*)pos_mass, (double*)vel, (double*)acc, (double*)temp, true);
}

extern "C" int nbody_download(nbody_ctx* c, void* pos_mass, void* vel, void* acc, void* temp)
{
    DeviceGuard device_guard_;
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
// One cross-GPU epoch barrier (exchange.cuh) queued on every shard's stream: when it retires on a device, every rank has
// finished the work queued before ITS barrier -- peer stores into this device's buffers have landed, peer reads of them are over.
int queue_exchange_barrier(nbody_ctx* c)
{
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

int exchange_positions(nbody_ctx* c, int which)
{
    if (c->world == 1) return NBODY_OK;
    if (c->p2p) return queue_exchange_barrier(c);
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

int start_force_timing(nbody_ctx* c, Shard& s)
{
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
    return NBODY_OK;
}

// The fp32 pairs-once pass (force_f32_sym.cuh): persistent CTAs work through this device's share of the unit table and add
// fixed-point partial sums to the accumulators of ALL bodies; after a cross-GPU barrier one O(N) kernel per device adds the
// peers' accumulators for the bodies it owns and runs the integrator epilogue.  Returns 1 when the pass does not apply
// (the caller takes the one-sided sweep), 0 when it was queued.
int launch_force_sym32(nbody_ctx* c, int mode, double dt, double G, double eps)
{
    if (c->sym32_run <= 0 || !(eps > 0.0) || !std::isfinite(eps)) return 1;
    const bool uniform = c->uniform;
    // |sum| <= sum_j |m_j| * max_r r / (r^2 + eps^2)^1.5 = sum|m| * 2 / (3 sqrt(3) eps^2); the equal-mass sweep counts every body as 1
    const double eps2 = double(float(eps) * float(eps));
    const double wmax = 2.0 / (3.0 * std::sqrt(3.0) * eps2);
    const double bound = 1.001 * wmax * (uniform ? double(c->n) : c->sum_abs_mass);
    if (!(bound > 0.0) || !std::isfinite(bound) || bound > 1e37) return 1;
    int e = 0;
    std::frexp(bound, &e);         // bound < 2^e
    const int k = std::max(-100, std::min(100, 61 - e));  // bound * 2^k < 2^61
    if (61 - e < -100) return 1;
    const bool fine = c->sym32_fine;
    int& slots = fine ? c->symw_slots[uniform ? 1 : 0][c->sym32_R / 2 - 1] : c->sym32_slots[uniform ? 1 : 0][c->sym32_R == 8 ? 1 : 0];
    if (slots == 0) {
        int per_sm = 0;
        cudaError_t err = cudaSuccess;
        CU(c, cudaSetDevice(c->shards[0].device));
        if (fine)
            pick_f32_symw_kernel(c->sym32_R, uniform, [&](auto kernel) { err = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, kSymwThreads, 0); });
        else
            pick_f32_sym_kernel(c->sym32_R, uniform, [&](auto kernel) {
                err = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, kF32Threads + 32, force_f32_sym_smem_bytes());
            });
        if (err != cudaSuccess || per_sm < 1) return fail(c, NBODY_ECUDA, "pairs-once kernel does not fit an SM: %s", cudaGetErrorString(err));
        slots = per_sm * c->sm_count;
    }
    for (Shard& s : c->shards) {
        CU(c, cudaSetDevice(s.device));
        if (c->timing && &s == &c->shards[0]) {
            const int rc = start_force_timing(c, s);
            if (rc != NBODY_OK) return rc;
        }
        Sym32Params sp;
        sp.pos = static_cast<const float*>(s.pos[c->cur]);
        sp.units = static_cast<const SymUnit*>(s.sym_units);
        sp.u_begin = c->sym32_split[s.rank];
        sp.u_end = c->sym32_split[s.rank + 1];
        sp.sched = s.sym_ctl;
        sp.accfix = s.accfix;
        sp.n_pad = c->n_pad;
        sp.eps2 = float(eps) * float(eps);
        sp.scale = std::ldexp(1.0f, k);
        sp.bound = std::nextafter(float(bound), INFINITY);
        if (c->world == 1) {
            // nothing to clear: the previous finish kernel rewound the counter and cleared the poison word (Sym32Sources)
        } else {
            CU(c, cudaMemsetAsync(s.sym_ctl, 0, sizeof(unsigned), s.stream));
            if (!c->p2p)  // (after the all-reduce below every device holds ALL bodies' totals and clears only its own)
                CU(c, cudaMemsetAsync(s.accfix, 0, (3 * c->n_pad + 1) * sizeof(unsigned long long), s.stream));
            else
                CU(c, cudaMemsetAsync(s.accfix + 3 * c->n_pad, 0, sizeof(unsigned long long), s.stream));
        }
        const unsigned units = sp.u_end - sp.u_begin;
        if (fine) {  // every warp a worker
            const unsigned grid = std::max(1u, std::min((units + kSymwWarps - 1) / kSymwWarps, unsigned(slots)));
            pick_f32_symw_kernel(c->sym32_R, uniform, [&](auto kernel) { kernel<<<grid, kSymwThreads, 0, s.stream>>>(sp); });
        } else {
            const unsigned grid = std::max(1u, std::min(units, unsigned(slots)));
            pick_f32_sym_kernel(c->sym32_R, uniform, [&](auto kernel) { kernel<<<grid, kF32Threads + 32, force_f32_sym_smem_bytes(), s.stream>>>(sp); });
        }
        c->launches++;
        CU(c, cudaGetLastError());
    }
    if (c->world > 1 && c->p2p) {  // every rank's accumulators are complete
        const int rc = queue_exchange_barrier(c);
        if (rc != NBODY_OK) return rc;
    } else if (c->world > 1) {  // no peer mappings: an exact all-reduce of the accumulators (and the poison word)
        if (!c->rank_mode) NC(c, nccl().GroupStart());
        for (Shard& s : c->shards) {
            if (!c->rank_mode) CU(c, cudaSetDevice(s.device));
            NC(c, nccl().AllReduce(s.accfix, s.accfix, 3 * c->n_pad + 1, ncclInt64_, ncclSum_, s.comm, s.stream));
        }
        if (!c->rank_mode) NC(c, nccl().GroupEnd());
    }
    for (Shard& s : c->shards) {
        CU(c, cudaSetDevice(s.device));
        ForceParams<float> p = make_params<float>(c, s, mode, dt, G, eps);
        if (uniform && c->outliers.empty()) {
            p.G = float(G) * float(c->uniform_mass);
        } else if (uniform) {
            p.n_outliers = int(c->outliers.size());
            p.bulk_mass = float(c->uniform_mass);
        }
        Sym32Sources src;
        src.n = 0;
        src.inv_scale = std::ldexp(1.0, -k);
        src.sched_rewind = c->world == 1 ? s.sym_ctl : nullptr;
        src.done = s.sym_ctl + 2;
        if (c->world == 1 || !c->p2p) {
            src.acc[src.n++] = s.accfix;
        } else {
            for (int r = 0; r < c->world; ++r) src.acc[src.n++] = static_cast<unsigned long long*>(s.peer_acc[r]);
        }
        const unsigned blocks = unsigned((s.n_local + kSym32FinishThreads - 1) / kSym32FinishThreads);
        if (uniform) f32_sym_finish_kernel<true><<<blocks, kSym32FinishThreads, 0, s.stream>>>(p, src, c->n_pad);
        else f32_sym_finish_kernel<false><<<blocks, kSym32FinishThreads, 0, s.stream>>>(p, src, c->n_pad);
        c->launches++;
        CU(c, cudaGetLastError());
        if (c->timing && &s == &c->shards[0]) {
            CU(c, cudaEventRecord(s.ev[s.ev_used + 1], s.stream));
            s.ev_used += 2;
        }
    }
    // the accumulators a peer has just read and cleared belong to the next pass: nobody may add to them before every finish
    // kernel has retired.  A fused step gets that barrier from its position exchange (nbody_step).
    if (c->world > 1 && c->p2p && mode != kForceKick2Kick1) return queue_exchange_barrier(c);
    return NBODY_OK;
}

template <typename F> void pick_f64_sym_fixed_kernel(bool uniform, int r, F&& f)
{
    if (uniform) {
        if (r == 1) return f(force_f64_sym_kernel<false, true, 1, true>);
        if (r == 2) return f(force_f64_sym_kernel<false, true, 2, true>);
        return f(force_f64_sym_kernel<false, true, 4, true>);
    }
    if (r == 1) return f(force_f64_sym_kernel<false, false, 1, true>);
    if (r == 2) return f(force_f64_sym_kernel<false, false, 2, true>);
    return f(force_f64_sym_kernel<false, false, 4, true>);
}

// The fp64 pairs-once pass with exact fixed-point accumulators (force_f64_sym.cuh, FIXED): softening > 0, any number of GPUs.
// The ranks share the rows (i-blocks) of the pair triangle by cost; each accumulates hi/lo contributions for ALL bodies; after a
// barrier every device adds the ranks' accumulators for the bodies it owns and runs the integrator.  Returns 1 when it does not apply.
int launch_force_sym64(nbody_ctx* c, int mode, double dt, double G, double eps)
{
    if (c->f64_sym_k <= 0 || !(eps > 0.0) || !std::isfinite(eps)) return 1;
    const bool uni = c->uniform && c->f64_math == 2;
    const double wmax = 2.0 / (3.0 * std::sqrt(3.0) * eps * eps);
    const double bound = 1.001 * wmax * (uni ? double(c->n) : c->sum_abs_mass);
    if (!(bound > 0.0) || !std::isfinite(bound) || bound > 1e250) return 1;
    int e = 0;
    std::frexp(bound, &e);
    const int k1 = 61 - e;  // bound * 2^k1 < 2^61
    if (k1 < -900 || k1 > 900) return 1;
    const int r = c->f64_sym_r, k = c->f64_sym_k;
    const int n_iblocks = int(c->n_pad / (size_t(kSymThreads) * r));
    // rows by cost: row I walks (n_iblocks - I) * r j-blocks
    std::vector<int> split(size_t(c->world) + 1, 0);
    {
        double total = 0.0;
        for (int I = 0; I < n_iblocks; ++I) total += double(n_iblocks - I);
        double run = 0.0;
        int I = 0;
        for (int g = 1; g < c->world; ++g) {
            while (I < n_iblocks && run + 0.5 * double(n_iblocks - I) < total * g / c->world) run += double(n_iblocks - I++);
            split[g] = I;
        }
        split[c->world] = n_iblocks;
    }
    const size_t lo_off = 3 * c->cap_pad + 1;
    const size_t words = lo_off + 3 * c->n_pad;
    for (Shard& s : c->shards) {
        CU(c, cudaSetDevice(s.device));
        if (c->timing && &s == &c->shards[0]) {
            const int rc = start_force_timing(c, s);
            if (rc != NBODY_OK) return rc;
        }
        if (c->world > 1 && !c->p2p) CU(c, cudaMemsetAsync(s.accfix, 0, words * sizeof(unsigned long long), s.stream));
        else CU(c, cudaMemsetAsync(s.accfix + 3 * c->n_pad, 0, sizeof(unsigned long long), s.stream));
        const int I0 = split[s.rank], I1 = split[s.rank + 1];
        if (I1 > I0) {
            SymParams sp;
            memset(&sp, 0, sizeof sp);
            sp.pos = static_cast<const double*>(s.pos[c->cur]);
            sp.n_pad = c->n_pad;
            sp.r = r;
            sp.n_iblocks = n_iblocks;
            sp.k = k;
            sp.eps2 = eps * eps;
            sp.i_block_begin = I0;
            sp.acc_hi = s.accfix;
            sp.acc_lo = s.accfix + lo_off;
            sp.scale_hi = std::ldexp(1.0, k1);
            sp.inv_scale_hi = std::ldexp(1.0, -k1);
            sp.scale_lo = std::ldexp(1.0, k1 + 50);
            sp.bound = bound;
            const dim3 sgrid(unsigned(sym_chunks_of_row(n_iblocks, r, k, I0)), unsigned(I1 - I0));
            pick_f64_sym_fixed_kernel(uni, r, [&](auto kernel) { kernel<<<sgrid, kSymThreads, 0, s.stream>>>(sp); });
            c->launches++;
            CUK(c, "force_f64_sym_kernel", s.device);
        }
    }
    if (c->world > 1 && c->p2p) {
        const int rc = queue_exchange_barrier(c);
        if (rc != NBODY_OK) return rc;
    } else if (c->world > 1) {
        if (!c->rank_mode) NC(c, nccl().GroupStart());
        for (Shard& s : c->shards) {
            if (!c->rank_mode) CU(c, cudaSetDevice(s.device));
            NC(c, nccl().AllReduce(s.accfix, s.accfix, words, ncclInt64_, ncclSum_, s.comm, s.stream));
        }
        if (!c->rank_mode) NC(c, nccl().GroupEnd());
    }
    for (Shard& s : c->shards) {
        CU(c, cudaSetDevice(s.device));
        ForceParams<double> p = make_params<double>(c, s, mode, dt, G, eps);
        if (uni) p.G = G * c->uniform_mass;
        Sym64Sources src;
        src.n = 0;
        src.inv_scale_hi = std::ldexp(1.0, -k1);
        src.inv_scale_lo = std::ldexp(1.0, -(k1 + 50));
        if (c->world == 1 || !c->p2p) {
            src.hi[0] = s.accfix;
            src.lo[0] = s.accfix + lo_off;
            src.n = 1;
        } else {
            for (int g = 0; g < c->world; ++g) {
                src.hi[src.n] = static_cast<unsigned long long*>(s.peer_acc[g]);
                src.lo[src.n] = src.hi[src.n] + lo_off;
                src.n++;
            }
        }
        f64_sym_fixed_finish_kernel<<<unsigned((s.n_local + 255) / 256), 256, 0, s.stream>>>(p, src, c->n_pad, p.G);
        c->launches++;
        CUK(c, "f64_sym_fixed_finish_kernel", s.device);
        if (c->timing && &s == &c->shards[0]) {
            CU(c, cudaEventRecord(s.ev[s.ev_used + 1], s.stream));
            s.ev_used += 2;
        }
    }
    if (c->world > 1 && c->p2p && mode != kForceKick2Kick1) return queue_exchange_barrier(c);
    return NBODY_OK;
}

int launch_force(nbody_ctx* c, int mode, double dt, double G, double eps)
{
    if (c->detect_pending) {
        const int rcd = resolve_mass_detection(c);
        if (rcd != NBODY_OK) return rcd;
    }
    c->sym32_used = false;
    if (c->prec == NBODY_F64) {
        const int rcs = launch_force_sym64(c, mode, dt, G, eps);
        if (rcs <= 0) {
            c->near_valid = false;
            return rcs;
        }
    }
    if (c->prec == NBODY_F32 && !(c->has_radius && env_int("NBODY_NEAR", 1) != 0)) {
        const int rcs = launch_force_sym32(c, mode, dt, G, eps);
        if (rcs <= 0) {
            c->near_valid = false;
            c->sym32_used = rcs == 0;
            return rcs;
        }
    }
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
            const int rct = start_force_timing(c, s);
            if (rct != NBODY_OK) return rct;
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
            const int tp = c->j_unit / 2;  // j-pairs per ring stage: the j unit is one tile
            const size_t sm = tp == kF32SmallTilePairs ? force_f32_smem_bytes<kF32SmallTilePairs>() : force_f32_smem_bytes();
            const dim3 blk(kF32Threads + 32);
            pick_f32_kernel(track_near, uniform, eps == 0.0, c->R, tp, [&](auto kernel) { kernel<<<grid, blk, sm, s.stream>>>(p); });
        } else {
            const ForceParams<double> p = make_params<double>(c, s, mode, dt, G, eps);
            const dim3 blk(kThreads + 32);
            const size_t sm = force_f64_smem_bytes();
            ForceParams<double> q = p;
            const bool uni = c->uniform && c->f64_math == 2;
            if (uni) q.G = G * c->uniform_mass;  // equal masses: the common m joins G (force_f64.cuh, UNIFORM)
            int sym_rc = 1;
            if (c->f64_sym_k > 0 && c->world == 1 && eps == 0.0) sym_rc = ensure_sym_buffers(c);  // (array form; 1 = over its budget)
            if (sym_rc < 0) return sym_rc;
            if (sym_rc == 0) {
                // every pair once, then one pass that adds the partial sums and runs the epilogue (force_f64_sym.cuh)
                SymParams sp;
                memset(&sp, 0, sizeof sp);
                sp.pos = q.pos_cur;
                sp.part_i = s.sym_part_i;
                sp.part_j = s.sym_part_j;
                sp.n_pad = c->n_pad;
                sp.r = c->f64_sym_r;
                sp.n_iblocks = int(c->n_pad / (size_t(kSymThreads) * sp.r));
                sp.k = c->f64_sym_k;
                sp.eps2 = q.eps2;
                const dim3 sgrid(unsigned(sym_chunks_of_row(sp.n_iblocks, sp.r, sp.k, 0)), unsigned(sp.n_iblocks));
                pick_f64_sym_kernel(eps == 0.0, uni, sp.r, [&](auto kernel) { kernel<<<sgrid, kSymThreads, 0, s.stream>>>(sp); });
                c->launches++;
                CU(c, cudaGetLastError());
                f64_sym_finish_kernel<<<unsigned(c->n_pad / 32), kSymFinishWarps * 32, 0, s.stream>>>(q, sp, q.G);
            } else {
                pick_f64_kernel(c->f64_math, eps == 0.0, uni, c->R, [&](auto kernel) { kernel<<<grid, blk, sm, s.stream>>>(q); });
            }
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
    p.rec_out = static_cast<Real*>(c->step_rec_out);  // nbody_step_host: the kernel writes the caller's record itself
    c->step_rec_out = nullptr;
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

// ... and of a system of up to medium_max_n bodies in one cooperative launch over all SMs (medium.cuh).
template <typename Real> int launch_medium_steps_t(nbody_ctx* c, Shard& s, double dt, double G, double eps, int nsteps)
{
    MediumParams<Real> p;
    p.pos[0] = static_cast<Real*>(s.pos[c->cur]);
    p.pos[1] = static_cast<Real*>(s.pos[c->cur ^ 1]);
    p.vel = static_cast<Real*>(s.vel);
    p.acc = static_cast<Real*>(s.acc);
    p.temp = c->has_temp ? static_cast<Real*>(s.temp) : nullptr;
    p.arrivals = s.medium_arrivals;
    p.n = int(c->n);
    p.nsteps = nsteps;
    p.G = Real(G);
    p.eps2 = Real(eps) * Real(eps);
    p.r2_floor = eps == 0.0 ? Real(1e-20) : Real(0);
    p.half_dt = Real(0.5) * Real(dt);
    p.dt = Real(dt);
    p.heat_decay = Real(c->heat_decay);
    CU(c, cudaMemsetAsync(s.medium_arrivals, 0, sizeof(unsigned), s.stream));
    void* args[] = {&p};
    CU(c, cudaLaunchCooperativeKernel(reinterpret_cast<void*>(medium_step_kernel<Real>), dim3(unsigned(c->medium_ctas)), dim3(kMediumThreads), args,
                                      medium_smem_bytes<Real>(), s.stream));
    c->launches++;
    c->cur ^= nsteps & 1;  // step s drifts into the other buffer
    return NBODY_OK;
}

template <typename F> void pick_sym_coop_kernel(int R, bool uniform, F&& f)
{
    if (R == 8) return uniform ? f(sym_coop_step_kernel<8, true>) : f(sym_coop_step_kernel<8, false>);
    if (R == 2) return uniform ? f(sym_coop_step_kernel<2, true>) : f(sym_coop_step_kernel<2, false>);
    return uniform ? f(sym_coop_step_kernel<4, true>) : f(sym_coop_step_kernel<4, false>);
}

// Mid-size fp32 systems on one GPU: all nsteps steps in ONE cooperative launch around the warp-unit pairs-once pass
// (sym_coop.cuh).  Returns 1 when it does not apply (the caller launches per step), 0 when it was queued.
int launch_sym_coop_steps(nbody_ctx* c, double dt, double G, double eps, int nsteps)
{
    if (c->prec != NBODY_F32 || c->world != 1 || c->timing || !c->sym_coop || c->n > c->sym_coop_max_n || !(eps > 0.0) || !std::isfinite(eps)) return 1;
    if (c->has_radius && env_int("NBODY_NEAR", 1) != 0) return 1;  // (the sweep that tracks nearest neighbours)
    if (c->detect_pending) {
        const int rcd = resolve_mass_detection(c);
        if (rcd != NBODY_OK) return rcd;
    }
    if (c->sym32_run <= 0 || !c->sym32_fine || (c->uniform && !c->outliers.empty())) return 1;
    const bool uniform = c->uniform;
    const double eps2 = double(float(eps) * float(eps));
    const double bound = 1.001 * (2.0 / (3.0 * std::sqrt(3.0) * eps2)) * (uniform ? double(c->n) : c->sum_abs_mass);  // launch_force_sym32's
    if (!(bound > 0.0) || !std::isfinite(bound) || bound > 1e37) return 1;
    int e = 0;
    std::frexp(bound, &e);
    if (61 - e < -100) return 1;
    const int k = std::max(-100, std::min(100, 61 - e));
    Shard& s = c->shards[0];
    CU(c, cudaSetDevice(s.device));
    const int ri = c->sym32_R / 2 - 1;
    int& slots = c->sym_coop_slots[uniform ? 1 : 0][ri];
    if (slots == 0) {
        int per_sm = 0;
        cudaError_t err = cudaSuccess;
        pick_sym_coop_kernel(c->sym32_R, uniform, [&](auto kernel) { err = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, kSymwThreads, 0); });
        if (err != cudaSuccess || per_sm < 1) return fail(c, NBODY_ECUDA, "sym_coop_step_kernel does not fit an SM: %s", cudaGetErrorString(err));
        slots = per_sm * c->sm_count;
    }
    if (size_t(slots) * kSymwThreads < c->n_pad) return 1;  // one owner thread per body slot
    SymCoopParams p;
    p.sym.pos = static_cast<const float*>(s.pos[c->cur]);
    p.sym.units = static_cast<const SymUnit*>(s.sym_units);
    p.sym.u_begin = c->sym32_split[0];
    p.sym.u_end = c->sym32_split[1];
    p.sym.sched = s.sym_ctl;
    p.sym.accfix = s.accfix;
    p.sym.n_pad = c->n_pad;
    p.sym.eps2 = float(eps) * float(eps);
    p.sym.scale = std::ldexp(1.0f, k);
    p.sym.bound = std::nextafter(float(bound), INFINITY);
    p.pos = static_cast<float*>(s.pos[c->cur]);
    p.vel = static_cast<float*>(s.vel);
    p.acc = static_cast<float*>(s.acc);
    p.temp = c->has_temp ? static_cast<float*>(s.temp) : nullptr;
    p.arrivals = s.medium_arrivals;
    p.n = unsigned(c->n);
    p.nsteps = nsteps;
    p.G = uniform ? float(G) * float(c->uniform_mass) : float(G);
    p.half_dt = 0.5f * float(dt);
    p.dt = float(dt);
    p.heat_decay = float(c->heat_decay);
    p.inv_scale = std::ldexp(1.0, -k);
    CU(c, cudaMemsetAsync(s.medium_arrivals, 0, sizeof(unsigned), s.stream));
    void* args[] = {&p};
    cudaError_t err = cudaSuccess;
    pick_sym_coop_kernel(c->sym32_R, uniform, [&](auto kernel) {
        err = cudaLaunchCooperativeKernel(reinterpret_cast<void*>(kernel), dim3(unsigned(slots)), dim3(kSymwThreads), args, 0, s.stream);
    });
    if (err != cudaSuccess) return fail(c, NBODY_ECUDA, "sym_coop_step_kernel: %s", cudaGetErrorString(err));
    c->launches++;
    CU(c, cudaMemsetAsync(s.accfix + 3 * c->n_pad, 0, sizeof(unsigned long long), s.stream));  // the (sticky) poison word
    c->near_valid = false;
    c->sym32_used = true;
    return NBODY_OK;
}

int launch_medium_steps(nbody_ctx* c, double dt, double G, double eps, int nsteps)
{
    Shard& s = c->shards[0];
    CU(c, cudaSetDevice(s.device));
    c->near_valid = false;
    return c->prec == NBODY_F32 ? launch_medium_steps_t<float>(c, s, dt, G, eps, nsteps)
                                : launch_medium_steps_t<double>(c, s, dt, G, eps, nsteps);
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


}  // namespace

extern "C" int nbody_step(nbody_ctx* c, double dt, double G, double softening, int nsteps)
{
    DeviceGuard device_guard_;
    int rc = check_ready(c, "step");
    if (rc != NBODY_OK) return rc;
    if (nsteps < 0 || softening < 0.0) return fail(c, NBODY_EINVAL, "step: nsteps=%d softening=%g", nsteps, softening);
    if (nsteps == 0) return NBODY_OK;
    if (c->small_path && c->world == 1 && c->n <= c->small_max_n && !c->timing) return launch_small_steps(c, dt, G, softening, nsteps);
    if (c->medium_ctas > 0 && c->n <= c->medium_max_n && !c->timing) return launch_medium_steps(c, dt, G, softening, nsteps);
    rc = launch_sym_coop_steps(c, dt, G, softening, nsteps);
    if (rc <= 0) return rc;
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

extern "C" int nbody_step_host(nbody_ctx* c, size_t n, void* pos_mass, void* vel, void* acc, void* temp, double dt, double G, double softening,
                               int nsteps)
{
    DeviceGuard device_guard_;
    // The JS shim's frame: typed arrays in, nsteps x integrate(dt), typed arrays out.  For a small
    // system on one device this is one copy in, pack, the step kernel(s), unpack, one copy out and ONE synchronise.
    if (!c) return NBODY_EINVAL;
    if (!pos_mass || !vel || !acc) return fail(c, NBODY_EINVAL, "step_host: pos_mass, vel and acc are required (in and out)");
    int rc = nbody_upload(c, n, pos_mass, vel, acc, temp);
    if (rc != NBODY_OK) return rc;
    Shard& s = c->shards[0];
    // the page's own sizes: the cluster kernel (small.cuh) also writes the record the caller reads, so the frame is
    // count reset + pack + ONE step kernel + one synchronise
    const bool fused_out = s.io_host && c->upload_deferred && n <= kSmallIoZeroCopyMaxN && c->small_path && c->world == 1 &&
                           n <= c->small_max_n && !c->timing && nsteps > 0 && softening >= 0.0;
    if (!fused_out) {
        rc = nbody_step(c, dt, G, softening, nsteps);
        if (rc == NBODY_OK) rc = nbody_download(c, pos_mass, vel, acc, temp);
        return rc;
    }
    c->step_rec_out = s.io_host;
    rc = nbody_step(c, dt, G, softening, nsteps);
    c->step_rec_out = nullptr;
    if (rc != NBODY_OK) return rc;
    CU(c, cudaSetDevice(s.device));
    CU(c, cudaStreamSynchronize(s.stream));
    const size_t bytes4 = 4 * n * c->esz;
    const char* rec = static_cast<const char*>(s.io_host);
    memcpy(pos_mass, rec, bytes4);
    memcpy(vel, rec + bytes4, bytes4);
    memcpy(acc, rec + 2 * bytes4, bytes4);
    if (temp) memcpy(temp, rec + 3 * bytes4, n * c->esz);
    return check_exchange(c);
}

extern "C" int nbody_kick1_drift(nbody_ctx* c, double dt)
{
    DeviceGuard device_guard_;
    int rc = check_ready(c, "kick1_drift");
    if (rc != NBODY_OK) return rc;
    return launch_kick1_drift(c, dt);
}

extern "C" int nbody_force(nbody_ctx* c, double G, double softening)
{
    DeviceGuard device_guard_;
    int rc = check_ready(c, "force");
    if (rc != NBODY_OK) return rc;
    if (softening < 0.0) return fail(c, NBODY_EINVAL, "force: softening=%g", softening);
    return launch_force(c, kForceOnly, 0.0, G, softening);
}

extern "C" int nbody_kick2(nbody_ctx* c, double dt, double heat_decay)
{
    DeviceGuard device_guard_;
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
    DeviceGuard device_guard_;
    if (!c) return NBODY_EINVAL;
    int rc = fold_timing(c, c->shards[0]);
    c->timing = enable != 0;
    return rc;
}

extern "C" int nbody_read_timing(nbody_ctx* c, long long* launches, double* force_ms)
{
    DeviceGuard device_guard_;
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
    DeviceGuard device_guard_;
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

// Host logic only (no device needed): the pairs-once unit table for n bodies and the ranks' shares of it, for the CPU tests.
// units_out: (i-tile, first position tile, tile count, diagonal flag) per unit.
static int selftest_units(bool fine, size_t n, int bodies_per_thread, int run, int world, unsigned* units_out, size_t cap, size_t* count,
                          unsigned* split_out);

extern "C" int nbody_selftest_sym32_units(size_t n, int bodies_per_thread, int run, int world, unsigned* units_out /* 4*cap */, size_t cap,
                                          size_t* count, unsigned* split_out /* world + 1 */)
{
    return selftest_units(false, n, bodies_per_thread, run, world, units_out, cap, count, split_out);
}

extern "C" int nbody_selftest_symw_units(size_t n, int bodies_per_thread, int run, int world, unsigned* units_out /* 4*cap */, size_t cap,
                                         size_t* count, unsigned* split_out /* world + 1 */)
{
    return selftest_units(true, n, bodies_per_thread, run, world, units_out, cap, count, split_out);
}

static int selftest_units(bool fine, size_t n, int bodies_per_thread, int run, int world, unsigned* units_out, size_t cap, size_t* count,
                          unsigned* split_out)
{
    const bool r_ok = fine ? (bodies_per_thread == 2 || bodies_per_thread == 4 || bodies_per_thread == 8) : (bodies_per_thread == 4 || bodies_per_thread == 8);
    if (n == 0 || !r_ok || run < 1 || run > kSymUnitMaxRun || world < 1 || !count) return NBODY_EINVAL;
    std::vector<SymUnit> units;
    std::vector<unsigned> split;
    if (fine) make_symw_units(n, bodies_per_thread, run, world, units, split);
    else make_sym32_units(n, bodies_per_thread, run, world, units, split);
    *count = units.size();
    if (units.size() > cap) return NBODY_ENOMEM;
    for (size_t k = 0; units_out && k < units.size(); ++k) {
        units_out[4 * k] = units[k].tile;
        units_out[4 * k + 1] = units[k].run & 0xffffffu;
        units_out[4 * k + 2] = (units[k].run >> 24) & 0x7fu;
        units_out[4 * k + 3] = units[k].run >> 31;
    }
    for (int r = 0; split_out && r <= world; ++r) split_out[r] = split[r];
    return NBODY_OK;
}
