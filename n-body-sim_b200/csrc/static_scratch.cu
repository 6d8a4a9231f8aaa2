// nbody_accel_f32_static: fp32 accelerations of up to NBODY_STATIC_MAX_BODIES bodies WITHOUT any allocation
// (include/nbody.h).  No counterpart in the reference; added for callers that
// cannot call cudaMalloc -- code under CUDA graph capture, tight serving loops, harnesses that forbid it.
//
// The kernel is the warp-unit pairs-once force pass (force_f32_symw.cuh, general masses, 4 bodies per thread); the
// variant here takes ALL of its scratch -- the tile-SoA copy of the positions, the 64-bit fixed-point accumulators with
// their poison word, the unit counter, the unit table and the accumulator scale -- from ONE file-scope device array.
// Every GPU that loads the library holds that array (0.76 MiB) until the process exits.
//
// In host code `g_scratch` is a host-side stand-in, not a device address: its device address is looked up once per
// device with cudaGetSymbolAddress (that device current) and passed to the kernels.  Device code names it directly.
// Every call shares the array and finds the previous call's contents in it: the pack kernel clears what must start at
// zero.  Calls that overlap on the GPU (two streams, two host threads on one device) race on it: keep them ordered.
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <mutex>

#include "../../include/nbody.h"
#include "force_f32_symw.cuh"

namespace {
using namespace nb;

constexpr size_t kMaxN = NBODY_STATIC_MAX_BODIES;
static_assert(kMaxN % kPosTile == 0, "whole position tiles");
constexpr int kR = 4;                                    // bodies per thread: an i-block is two groups of 64
constexpr size_t kMaxGroups = kMaxN / kSymwGroup;
constexpr size_t kMaxBlocks = kMaxGroups / (kR / 2);
// units with one group each: per i-block one diagonal unit + the groups above it
constexpr size_t kMaxUnits = kMaxBlocks * (1 + kMaxGroups) - (kR / 2) * kMaxBlocks * (kMaxBlocks + 1) / 2;

// layout of the scratch array (every region 256-byte aligned)
constexpr size_t kOffHead = 0;                                            // StaticHead
constexpr size_t kOffAcc = 256;                                           // [3][n_pad] accumulators + the poison word
constexpr size_t kOffPos = kOffAcc + (3 * kMaxN + 32) * sizeof(unsigned long long);
constexpr size_t kOffUnits = kOffPos + 4 * kMaxN * sizeof(float);
constexpr size_t kScratchBytes = kOffUnits + ((kMaxUnits * sizeof(SymUnit) + 255) / 256) * 256;
static_assert(kOffPos % 256 == 0 && kOffUnits % 256 == 0, "aligned regions");
static_assert(kScratchBytes <= (size_t(1) << 20), "the static scratch stays modest");

struct StaticHead {
    unsigned sched;      // the unit counter of the persistent warps
    unsigned unused;
    float scale, bound;  // fixed-point scale 2^k and the bound on any partial sum, from sum |m| (set by static_pack_kernel)
    double inv_scale;
};

static __device__ __align__(256) unsigned char g_scratch[kScratchBytes];

// AoS (x, y, z, m) -> the tile-SoA layout of common.cuh, pads massless and far away; clears the accumulators, the poison
// word and the counter; adds |m| in a fixed order and derives the accumulator scale from it (launch_force_sym32's rule).
// One CTA: n <= 16,384.
__global__ void __launch_bounds__(1024) static_pack_kernel(const float4* __restrict__ pm, unsigned n, unsigned n_pad, float eps2)
{
    __shared__ double part[1024];
    float* pos = reinterpret_cast<float*>(g_scratch + kOffPos);
    unsigned long long* acc = reinterpret_cast<unsigned long long*>(g_scratch + kOffAcc);
    double mine = 0.0;
    for (unsigned i = threadIdx.x; i < n_pad; i += blockDim.x) {
        const float4 q = i < n ? pm[i] : make_float4(1e18f * float(1 + i - n), 0.f, 0.f, 0.f);
        pos[pm_index<float>(i, 0)] = q.x;
        pos[pm_index<float>(i, 1)] = q.y;
        pos[pm_index<float>(i, 2)] = q.z;
        pos[pm_index<float>(i, 3)] = q.w;
        mine += fabs(double(q.w));
        acc[i] = acc[n_pad + i] = acc[2 * size_t(n_pad) + i] = 0ull;
    }
    part[threadIdx.x] = mine;
    __syncthreads();
    for (unsigned s = 512; s > 0; s >>= 1) {
        if (threadIdx.x < s) part[threadIdx.x] += part[threadIdx.x + s];
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        StaticHead* h = reinterpret_cast<StaticHead*>(g_scratch + kOffHead);
        acc[3 * size_t(n_pad)] = 0ull;
        h->sched = 0u;
        // |sum| <= sum_j |m_j| * max_r r / (r^2 + eps^2)^1.5 = sum|m| * 2 / (3 sqrt(3) eps^2)
        const double bound = 1.001 * (2.0 / (3.0 * sqrt(3.0) * double(eps2))) * part[0];
        int e = 0;
        frexp(bound, &e);
        int k = 61 - e;
        k = k < -100 ? -100 : (k > 100 ? 100 : k);
        const bool ok = bound > 0.0 && bound < 1e37;
        h->scale = ok ? ldexpf(1.0f, k) : 0.0f;           // (all masses zero, or out of range: zero or poisoned sums)
        h->bound = ok ? nextafterf(float(bound), INFINITY) : (bound > 0.0 ? -1.0f : 0.0f);
        h->inv_scale = ldexp(1.0, -k);
    }
}

// The unit table of make_symw_units (nbody_capi.cu) for run = 1, built where it is used: one thread per i-block.
__global__ void static_units_kernel(unsigned n_groups)
{
    constexpr unsigned h = kR / 2;
    const unsigned n_blocks = (n_groups + h - 1) / h;
    const unsigned b = blockIdx.x * blockDim.x + threadIdx.x;
    SymUnit* units = reinterpret_cast<SymUnit*>(g_scratch + kOffUnits);
    if (b >= n_blocks) return;
    // blocks before b are complete: each has 1 + n_groups - (k + 1) h units
    size_t at = size_t(b) * (1 + n_groups) - size_t(h) * b * (b + 1) / 2;
    const unsigned s0 = b * h, s1 = min((b + 1) * h, n_groups);
    units[at++] = SymUnit{b, s0 | (s1 - s0) << 24 | kSymUnitDiag};
    for (unsigned g = s1; g < n_groups; ++g) units[at++] = SymUnit{b, g | 1u << 24};
}

// The variant of force_f32_symw_kernel<4, false>: the same body, every pointer into the static array, the accumulator
// scale read from where static_pack_kernel left it.
// `scratch` is the array's device address as the host looked it up (cudaGetSymbolAddress): the pointer a host-launched
// kernel must be given; the host-side name of the array is not one.
__global__ void __launch_bounds__(kSymwThreads, 3) force_f32_symw_static_kernel(unsigned char* scratch, const unsigned n_units, const unsigned n_pad,
                                                                               const float eps2)
{
    const StaticHead* h = reinterpret_cast<const StaticHead*>(scratch + kOffHead);
    Sym32Params p;
    p.pos = reinterpret_cast<const float*>(scratch + kOffPos);
    p.units = reinterpret_cast<const SymUnit*>(scratch + kOffUnits);
    p.u_begin = 0;
    p.u_end = n_units;
    p.sched = &reinterpret_cast<StaticHead*>(scratch + kOffHead)->sched;
    p.accfix = reinterpret_cast<unsigned long long*>(scratch + kOffAcc);
    p.n_pad = n_pad;
    p.eps2 = eps2;
    p.scale = h->scale;
    p.bound = h->bound;
    force_f32_symw_body<kR, false>(p);
}

// acc_i = G * sum (idx:280: f = G / ...): one rounding from the integer sums; a poisoned pass is NaN everywhere.
__global__ void static_finish_kernel(const unsigned char* __restrict__ scratch, float4* __restrict__ out, unsigned n, unsigned n_pad, float G)
{
    const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const StaticHead* h = reinterpret_cast<const StaticHead*>(scratch + kOffHead);
    const unsigned long long* acc = reinterpret_cast<const unsigned long long*>(scratch + kOffAcc);
    const size_t sl = fix_slot(i);
    float a[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) a[c] = float(double(static_cast<long long>(acc[c * size_t(n_pad) + sl])) * h->inv_scale) * G;
    if (acc[3 * size_t(n_pad)] != 0ull) a[0] = a[1] = a[2] = __int_as_float(0x7fc00000);
    out[i] = make_float4(a[0], a[1], a[2], 0.f);
}

struct PerDevice {
    unsigned char* scratch = nullptr;  // cudaGetSymbolAddress(g_scratch) with this device current
    int slots = 0;                     // resident CTAs of the force kernel: the persistent grid
};
constexpr int kMaxDevices = 64;
PerDevice g_dev[kMaxDevices];
std::mutex g_mutex;

int say(char* err, size_t cap, int code, const char* fmt, ...)
{
    if (err && cap) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(err, cap, fmt, ap);
        va_end(ap);
    }
    return code;
}

}  // namespace

extern "C" int nbody_accel_f32_static(const float* pos_mass_dev, float* acc_dev, size_t n, double G, double softening, void* stream, char* err,
                                      size_t err_cap)
{
    if (err && err_cap) err[0] = 0;
    if (n == 0) return NBODY_OK;
    if (n > kMaxN)
        return say(err, err_cap, NBODY_EINVAL, "nbody_accel_f32_static: n=%zu exceeds NBODY_STATIC_MAX_BODIES=%zu (the static scratch is sized for that)", n,
                   kMaxN);
    if (!pos_mass_dev || !acc_dev) return say(err, err_cap, NBODY_EINVAL, "nbody_accel_f32_static: pos_mass and acc (device pointers) are required");
    if (!(softening > 0.0) || !std::isfinite(softening) || !std::isfinite(G))
        return say(err, err_cap, NBODY_EINVAL, "nbody_accel_f32_static: softening=%g G=%g (softening > 0: the fixed-point bound needs it)", softening, G);
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices) return say(err, err_cap, NBODY_ECUDA, "nbody_accel_f32_static: cudaGetDevice: %s", cudaGetErrorString(e));
    PerDevice d;
    {
        std::lock_guard<std::mutex> lock(g_mutex);
        if (!g_dev[dev].scratch) {  // set-up, once per device, with that device current
            void* ptr = nullptr;
            e = cudaGetSymbolAddress(&ptr, g_scratch);
            if (e != cudaSuccess || !ptr) return say(err, err_cap, NBODY_ECUDA, "nbody_accel_f32_static: cudaGetSymbolAddress on device %d: %s", dev, cudaGetErrorString(e));
            int per_sm = 0, sms = 0;
            e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, force_f32_symw_static_kernel, kSymwThreads, 0);
            if (e == cudaSuccess) e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
            if (e != cudaSuccess || per_sm < 1 || sms < 1) return say(err, err_cap, NBODY_ECUDA, "nbody_accel_f32_static: occupancy query on device %d: %s", dev, cudaGetErrorString(e));
            g_dev[dev].slots = per_sm * sms;
            g_dev[dev].scratch = static_cast<unsigned char*>(ptr);
        }
        d = g_dev[dev];
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const unsigned nn = unsigned(n);
    const unsigned n_pad = unsigned((n + kPosTile - 1) / kPosTile * kPosTile);
    const unsigned NG = unsigned((n + kSymwGroup - 1) / kSymwGroup), h = kR / 2, NB = (NG + h - 1) / h;
    const unsigned n_units = NB * (1 + NG) - h * (NB - 1) * NB / 2 - NG;
    const float eps2 = float(softening) * float(softening);
    static_pack_kernel<<<1, 1024, 0, st>>>(reinterpret_cast<const float4*>(pos_mass_dev), nn, n_pad, eps2);
    static_units_kernel<<<(NB + 127) / 128, 128, 0, st>>>(NG);
    const unsigned grid = std::max(1u, std::min((n_units + kSymwWarps - 1) / kSymwWarps, unsigned(d.slots)));
    force_f32_symw_static_kernel<<<grid, kSymwThreads, 0, st>>>(d.scratch, n_units, n_pad, eps2);
    static_finish_kernel<<<(nn + 255) / 256, 256, 0, st>>>(d.scratch, reinterpret_cast<float4*>(acc_dev), nn, n_pad, float(G));
    e = cudaGetLastError();
    if (e != cudaSuccess) return say(err, err_cap, NBODY_ECUDA, "nbody_accel_f32_static: launch on device %d: %s", dev, cudaGetErrorString(e));
    return NBODY_OK;
}
