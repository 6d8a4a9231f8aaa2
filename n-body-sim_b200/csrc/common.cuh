// Shared definitions of the device-side state layout and kernel parameters.
//
// HBM layout (DESIGN.md "Data layout"):
//   fp32  pos_mass: TILE-SoA.  Tile t = bodies [512 t, 512 t + 512) is 8 KiB: four planes
//         x[512] y[512] z[512] m[512], so that a j-tile lands in shared memory by ONE
//         bulk copy (the equal-mass kernels skip the mass plane: 6 KiB).  Inside a plane
//         every group of 64 bodies is stored as 32 adjacent couples (b, b + 32):
//           slot(o) = (o & ~63) | ((o & 31) << 1) | ((o >> 5) & 1),   o = body & 511
//         Any two adjacent floats are two j-bodies of one aligned 64-bit register pair
//         for the packed FADD2/FFMA2/FMUL2 pipe (one broadcast LDS.128 per plane = two
//         couples, force_f32.cuh), and in the pairs-once kernel (force_f32_sym.cuh) lane l
//         of a warp fetches the couple (l + q) & 31 of a group with one conflict-free
//         LDS.64 per plane while the couple's sums rotate through the warp.
//   fp64  pos_mass: double4 (x, y, z, m) per body.
//   vel, acc: Real4 per body (x, y, z, unused), indexed from the shard's i_begin.
//   partial : Real4 [j_chunks][local bodies]; per-chunk acceleration sums.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

// NB_DEBUG_CHECKS (make variant NAME=checked EXTRA=-DNB_DEBUG_CHECKS): device-side assert()s on the indices the hand-rolled
// machinery computes.  compute-sanitizer is closed on the build pool (profiles/r2_sanitizer_closed.txt); an out-of-range index
// then traps in the checked build instead of corrupting memory.  The release build compiles them away.
#ifdef NB_DEBUG_CHECKS
#include <cassert>
#define NB_CHECK(cond) assert(cond)
#else
#define NB_CHECK(cond) ((void)0)
#endif

namespace nb {

// fp64 kernel: 8 compute warps + 1 producer warp, one i-body per thread
constexpr int kThreads = 256;       // compute threads per CTA
#ifndef NB_F64_STAGES
#define NB_F64_STAGES 4
#endif
#ifndef NB_F64_TILE_BYTES
#define NB_F64_TILE_BYTES 16384
#endif
constexpr int kStages = NB_F64_STAGES;          // j-tile ring depth
constexpr int kTileBytes = NB_F64_TILE_BYTES;   // bytes per j-tile stage (512 fp64 bodies)
// fp32 kernel (shape chosen by the force_lab sweeps, profiles/r1_force_lab_sweep*.txt):
// 4 compute warps + 1 producer warp, R i-bodies per thread, 3 CTAs per SM at 128 registers
constexpr int kF32Threads = 128;    // compute threads per CTA
constexpr int kF32TilePairs = 256;  // j-PAIRS per ring stage (8 KiB, 512 bodies)
constexpr int kF32Stages = 4;
// The j extent and every j-chunk are multiples of the kernel's j unit.  fp32: one ring
// stage (tiles are always full).  fp64: 64 bodies -- its ring takes partial tiles, and the
// reference's own N=151 scene then splits into several CTAs instead of one.
constexpr int kJUnit = 512;
constexpr int kJUnitF64 = 64;
// Peer-store exchange (exchange.cuh): a position update is stored straight into every
// rank's copy of the next position buffer over NVLink; at most this many ranks.
constexpr int kMaxPeers = 8;

// Destinations of one position update: the `other` position buffer of this rank and,
// in peer-store mode, of every peer (peer-mapped or IPC-opened device pointers).
template <typename Real> struct PosOut {
    Real* dst[kMaxPeers];
    int n;
};

// Epoch flags of the peer-store exchange barrier (exchange.cuh).
struct PeerFlags {
    unsigned long long* flags[kMaxPeers];  // flags[r] = rank r's flag array (kMaxPeers entries), as mapped here
};

template <typename Real> struct Vec4;
template <> struct Vec4<float> { using type = float4; };
template <> struct Vec4<double> { using type = double4; };

// Scalar element index of component c (0=x,1=y,2=z,3=m) of body i in pos_mass.
template <typename Real> __host__ __device__ __forceinline__ size_t pm_index(size_t i, int c);
constexpr int kPosTile = 512;  // bodies per fp32 position tile (4 planes of kPosTile floats)
// slot of body o (0..511) inside a plane of its tile, and back
__host__ __device__ __forceinline__ unsigned tile_slot(unsigned o) { return (o & ~63u) | ((o & 31u) << 1) | ((o >> 5) & 1u); }
__host__ __device__ __forceinline__ unsigned tile_body(unsigned s) { return (s & ~63u) | ((s & 1u) << 5) | ((s >> 1) & 31u); }
template <> __host__ __device__ __forceinline__ size_t pm_index<float>(size_t i, int c)
{
    return (i >> 9) * size_t(4 * kPosTile) + size_t(c) * kPosTile + tile_slot(unsigned(i) & 511u);
}
template <> __host__ __device__ __forceinline__ size_t pm_index<double>(size_t i, int c)
{
    return 4 * i + c;
}

// Equal-mass sweep with a few exceptions (force_f32.cuh, UNIFORM): body `index` carries
// `dm` = its mass minus the common mass, added per i-body after the unit-mass sums.
struct MassOutlier {
    unsigned index;
    float dm;
};
constexpr int kMaxOutliers = 64;

enum EpilogueMode : int {
    kForceOnly = 0,        // acc = G * sum                                   (idx:264-291)
    kForceKick2 = 1,       // + vel += acc*h, thermal                         (idx:258-261)
    kForceKick2Kick1 = 2   // + next step's vel += acc*h; pos += vel*dt       (idx:250-253)
};

template <typename Real> struct ForceParams {
    const Real* pos_cur;    // full padded position array (all shards' bodies)
    PosOut<Real> out;       // written in mode kForceKick2Kick1 (own range, on every rank in peer-store mode)
    Real* vel;              // local
    Real* acc;              // local
    Real* partial;          // [j_chunks][n_local] Real4
    Real* temp;             // local, may be null
    Real* near;             // local; NEAR kernels only: max_j!=i rsqrt(d^2+eps2) (contacts.cuh)
    const MassOutlier* outliers;  // UNIFORM fp32 kernels only: the bodies whose mass is not the common one
    int n_outliers;         // 0: `G` already carries the common mass
    Real bulk_mass;         // n_outliers > 0: the common mass (G stays the plain constant)
    unsigned* tile_counter; // one per local i-tile; zero between launches
    size_t n_j;             // j extent: n rounded up to kJUnit (independent of the GPU count)
    size_t i_begin;         // first global body of this shard (multiple of i-tile)
    size_t n_local;         // padded local body count (multiple of i-tile)
    size_t chunk_len;       // bodies per j-chunk (multiple of kJUnit)
    int j_chunks;
    int mode;
    Real G, eps2, half_dt, dt, heat_decay;  // heat_decay <= 0: no thermal update
};

}  // namespace nb
