// Shared definitions of the device-side state layout and kernel parameters.
//
// HBM layout (DESIGN.md "Data layout"):
//   fp32  pos_mass: PAIR-INTERLEAVED.  Bodies 2p and 2p+1 share 32 bytes:
//           float4 A_p = (x0, x1, y0, y1), float4 B_p = (z0, z1, m0, m1)
//         so that a j-tile lands in shared memory by one bulk copy and one
//         LDS.128 yields two aligned 64-bit register pairs that feed the packed
//         FADD2/FFMA2/FMUL2 pipe directly (two j-bodies per instruction).
//   fp64  pos_mass: double4 (x, y, z, m) per body.
//   vel, acc: Real4 per body (x, y, z, unused), indexed from the shard's i_begin.
//   partial : Real4 [j_chunks][local bodies]; per-chunk acceleration sums.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

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
template <> __host__ __device__ __forceinline__ size_t pm_index<float>(size_t i, int c)
{
    return 8 * (i >> 1) + 2 * c + (i & 1);
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
