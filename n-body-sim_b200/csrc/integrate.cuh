// O(N) kernels: the stand-alone halves of PhysicsSystem.integrate
// (/root/reference/index.html:250-253 and 258-261), layout conversion for
// upload/download, and the fp64 energy/momentum diagnostic (not in the
// reference).
#pragma once
#include "common.cuh"
#include "epilogue.cuh"

namespace nb {

// vel += acc*halfDt ; pos += vel*dt  (idx:250-253), bodies [i_begin, i_begin+n_local).
// Out of place: reads pos_in, writes the updated positions to every destination in `out`
// (this rank's other buffer and, in peer-store mode, every peer's).
template <typename Real>
__global__ void kick1_drift_kernel(const Real* pos_in, PosOut<Real> out, Real* vel, const Real* acc, size_t i_begin,
                                   size_t n_local, Real half_dt, Real dt)
{
    using V4 = typename Vec4<Real>::type;
    const size_t il = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (il >= n_local) return;
    V4 v = reinterpret_cast<V4*>(vel)[il];
    const V4 a = reinterpret_cast<const V4*>(acc)[il];
    v.x = add_rn(v.x, mul_rn(a.x, half_dt));
    v.y = add_rn(v.y, mul_rn(a.y, half_dt));
    v.z = add_rn(v.z, mul_rn(a.z, half_dt));
    reinterpret_cast<V4*>(vel)[il] = v;
    const size_t g = i_begin + il;
    const size_t ix = pm_index<Real>(g, 0), iy = pm_index<Real>(g, 1), iz = pm_index<Real>(g, 2);
    const Real nx = add_rn(pos_in[ix], mul_rn(v.x, dt));
    const Real ny = add_rn(pos_in[iy], mul_rn(v.y, dt));
    const Real nz = add_rn(pos_in[iz], mul_rn(v.z, dt));
    for (int k = 0; k < out.n; ++k) {
        Real* o = out.dst[k];
        o[ix] = nx;
        o[iy] = ny;
        o[iz] = nz;
    }
}

// vel += acc*halfDt ; updateThermal  (idx:258-261)
template <typename Real>
__global__ void kick2_kernel(Real* vel, const Real* acc, Real* temp, size_t n_local, Real half_dt,
                             Real heat_decay)
{
    using V4 = typename Vec4<Real>::type;
    const size_t il = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (il >= n_local) return;
    V4 v = reinterpret_cast<V4*>(vel)[il];
    const V4 a = reinterpret_cast<const V4*>(acc)[il];
    v.x = add_rn(v.x, mul_rn(a.x, half_dt));
    v.y = add_rn(v.y, mul_rn(a.y, half_dt));
    v.z = add_rn(v.z, mul_rn(a.z, half_dt));
    reinterpret_cast<V4*>(vel)[il] = v;
    if (temp != nullptr && heat_decay > Real(0)) temp[il] = thermal(temp[il], heat_decay);
}

// Host AoS (x,y,z,m)[n] staged on the device -> device pos_mass layout.  Pads are massless
// bodies parked far away at x = pad_x (1e18; 1e150 in fp64): the equal-mass kernels ignore masses, and a
// pad at the origin would meet a real body there as 0 * inf when softening is 0.  In fp64 pad k sits at
// pad_x * (1 + k), every pad somewhere else: two pads at one point are 0 * inf for each other when softening
// is 0 (only the self pair is selected away), and a NaN acceleration would move the pad onto the bodies'
// sums at the next fused step.  (fp32 clamps r^2 instead: force_f32.cuh, ZERO_SOFT.)
// detect != nullptr (fp32): also find the bodies whose mass is not `m0` -- detect[0] counts them
// (a non-finite mass pushes the count past any limit) and the first kMaxOutliers found are
// appended to `outliers` in no particular order; the host sorts them (nbody_upload).
// mass_sums != nullptr: block b also leaves sum |m| of its 256 bodies in mass_sums[b], added in a fixed order (the host
// adds the blocks in index order): the bound behind the fixed-point scale of the pairs-once pass (force_f32_sym.cuh)
// must be the same number on every rank and in every run.
template <typename Real>
__global__ void pack_pos_kernel(const Real* aos, Real* pos_a, Real* pos_b, size_t n, size_t n_pad, Real pad_x, Real m0,
                                unsigned* detect, MassOutlier* outliers, double* mass_sums = nullptr)
{
    const size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    double am = 0.0;
    if (i < n_pad) {
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const Real v = i < n ? aos[4 * i + c] : (c == 0 ? (sizeof(Real) == 8 ? pad_x * Real(1 + (i - n)) : pad_x) : Real(0));
            pos_a[pm_index<Real>(i, c)] = v;
            pos_b[pm_index<Real>(i, c)] = v;  // mass is constant: both buffers carry it
            if (c == 3) am = fabs(double(v));
            if (c == 3 && detect != nullptr && i < n && !(v == m0)) {
                const bool finite = v - v == Real(0);
                const unsigned k = atomicAdd(detect, finite ? 1u : unsigned(kMaxOutliers) + 1u);
                if (finite && k < unsigned(kMaxOutliers)) outliers[k] = MassOutlier{unsigned(i), float(v - m0)};
            }
        }
    }
    if (mass_sums == nullptr) return;  // (uniform over the grid)
    __shared__ double wsum[8];
    for (int o = 16; o > 0; o >>= 1) am += __shfl_xor_sync(0xffffffffu, am, o);
    if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = am;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < int(blockDim.x >> 5); ++w) t += wsum[w];
        mass_sums[blockIdx.x] = t;
    }
}

// Several devices with peer mappings: every shard packs ITS OWN slice [i_begin, i_begin + n_local) of the host array -- `aos`
// holds that slice only -- into BOTH position buffers of EVERY rank (peer stores), so the host->device traffic of an upload
// is one copy of the array in total instead of one per device.  Same pads, same mass detection, same block sums as above.
template <typename Real> struct PackDst {
    Real* p[2 * kMaxPeers];
    int n;
};
template <typename Real>
__global__ void pack_pos_shard_kernel(const Real* aos, PackDst<Real> dst, size_t i_begin, size_t n_local, size_t n, Real pad_x, Real m0,
                                      unsigned* detect, MassOutlier* outliers, double* mass_sums)
{
    const size_t il = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    const size_t i = i_begin + il;
    double am = 0.0;
    if (il < n_local) {
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const Real v = i < n ? aos[4 * il + c] : (c == 0 ? (sizeof(Real) == 8 ? pad_x * Real(1 + (i - n)) : pad_x) : Real(0));
            const size_t at = pm_index<Real>(i, c);
            for (int k = 0; k < dst.n; ++k) dst.p[k][at] = v;
            if (c == 3) am = fabs(double(v));
            if (c == 3 && detect != nullptr && i < n && !(v == m0)) {
                const bool finite = v - v == Real(0);
                const unsigned k = atomicAdd(detect, finite ? 1u : unsigned(kMaxOutliers) + 1u);
                if (finite && k < unsigned(kMaxOutliers)) outliers[k] = MassOutlier{unsigned(i), float(v - m0)};
            }
        }
    }
    if (mass_sums == nullptr) return;
    __shared__ double wsum[8];
    for (int o = 16; o > 0; o >>= 1) am += __shfl_xor_sync(0xffffffffu, am, o);
    if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = am;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < int(blockDim.x >> 5); ++w) t += wsum[w];
        mass_sums[blockIdx.x] = t;
    }
}

// one shard's own range of the device positions -> the same range of the host AoS (nbody_download_own)
template <typename Real>
__global__ void unpack_pos_range_kernel(const Real* pos, Real* aos, size_t i_begin, size_t count)
{
    const size_t k = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (k >= count) return;
#pragma unroll
    for (int c = 0; c < 4; ++c) aos[4 * k + c] = pos[pm_index<Real>(i_begin + k, c)];
}

// Small systems (nbody_upload / nbody_download with at most kSmallIoMaxN bodies on one device): the host arrays travel as
// ONE record [pos_mass 4n | vel 4n | acc 4n | temp n] through pinned memory, so a call is one copy + one kernel instead of
// four copies, three memsets and a kernel -- the per-frame flow of the JS shim (upload, step, download every frame at the
// reference's N <= 300, idx:57) is bound by exactly those driver calls.  pack_state_kernel = pack_pos_kernel + the local
// arrays with their pads zeroed; unpack_state_kernel is its inverse.
constexpr size_t kSmallIoMaxN = 4096;
constexpr size_t kSmallIoZeroCopyMaxN = 1024;  // up to here the two kernels read / write the pinned record in place (<= 104 KiB over PCIe)

template <typename Real> __host__ __device__ constexpr size_t small_io_elems(size_t n) { return 13 * n; }

template <typename Real>
__global__ void pack_state_kernel(const Real* rec, Real* pos_a, Real* pos_b, Real* vel, Real* acc, Real* temp, size_t n, size_t n_pad,
                                  Real pad_x, Real m0, unsigned* detect, MassOutlier* outliers)
{
    const size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n_pad) return;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        const Real v = i < n ? rec[4 * i + c] : (c == 0 ? (sizeof(Real) == 8 ? pad_x * Real(1 + (i - n)) : pad_x) : Real(0));
        pos_a[pm_index<Real>(i, c)] = v;
        pos_b[pm_index<Real>(i, c)] = v;
        if (c == 3 && detect != nullptr && i < n && !(v == m0)) {
            const bool finite = v - v == Real(0);
            const unsigned k = atomicAdd(detect, finite ? 1u : unsigned(kMaxOutliers) + 1u);
            if (finite && k < unsigned(kMaxOutliers)) outliers[k] = MassOutlier{unsigned(i), float(v - m0)};
        }
        vel[4 * i + c] = i < n ? rec[4 * n + 4 * i + c] : Real(0);
        acc[4 * i + c] = i < n ? rec[8 * n + 4 * i + c] : Real(0);
    }
    temp[i] = i < n ? rec[12 * n + i] : Real(0);
}

template <typename Real>
__global__ void unpack_state_kernel(const Real* pos, const Real* vel, const Real* acc, const Real* temp, Real* rec, size_t n)
{
    const size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        rec[4 * i + c] = pos[pm_index<Real>(i, c)];
        rec[4 * n + 4 * i + c] = vel[4 * i + c];
        rec[8 * n + 4 * i + c] = acc[4 * i + c];
    }
    rec[12 * n + i] = temp[i];
}

template <typename Real>
__global__ void unpack_pos_kernel(const Real* pos, Real* aos, size_t n)
{
    const size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
#pragma unroll
    for (int c = 0; c < 4; ++c) aos[4 * i + c] = pos[pm_index<Real>(i, c)];
}

// ---- energy / momentum, fp64 accumulation whatever the state precision ------
// out[block*5 + {0..4}] = {kinetic, potential(pair sum, unscaled by -G), px, py, pz}
// for the i-bodies of this block against all n bodies; the host adds the block
// rows in index order.
constexpr int kEnergyThreads = 256;

template <typename Real>
__global__ void __launch_bounds__(kEnergyThreads)
    energy_kernel(const Real* pos, const Real* vel, size_t n, size_t i_begin, size_t n_local_real,
                  double eps2, double* out)
{
    using V4 = typename Vec4<Real>::type;
    __shared__ double4 tile[kEnergyThreads];
    __shared__ double red[5][kEnergyThreads / 32];
    const int tid = threadIdx.x;
    const size_t il = size_t(blockIdx.x) * kEnergyThreads + tid;
    const bool live = il < n_local_real;
    const size_t gi = i_begin + il;
    double xi = 0, yi = 0, zi = 0, mi = 0;
    double ke = 0, px = 0, py = 0, pz = 0, pot = 0;
    if (live) {
        xi = pos[pm_index<Real>(gi, 0)];
        yi = pos[pm_index<Real>(gi, 1)];
        zi = pos[pm_index<Real>(gi, 2)];
        mi = pos[pm_index<Real>(gi, 3)];
        const V4 v = reinterpret_cast<const V4*>(vel)[il];
        const double vx = v.x, vy = v.y, vz = v.z;
        ke = 0.5 * mi * (vx * vx + vy * vy + vz * vz);
        px = mi * vx;
        py = mi * vy;
        pz = mi * vz;
    }
    for (size_t jb = 0; jb < n; jb += kEnergyThreads) {
        const size_t j = jb + tid;
        double4 b = make_double4(0, 0, 0, 0);
        if (j < n)
            b = make_double4(pos[pm_index<Real>(j, 0)], pos[pm_index<Real>(j, 1)],
                             pos[pm_index<Real>(j, 2)], pos[pm_index<Real>(j, 3)]);
        __syncthreads();
        tile[tid] = b;
        __syncthreads();
        const int cnt = static_cast<int>(n - jb < kEnergyThreads ? n - jb : kEnergyThreads);
#pragma unroll 4
        for (int q = 0; q < cnt; ++q) {
            const double4 o = tile[q];
            const double dx = o.x - xi, dy = o.y - yi, dz = o.z - zi;
            const double r2 = dx * dx + dy * dy + dz * dz + eps2;
            const double term = o.w * rsqrt(r2);
            pot += (jb + q == gi) ? 0.0 : term;
        }
    }
    pot *= 0.5 * mi;  // each unordered pair is visited twice
    double vals[5] = {ke, live ? pot : 0.0, px, py, pz};
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        double v = vals[k];
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if ((tid & 31) == 0) red[k][tid >> 5] = v;
    }
    __syncthreads();
    if (tid < 5) {
        double s = 0;
        for (int w = 0; w < kEnergyThreads / 32; ++w) s += red[tid][w];
        out[size_t(blockIdx.x) * 5 + tid] = s;
    }
}

}  // namespace nb
