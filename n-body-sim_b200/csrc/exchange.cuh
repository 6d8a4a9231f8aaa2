// Peer-store position exchange over NVLink / NVSwitch (SURVEY.md 8e "B200-native
// form"): the kernels that update positions (the fused force epilogue and
// kick1_drift) store each new position directly into EVERY rank's copy of the next
// position buffer (PosOut, peer-mapped or IPC-opened pointers), so the all-gather
// is fused into the compute kernel and overlaps it CTA by CTA.  What remains is one
// cross-GPU epoch barrier per update, below: a 1-warp kernel queued behind the
// update on every rank's stream.
//
// Protocol.  flags[r] on rank me = last epoch rank r has published to me.  After the
// update kernel of epoch e (stream order => its peer stores have completed), lane r
// release-stores e into rank r's flags[me], then spins until its own flags[r] >= e.
// When the barrier kernel retires, every rank's slice of the next buffer has landed
// here, and every rank has finished READING the buffer that the following update
// will overwrite (double buffering: update e+1 writes the buffer update e read).
// Each rank runs on its own GPU, so the spin never waits on work queued behind it.
// A wall-clock timeout (globaltimer) turns a lost peer into an error flag instead of
// a hung GPU.
#pragma once
#include <cstdint>

#include "common.cuh"

namespace nb {

__device__ __forceinline__ unsigned long long globaltimer_ns()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

__global__ void exchange_barrier_kernel(PeerFlags pf, int me, int world, unsigned long long epoch, unsigned long long timeout_ns,
                                        unsigned* err)
{
    const int r = threadIdx.x;
    if (r >= world) return;
    // publish: my update of this epoch is complete (kernel boundary) -> tell rank r
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(pf.flags[r] + me), "l"(epoch) : "memory");
    // wait: rank r's update of this epoch has landed in my buffers
    const unsigned long long* mine = pf.flags[me] + r;
    const unsigned long long t0 = globaltimer_ns();
    unsigned long long seen;
    for (;;) {
        asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(seen) : "l"(mine) : "memory");
        if (seen >= epoch) break;
        if (globaltimer_ns() - t0 > timeout_ns) {
            atomicExch(err, 1u + unsigned(r));
            break;
        }
        __nanosleep(200);
    }
}

}  // namespace nb
