// xchg.cuh — one small exchange over peer memory, as a device function a whole CTA executes.
//
// The sharded resample needs a few tiny all-gathers (<= 4 KB per rank) and barriers per update.  Through NCCL each costs
// a launch plus ~20 us of protocol at 8 devices; here ONE CTA per rank STORES its chunk straight into every peer's copy
// of the array over NVLink (the buffers are CUDA-IPC mapped), publishes an arrival counter in every peer's flag array
// behind a system-scope fence, and spins on its own flags until all peers have arrived: the stores of a peer are
// visible once its counter is.  Every rank issues the same sequence of exchanges, so the counter is simply the exchange
// number.  The wait is bounded (~20 s of SM clocks): a missing peer raises *error (sticky; the host turns it into
// SLAMGPU_ECOMM at its next synchronisation) instead of hanging the device.
//
// Called either by the stand-alone kernel k_xchg (comm.cu: barriers, the tail record) or from INSIDE the cooperative
// weights kernel (kernels.cuh k_weights_fused), where CTA 0 runs it between the arrival and the release of a grid-wide
// barrier — compute and exchange in one launch.
#pragma once

#include "types.cuh"

namespace slam {

__device__ __forceinline__ void xchg_body(const XchgArgs &a) {
    const int tid = threadIdx.x, nt = blockDim.x;
    for (int k = 0; k < a.narr; k++) {
        const float *src = a.slab[a.rank] + a.arr_off[k] + a.b0;
        for (int r = 0; r < a.world; r++) {
            if (r == a.rank) continue;
            float *dst = a.slab[r] + a.arr_off[k] + a.b0;
            for (int i = tid; i < a.count; i += nt) dst[i] = __ldcg(src + i);
        }
    }
    if (a.blob_words && tid < a.blob_words * a.world) {
        const int r = tid / a.blob_words, k = tid % a.blob_words;
        reinterpret_cast<unsigned *>(a.slab[r])[a.blob_off + a.rank * a.blob_words + k] = __ldcg(a.blob_src + k);
    }
    if (a.halo && a.rank + 1 < a.world && tid < 2)  // the next rank's search needs this rank's last two uniforms
        a.unif[a.rank + 1][tid] = a.unif[a.rank][a.P + tid];
    __threadfence_system();
    __syncthreads();
    if (tid < a.world) {
        __threadfence_system();  // cumulative: the whole CTA's stores (ordered by the barrier) precede the counter
        volatile unsigned long long *theirs = a.flags[tid] + a.rank;
        *theirs = a.epoch;
        volatile unsigned long long *mine = a.flags[a.rank] + tid;
        const long long t0 = clock64();
        while (*mine < a.epoch) {
            if (clock64() - t0 > 40000000000ll) {
                *a.error = 1;
                break;
            }
        }
    }
    __threadfence_system();
    __syncthreads();
}

}  // namespace slam
