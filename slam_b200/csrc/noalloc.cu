// noalloc.cu — slamgpu_estimate_noalloc (include/slamgpu.h): the pose estimate of k_estimate (kernels.cuh; K8,
// ParticleSLAMWrapper.cpp:56-77) for callers that may not allocate: code under CUDA graph capture, tight serving
// loops, harnesses that forbid cudaMalloc.  No counterpart in the
// reference.
//
// k_estimate needs scratch memory (one partial per CTA and the ticket that elects the combining CTA), which a handle
// gets from cudaMalloc at slamgpu_create.  Here that scratch is a file-scope __device__ array: every GPU that loads
// the library holds it (sizeof(NaScratch) = 33 KB) until the process exits.  Its size bounds the call:
// SLAMGPU_NOALLOC_MAX_PARTICLES (2^22) particles, one 32-byte partial per NA_CHUNK = 4096 of them.
//
// In host code the array's NAME is only a stand-in (the address of a host shadow symbol), not a device address: a
// kernel handed it would fault or, where the GPU can reach host memory, quietly use the wrong memory.  The device
// address comes from cudaGetSymbolAddress, once per device with that device current (scratch_for_current_device).
//
// Every call on a device shares the one array and finds the previous call's contents in it.  The ticket is cleared by
// a stream-ordered memset in front of every launch; the partials are written before they are read.  Two calls that
// overlap on the GPU (different streams, no dependency between them) race on the array: callers serialise them.
#include "../../include/slamgpu.h"

#include <cuda_runtime.h>

#include <math.h>
#include <mutex>
#include <string>

#include "handle.h"

namespace {

constexpr int NA_THREADS = 256;
constexpr long long NA_CHUNK = 4096;  // particles per CTA: 16 per thread
constexpr long long NA_MAX_PARTICLES = SLAMGPU_NOALLOC_MAX_PARTICLES;
constexpr int NA_MAX_CTAS = (int) (NA_MAX_PARTICLES / NA_CHUNK);  // 1024

struct NaPartial {  // the EstPartial of types.cuh: double sums, the first maximum weight and where it is
    double sx, sy;
    float wmax;
    float pad;
    long long imax;
};
struct NaScratch {
    unsigned ticket;             // CTAs that have written their partial in the launch in flight
    unsigned pad[63];            // the partials start on their own 256-byte line
    NaPartial part[NA_MAX_CTAS];
};
static_assert(sizeof(NaScratch) == 256 + 32 * NA_MAX_CTAS, "scratch layout");
static_assert(NA_MAX_PARTICLES % NA_CHUNK == 0, "whole chunks");

// The scratch.  Host code must not pass `g_na_scratch` to a kernel: see the header comment.
static __device__ __align__(256) unsigned char g_na_scratch[sizeof(NaScratch)];

__device__ __forceinline__ void na_combine(NaPartial &a, const NaPartial &b) {
    a.sx += b.sx;
    a.sy += b.sy;
    if (b.wmax > a.wmax || (b.wmax == a.wmax && b.imax < a.imax)) {  // the FIRST maximum wins
        a.wmax = b.wmax;
        a.imax = b.imax;
    }
}
__device__ __forceinline__ NaPartial na_empty() {
    NaPartial v;
    v.sx = 0.0;
    v.sy = 0.0;
    v.wmax = -INFINITY;
    v.pad = 0.0f;
    v.imax = 0x7fffffffffffffffLL;
    return v;
}
__device__ __forceinline__ NaPartial na_block_reduce(NaPartial v) {
    __shared__ NaPartial sh[NA_THREADS / 32];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        NaPartial o;
        o.sx = __shfl_down_sync(0xffffffffu, v.sx, d);
        o.sy = __shfl_down_sync(0xffffffffu, v.sy, d);
        o.wmax = __shfl_down_sync(0xffffffffu, v.wmax, d);
        o.imax = __shfl_down_sync(0xffffffffu, v.imax, d);
        na_combine(v, o);
    }
    __syncthreads();  // sh may still be read by the previous reduction of this CTA
    if (lane == 0) sh[wid] = v;
    __syncthreads();
    if (wid == 0) {
        v = lane < NA_THREADS / 32 ? sh[lane] : na_empty();
#pragma unroll
        for (int d = NA_THREADS / 64; d > 0; d >>= 1) {
            NaPartial o;
            o.sx = __shfl_down_sync(0xffffffffu, v.sx, d);
            o.sy = __shfl_down_sync(0xffffffffu, v.sy, d);
            o.wmax = __shfl_down_sync(0xffffffffu, v.wmax, d);
            o.imax = __shfl_down_sync(0xffffffffu, v.imax, d);
            na_combine(v, o);
        }
    }
    return v;  // valid in thread 0
}

// The variant of k_estimate: CTA b reduces particles [b * NA_CHUNK, (b + 1) * NA_CHUNK) to part[b]; the CTA that
// draws the last ticket combines the partials in index order (the same order whichever CTA that is) and writes
// out[0..3] = mean x, mean y, heading of the first maximum-weight particle, that weight.
__global__ void __launch_bounds__(NA_THREADS) k_estimate_noalloc(const float *__restrict__ xv, const float *__restrict__ w,
                                                                 long long P, NaScratch *sc, double *__restrict__ out) {
    __shared__ bool last;
    NaPartial v = na_empty();
    const long long base = (long long) blockIdx.x * NA_CHUNK;
    const long long end = base + NA_CHUNK < P ? base + NA_CHUNK : P;
    for (long long p = base + threadIdx.x; p < end; p += NA_THREADS) {
        const float wp = w[p];
        v.sx += (double) xv[p];
        v.sy += (double) xv[P + p];
        if (wp > v.wmax) {  // within a thread the indices only grow: the first maximum wins by itself
            v.wmax = wp;
            v.imax = p;
        }
    }
    v = na_block_reduce(v);
    if (threadIdx.x == 0) {
        sc->part[blockIdx.x] = v;
        __threadfence();
        last = atomicAdd(&sc->ticket, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!last) return;
    __threadfence();
    v = na_empty();
    for (int i = threadIdx.x; i < (int) gridDim.x; i += NA_THREADS) {
        const volatile NaPartial *q = sc->part + i;  // written by other CTAs of this launch: no cached / hoisted reads
        NaPartial o;
        o.sx = q->sx;
        o.sy = q->sy;
        o.wmax = q->wmax;
        o.imax = q->imax;
        na_combine(v, o);
    }
    v = na_block_reduce(v);
    if (threadIdx.x == 0) {
        out[0] = v.sx / (double) P;
        out[1] = v.sy / (double) P;
        out[2] = (v.imax >= 0 && v.imax < P) ? (double) xv[2 * P + v.imax] : 0.0;
        out[3] = (double) v.wmax;
        sc->ticket = 0;
    }
}

// The device address of g_na_scratch on the CURRENT device, resolved once per device.  cudaGetSymbolAddress allocates
// nothing: the array came with the module.
constexpr int NA_MAX_DEVICES = 64;
std::mutex g_na_mutex;
NaScratch *g_na_ptr[NA_MAX_DEVICES] = {nullptr};

int scratch_for_current_device(int dev, NaScratch **out) {
    std::lock_guard<std::mutex> lock(g_na_mutex);
    if (!g_na_ptr[dev]) {
        void *ptr = nullptr;
        CU(cudaGetSymbolAddress(&ptr, g_na_scratch));
        if (!ptr) return slam_fail(SLAMGPU_ECUDA, "cudaGetSymbolAddress returned a null scratch address");
        g_na_ptr[dev] = static_cast<NaScratch *>(ptr);
    }
    *out = g_na_ptr[dev];
    return SLAMGPU_OK;
}

// Device (or managed) memory of the current device?  A host pointer would make the kernel fault, or read host memory.
int check_device_pointer(const void *p, int dev, const char *what) {
    cudaPointerAttributes at;
    CU(cudaPointerGetAttributes(&at, p));
    if (at.type == cudaMemoryTypeManaged) return SLAMGPU_OK;
    if (at.type != cudaMemoryTypeDevice)
        return slam_fail(SLAMGPU_EINVAL, std::string("slamgpu_estimate_noalloc: ") + what + " is not device memory");
    if (at.device != dev)
        return slam_fail(SLAMGPU_EINVAL, std::string("slamgpu_estimate_noalloc: ") + what +
                                             " belongs to device " + std::to_string(at.device) + ", the current device is " +
                                             std::to_string(dev));
    return SLAMGPU_OK;
}

}  // namespace

extern "C" int slamgpu_estimate_noalloc(const float *d_xv, const float *d_w, int64_t n_particles, double *d_out,
                                        void *cuda_stream) {
    if (!d_xv || !d_w || !d_out) return slam_fail(SLAMGPU_EINVAL, "slamgpu_estimate_noalloc: null pointer");
    if (n_particles <= 0) return slam_fail(SLAMGPU_EINVAL, "slamgpu_estimate_noalloc: n_particles must be positive");
    if (n_particles > NA_MAX_PARTICLES)
        return slam_fail(SLAMGPU_EINVAL, "slamgpu_estimate_noalloc: n_particles " + std::to_string(n_particles) +
                                             " exceeds SLAMGPU_NOALLOC_MAX_PARTICLES (" + std::to_string(NA_MAX_PARTICLES) +
                                             "): the static scratch holds one partial per 4096 particles");
    int dev = 0;
    CU(cudaGetDevice(&dev));
    if (dev < 0 || dev >= NA_MAX_DEVICES) return slam_fail(SLAMGPU_EINVAL, "slamgpu_estimate_noalloc: device ordinal above 63");
    RET(check_device_pointer(d_xv, dev, "d_xv"));
    RET(check_device_pointer(d_w, dev, "d_w"));
    RET(check_device_pointer(d_out, dev, "d_out"));
    NaScratch *sc = nullptr;
    RET(scratch_for_current_device(dev, &sc));
    cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
    // the array keeps the last call's contents: the ticket must start at zero (a launch that died half way leaves it not)
    CU(cudaMemsetAsync(&sc->ticket, 0, sizeof(unsigned), stream));
    const unsigned grid = (unsigned) ((n_particles + NA_CHUNK - 1) / NA_CHUNK);
    k_estimate_noalloc<<<grid, NA_THREADS, 0, stream>>>(d_xv, d_w, (long long) n_particles, sc, d_out);
    CU(cudaGetLastError());
    return SLAMGPU_OK;
}
