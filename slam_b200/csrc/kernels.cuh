// kernels.cuh — the sm_100a kernels of the FastSLAM hot path.
//
// One thread owns one particle; state is structure-of-arrays so that every warp access is one contiguous
// segment (pose planes: 128 B; landmark rows: float2 -> 256 B, float4 -> 512 B per warp).  Landmark rows live in a
// pool of (nf_cap + spare) physical rows behind a host-side logical -> physical table; resampling does NOT copy the
// map: it appends keep[] to a log, and a row is brought up to date (k_chain + k_materialise_rows below) only when
// it is next touched, or when its backlog would outgrow the log.
//
// All contractions are 2x2 / 3x3 per particle: CUDA-core fp32, no tensor cores (nothing here is GEMM shaped).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "slam_math.cuh"
#include "slam_rng.cuh"
#include "types.cuh"
#include "xchg.cuh"

namespace slam {

__device__ __forceinline__ void load_pose(const PoseSoA &s, size_t P, size_t p, float *x, float *C) {
#pragma unroll
    for (int c = 0; c < 3; c++) x[c] = s.xv[c * P + p];
    if (C) {
#pragma unroll
        for (int j = 0; j < 9; j++) C[j] = s.Pv[j * P + p];
    }
}
__device__ __forceinline__ void store_pose(const PoseSoA &s, size_t P, size_t p, const float *x, const float *C) {
#pragma unroll
    for (int c = 0; c < 3; c++) s.xv[c * P + p] = x[c];
    if (C) {
#pragma unroll
        for (int j = 0; j < 9; j++) s.Pv[j * P + p] = C[j];
    }
}

// ---------------------------------------------------------------------------------------------------------
// K1  predict: FastSLAM{1,2}::predict over all particles (fastslam2.cpp:51-60, fastslam1.cpp:57-66).
// noise: [2][P] staged standard normals or the device generator (slam_rng.cuh).  96 B of state traffic per particle
// (FastSLAM2).
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_predict(PoseSoA s, size_t P, PredictArgs a, const NoiseSrc noise) {
    const size_t p = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P) return;
    const bool need_cov = (a.method == 2) || a.use_heading;
    float x[3], C[9];
    load_pose(s, P, p, x, need_cov ? C : nullptr);
    float g2[2];
    const bool noisy = noise.rng || noise.gauss;
    if (noisy) noise_gauss2(noise, P, p, g2);
    predict_particle(a.method, x, C, a.V, a.G, a.Q, a.dt, a.phi_true, a.wheel_base, a.sigma_phi, a.use_heading != 0,
                     noisy ? g2 : nullptr);
    store_pose(s, P, p, x, need_cov ? C : nullptr);
}

// K1 batched: up to SLAM_PREDICT_QUEUE consecutive noise-free predict ticks applied with the pose held in registers:
// 96 B of traffic per particle for the whole run of ticks instead of per tick.  The host queues ticks
// (slamgpu_predict returns at once) and flushes the queue before anything else looks at the poses, so callers keep
// the reference's one-call-per-control-tick loop (fastslam2wrapper.cpp:51-64).
#define SLAM_PREDICT_QUEUE 16
struct PredictBatch {
    PredictArgs a[SLAM_PREDICT_QUEUE];
    int n;
};
struct PredictBatchFast {  // SLAMGPU_MATH_FAST: control-only terms precomputed on the host (slam_math.cuh PredictTick)
    PredictTick t[SLAM_PREDICT_QUEUE];
    int n, with_cov, use_heading;
};
__global__ void __launch_bounds__(256) k_predict_batch_fast(PoseSoA s, size_t P, const PredictBatchFast b) {
    const size_t p = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P) return;
    const bool need_cov = b.with_cov || b.use_heading;
    float x[3], C[9];
    load_pose(s, P, p, x, need_cov ? C : nullptr);
    for (int t = 0; t < b.n; t++) predict_tick_fast(x, C, b.t[t], b.with_cov != 0, b.use_heading != 0);
    store_pose(s, P, p, x, need_cov ? C : nullptr);
}
__global__ void __launch_bounds__(256) k_predict_batch(PoseSoA s, size_t P, const PredictBatch b) {
    const size_t p = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P) return;
    const bool need_cov = (b.a[0].method == 2) || b.a[0].use_heading;
    float x[3], C[9];
    load_pose(s, P, p, x, need_cov ? C : nullptr);
    for (int t = 0; t < b.n; t++) {
        const PredictArgs &a = b.a[t];
        predict_particle(a.method, x, C, a.V, a.G, a.Q, a.dt, a.phi_true, a.wheel_base, a.sigma_phi,
                         a.use_heading != 0, nullptr);
    }
    store_pose(s, P, p, x, need_cov ? C : nullptr);
}

// ---------------------------------------------------------------------------------------------------------
// K3  FastSLAM2 per-particle update: sampleProposal (fastslam2.cpp:290-368) -> likelihoodGivenXv (370-400) ->
// featureUpdate (core.cpp:132-175), fused.  The observation set (z, physical landmark row) is staged in shared
// memory once per CTA.  Pass 1 walks the K observed landmarks sequentially refining the pose proposal (read only);
// pass 2 re-reads them at the sampled pose, multiplies the likelihood and writes the updated landmark back.
// Algorithmic traffic: 48 B per particle-landmark + 116 B per particle.
// ---------------------------------------------------------------------------------------------------------
struct ObsSet {           // device arrays written by the host before the launch
    const float2 *z;      // [K] range, bearing
    const int *row;       // [K] PHYSICAL landmark row
    const int *row_dst;   // [K] fused re-columning (K3 REMAP): the physical row the updated landmark is written to;
                          //     row_dst[i] != row[i] marks a row whose columns still belong to the previous resample
};

__device__ __forceinline__ void lm_load(const MapRing &m, size_t P, int row, size_t p, float *xf, float *Pf) {
    const float2 a = m.mean[(size_t) row * P + p];
    const float4 c = m.cov[(size_t) row * P + p];
    xf[0] = a.x; xf[1] = a.y;
    Pf[0] = c.x; Pf[1] = c.z; Pf[2] = c.y; Pf[3] = c.w;  // column-major linear -> row-major
}
__device__ __forceinline__ void lm_store(const MapRing &m, size_t P, int row, size_t p, const float *xf,
                                         const float *Pf) {
    __stcs(&m.mean[(size_t) row * P + p], make_float2(xf[0], xf[1]));
    __stcs(&m.cov[(size_t) row * P + p], make_float4(Pf[0], Pf[2], Pf[1], Pf[3]));
}

extern __shared__ __align__(16) unsigned char slam_smem[];

__device__ __forceinline__ void stage_obs(const ObsSet &o, int K, float2 *&sz, int *&srow) {
    sz = reinterpret_cast<float2 *>(slam_smem);
    srow = reinterpret_cast<int *>(sz + K);
    for (int i = threadIdx.x; i < K; i += blockDim.x) {
        sz[i] = o.z[i];
        srow[i] = o.row[i];
    }
    __syncthreads();
}

__global__ void __launch_bounds__(128) k_fs2_update(PoseSoA s, MapRing m, size_t P, ObsSet o, int K, float4 Rm,
                                                    const NoiseSrc noise) {
    float2 *sz;
    int *srow;
    stage_obs(o, K, sz, srow);
    const size_t p = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P) return;
    const float R[4] = {Rm.x, Rm.y, Rm.z, Rm.w};
    float x[3], C[9], x0[3], C0[9];
    load_pose(s, P, p, x, C);
#pragma unroll
    for (int j = 0; j < 3; j++) x0[j] = x[j];
#pragma unroll
    for (int j = 0; j < 9; j++) C0[j] = C[j];
    float g3[3];
    noise_gauss3(noise, P, p, g3);
    float wgt = s.w[p];

    // pass 1: proposal, sequential over observations, next landmark prefetched while this one is processed
    float xf[2], Pf[4], nxf[2], nPf[4];
    lm_load(m, P, srow[0], p, nxf, nPf);
    for (int i = 0; i < K; i++) {
        xf[0] = nxf[0]; xf[1] = nxf[1];
#pragma unroll
        for (int j = 0; j < 4; j++) Pf[j] = nPf[j];
        if (i + 1 < K) lm_load(m, P, srow[i + 1], p, nxf, nPf);
        const float z[2] = {sz[i].x, sz[i].y};
        proposal_step(x, C, R, xf, Pf, z);
    }
    // sample the pose (fastslam2.cpp:352-357) and the two pose-space Gaussian arguments (359-362)
    float xs[3];
    multivariate_gauss<3>(x, C, g3, xs);
    const float v1[3] = {x0[0] - xs[0], x0[1] - xs[1], trig_offset(x0[2] - xs[2])};
    const float v2[3] = {x[0] - xs[0], x[1] - xs[1], trig_offset(x[2] - xs[2])};
    const float prior = gauss_evaluate<3>(v1, C0);
    const float prop = gauss_evaluate<3>(v2, C);

    // pass 2: likelihood at the sampled pose + landmark EKF, independent per observation
    float like = 1.0f;
    lm_load(m, P, srow[0], p, nxf, nPf);
    for (int i = 0; i < K; i++) {
        xf[0] = nxf[0]; xf[1] = nxf[1];
#pragma unroll
        for (int j = 0; j < 4; j++) Pf[j] = nPf[j];
        if (i + 1 < K) lm_load(m, P, srow[i + 1], p, nxf, nPf);
        const float z[2] = {sz[i].x, sz[i].y};
        like = like * likelihood_and_feature_update(xs, R, xf, Pf, z);
        lm_store(m, P, srow[i], p, xf, Pf);
    }
    wgt = wgt * like * prior / prop;  // fastslam2.cpp:367
    s.w[p] = wgt;
    const float Z[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};  // Pv is zeroed after sampling, fastslam2.cpp:355-357
    store_pose(s, P, p, xs, Z);
}

__device__ __forceinline__ void lm_load_at(const float2 *mp, const float4 *cp, float *xf, float *Pf) {
    const float2 a = *mp;
    const float4 c = *cp;
    xf[0] = a.x; xf[1] = a.y;
    Pf[0] = c.x; Pf[1] = c.z; Pf[2] = c.y; Pf[3] = c.w;
}

// K3 (fast formulation, SLAMGPU_MATH_FAST): same kernel structure, ~4x fewer instructions (slam_math.cuh).
// A particle whose pose covariance is not numerically positive definite (e.g. Pv == 0 right after a pose sample)
// is NOT processed by the fast kernel: its index goes to a list and k_fs2_update_listed, launched right behind,
// gives it the reference-order routine (so Eigen's behaviour on a failed LLT is still reproduced).  Keeping that
// routine out of the fast kernel keeps its ~80 registers out of the fast kernel's allocation.
template <bool REMAP>
__device__ __forceinline__ void fs2_update_particle_reference_order(PoseSoA s, MapRing m, size_t P, size_t p,
                                                                    const float2 *sz, const unsigned long long *soff,
                                                                    const unsigned long long *doff,
                                                                    const float2 *smean, const float4 *scov,
                                                                    int K, float4 Rm, const NoiseSrc &noise) {
    const float R[4] = {Rm.x, Rm.y, Rm.z, Rm.w};
    float x[3], C[9], x0[3], C0[9];
    load_pose(s, P, p, x, C);
    for (int j = 0; j < 3; j++) x0[j] = x[j];
    for (int j = 0; j < 9; j++) C0[j] = C[j];
    float g3[3];
    noise_gauss3(noise, P, p, g3);
    float xf[2], Pf[4];
    for (int i = 0; i < K; i++) {
        const bool stale = REMAP && soff[i] != doff[i];
        lm_load_at((stale ? smean : m.mean + p) + soff[i], (stale ? scov : m.cov + p) + soff[i], xf, Pf);
        const float z[2] = {sz[i].x, sz[i].y};
        proposal_step(x, C, R, xf, Pf, z);
    }
    float xs[3];
    multivariate_gauss<3>(x, C, g3, xs);
    const float v1[3] = {x0[0] - xs[0], x0[1] - xs[1], trig_offset(x0[2] - xs[2])};
    const float v2[3] = {x[0] - xs[0], x[1] - xs[1], trig_offset(x[2] - xs[2])};
    const float prior = gauss_evaluate<3>(v1, C0);
    const float prop = gauss_evaluate<3>(v2, C);
    float like = 1.0f;
    for (int i = 0; i < K; i++) {
        const bool stale = REMAP && soff[i] != doff[i];
        const size_t dst = (REMAP ? doff[i] : soff[i]) + p;
        lm_load_at((stale ? smean : m.mean + p) + soff[i], (stale ? scov : m.cov + p) + soff[i], xf, Pf);
        const float z[2] = {sz[i].x, sz[i].y};
        like = like * likelihood_and_feature_update(xs, R, xf, Pf, z);
        __stcs(m.mean + dst, make_float2(xf[0], xf[1]));
        __stcs(m.cov + dst, make_float4(Pf[0], Pf[2], Pf[1], Pf[3]));
    }
    s.w[p] = s.w[p] * like * prior / prop;
    const float Z[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    store_pose(s, P, p, xs, Z);
}

// Observation staging for the fast kernels: z and the element offset row*P of each observed landmark row, padded by
// one copy of the last entry so that the "next landmark" prefetch needs no branch.
__device__ __forceinline__ void stage_obs_off(const ObsSet &o, int K, size_t P, float2 *&sz, unsigned long long *&soff) {
    soff = reinterpret_cast<unsigned long long *>(slam_smem);
    sz = reinterpret_cast<float2 *>(soff + (K + 1));
    for (int i = threadIdx.x; i <= K; i += blockDim.x) {
        const int j = min(i, K - 1);
        sz[i] = o.z[j];
        soff[i] = (unsigned long long) o.row[j] * (unsigned long long) P;
    }
    __syncthreads();
}
#define SLAM_OBS_OFF_SMEM(K) ((sizeof(unsigned long long) + sizeof(float2)) * ((K) + 1))

// Shared-memory staging of the fast kernels: source / destination element offsets (row * P), z, and (sin, cos) of the
// observed bearing (slam_math.cuh bearing_innovation: the pass over a fixed pose composes the ray's sine and cosine
// from these and the pose's own instead of evaluating an arctangent per particle-landmark).
template <bool REMAP>
__device__ __forceinline__ void stage_obs_fs2(unsigned char *base, const ObsSet &o, int K, size_t P, float2 *&sz,
                                              unsigned long long *&soff, unsigned long long *&doff, float2 *&ssc) {
    soff = reinterpret_cast<unsigned long long *>(base);
    doff = soff + K;
    sz = reinterpret_cast<float2 *>(doff + K);
    ssc = sz + K;
    for (int i = threadIdx.x; i < K; i += blockDim.x) {
        const float2 z = o.z[i];
        sz[i] = z;
        float sb, cb;
        sincosf(z.y, &sb, &cb);
        ssc[i] = make_float2(sb, cb);
        soff[i] = (unsigned long long) o.row[i] * (unsigned long long) P;
        doff[i] = (unsigned long long) (REMAP ? o.row_dst[i] : o.row[i]) * (unsigned long long) P;
    }
    __syncthreads();
}
#define SLAM_OBS_FS2_SMEM(K) ((2 * sizeof(unsigned long long) + 2 * sizeof(float2)) * (K))

// The same observation set as the fast FastSLAM2 update kernel wants it: physical ROW NUMBERS instead of 64-bit element
// offsets (an address is then base + row * pitch, one 32 x 32 + 64-bit multiply-add per array instead of a 64-bit add
// and two 64-bit shifted adds; P * 16 < 2^32 is checked by slamgpu_create), and what pass 2 reads per observation in
// ONE 16-byte word: (z_range, sin z_b, cos z_b, destination row).  Same footprint as stage_obs_fs2 (32 B / observation).
template <bool REMAP>
__device__ __forceinline__ void stage_obs_fast(unsigned char *base, const ObsSet &o, int K, float4 *&sp2, uint2 *&srow,
                                               float2 *&sz) {
    sp2 = reinterpret_cast<float4 *>(base);
    srow = reinterpret_cast<uint2 *>(sp2 + K);
    sz = reinterpret_cast<float2 *>(srow + K);
    for (int i = threadIdx.x; i < K; i += blockDim.x) {
        const float2 z = o.z[i];
        sz[i] = z;
        float sb, cb;
        sincosf(z.y, &sb, &cb);
        const unsigned rs = (unsigned) o.row[i], rd = (unsigned) (REMAP ? o.row_dst[i] : o.row[i]);
        sp2[i] = make_float4(z.x, sb, cb, __uint_as_float(rd));
        srow[i] = make_uint2(rs, rd);
    }
    __syncthreads();
}
// base + row * pitch in one 32 x 32 + 64-bit multiply-add (IMAD.WIDE.U32)
__device__ __forceinline__ const char *row_addr(const char *base, unsigned row, unsigned pitch) {
    unsigned long long a;
    asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(a) : "r"(row), "r"(pitch), "l"((unsigned long long) base));
    return reinterpret_cast<const char *>(a);
}

// anc: the resample's keep[] (or a composed ancestor map) for this device's slots; -1 means "the last particle"
// exactly as in k_gather_pose (core.cpp:800-806 leaves -1 where select ran past the CDF).
__device__ __forceinline__ long long remap_column(const int *__restrict__ anc, size_t p, long long n_global) {
    const int k = anc[p];
    return k < 0 ? n_global - 1 : (long long) k;
}
// The ancestor's column as pointers into the pool of the rank that holds it (global index q; rank = q / P).
__device__ __forceinline__ void remap_source(const PeerRows &pr, long long q, size_t P, const float2 *&smean,
                                             const float4 *&scov) {
    const int r = (pr.world > 1) ? (int) (q / (long long) P) : 0;
    const long long ql = q - (long long) r * (long long) P;
    smean = pr.mean[r] + ql;
    scov = pr.cov[r] + ql;
}

// Sharded resampling pulls particles across the shard edges, so the slots whose ancestor column lives on a peer sit
// together at one end of the shard: in launch order their CTAs would run together, NVLink-bound, as the kernel's tail
// (N=8: the update kernel of the worst rank took ~2.5x its HBM-bound time).  spread_block() walks the grid in chunks of
// SLAM_SPREAD_CHUNK_SLOTS consecutive slots with a stride coprime to the chunk count, so peer reads are interleaved
// with local, HBM-bound CTAs for the whole kernel and hide behind them; a chunk stays contiguous (DRAM pages).
#define SLAM_SPREAD_CHUNK_SLOTS 2048
__device__ __forceinline__ unsigned spread_block(unsigned b, unsigned nblocks, unsigned stride, unsigned per_chunk) {
    if (stride == 0) return b;
    const unsigned nch = nblocks / per_chunk, c = b / per_chunk;
    return (unsigned) (((unsigned long long) c * stride) % nch) * per_chunk + (b % per_chunk);
}

template <bool REMAP>
__global__ void __launch_bounds__(128) k_fs2_update_listed(PoseSoA s, MapRing m, size_t P, ObsSet o, int K, float4 Rm,
                                                           const NoiseSrc noise,
                                                           const int *__restrict__ anc, PeerRows pr, long long n_global,
                                                           const int *__restrict__ list, const int *__restrict__ count) {
    float2 *sz, *ssc;
    unsigned long long *soff, *doff;
    stage_obs_fs2<REMAP>(slam_smem, o, K, P, sz, soff, doff, ssc);
    const int n = *count;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const size_t p = (size_t) list[i];
        const float2 *smean = nullptr;
        const float4 *scov = nullptr;
        if (REMAP) remap_source(pr, remap_column(anc, p, n_global), P, smean, scov);
        fs2_update_particle_reference_order<REMAP>(s, m, P, p, sz, soff, doff, smean, scov, K, Rm, noise);
    }
}

// Per-thread asynchronous landmark ring.  Each thread copies ITS OWN column of the next landmark rows from HBM into its
// own shared-memory slots with cp.async (LDGSTS) and reads them back later: shared memory serves as a per-thread
// extension of the register file for loads in flight, so the prefetch depth (SLAM_RING_DEPTH rows = 24 B x depth per
// thread, ~64 KB in flight per SM) no longer costs registers and no CTA-wide barrier is needed (nobody reads
// another thread's slot).  With one row in flight per thread the kernel sat on long-scoreboard stalls at 58 % of the
// HBM peak (ncu r01 v2, profiles/).
#define SLAM_RING_DEPTH 4
#define SLAM_UPDATE_THREADS 128
__device__ __forceinline__ void cp_async_8(void *smem_dst, const void *gsrc) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((unsigned) __cvta_generic_to_shared(smem_dst)), "l"(gsrc)
                 : "memory");
}
__device__ __forceinline__ void cp_async_16(void *smem_dst, const void *gsrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned) __cvta_generic_to_shared(smem_dst)), "l"(gsrc)
                 : "memory");
}
// 16 bytes through L1 as well.  Used for rows read at an ANCESTOR's column: after many resamples a whole shard descends
// from a handful of columns, so thousands of warps ask for the same 32-byte sectors — on a peer device when the
// ancestor migrated.  Through L1 those reads are served on the SM after the first one; with .cg every warp's copy went
// over NVLink to the same address (measured at 2 GPUs: +2 ms on the update kernel of the rank whose ancestors had
// moved, every time a long-unobserved row came back into view).
__device__ __forceinline__ void cp_async_16_ca(void *smem_dst, const void *gsrc) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"((unsigned) __cvta_generic_to_shared(smem_dst)), "l"(gsrc)
                 : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
#define SLAM_FS2_FAST_SMEM(K) \
    ((sizeof(float4) + sizeof(float2)) * SLAM_RING_DEPTH * SLAM_UPDATE_THREADS + SLAM_OBS_FS2_SMEM(K))

// REMAP = the re-columning of deferred map resampling (K7) fused into the update: an observed row whose columns still
// belong to the previous resample epoch is READ at column anc[p] of its old physical row and the updated landmark is
// WRITTEN to column p of a fresh physical row, so the row needs no separate k_materialise_rows pass (one read + one
// write of 24 B per particle-landmark saved on every update that follows a resample).
// Waits until this warp's chunk of a noise buffer that is still being uploaded has landed (GaussArrival), then reads
// the thread's three Gaussians with loads the compiler cannot move above the wait.
__device__ __forceinline__ void gauss_arrived_load(const GaussArrival &ga, const NoiseSrc &noise, size_t P, size_t p,
                                                   float g3[3]) {
    const float *gauss = noise.gauss;
    if (noise.rng) {
        rng_gauss3(noise.seed, noise.slot0 + p, noise.draw, g3);
    } else if (ga.flags) {
        const unsigned am = __activemask();  // lanes that left for the fallback list are not here
        if ((int) (threadIdx.x & 31) == __ffs(am) - 1) {
            const unsigned *f = ga.flags + (p >> ga.shift);
            unsigned v;
            const long long t0 = clock64();
            while (true) {
                // acquire at system scope: the flag is written by the copy engine behind the chunk it announces, and the
                // data loads below must observe what that write follows
                asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(f) : "memory");
                if (v) break;
                if (clock64() - t0 > 4000000000ll) {  // ~2 s: the copy stream died
                    *ga.late = 1;
                    break;
                }
                __nanosleep(256);
            }
        }
        __syncwarp(am);
#pragma unroll
        for (int c = 0; c < 3; c++)
            asm volatile("ld.global.cg.f32 %0, [%1];" : "=f"(g3[c]) : "l"(gauss + c * P + p) : "memory");
    } else {
        g3[0] = gauss[p];
        g3[1] = gauss[P + p];
        g3[2] = gauss[2 * P + p];
    }
}

// REMAP: 0 = every observed row is current; 1 = some are stale (checked per row: source row != destination row); 2 = ALL
// K are stale — the step after a resample, the common case — so the per-row check and the second pair of copies drop
// out.  KFULL: K is a multiple of the ring depth (and >= it), so the "is there a row i" tests of the unrolled walk drop
// out and "is there a row to prefetch" is one test per D rows.  The kernel is fp32-issue bound (ncu r02: issue slots
// 83 % busy), so integer bookkeeping per particle-landmark is paid for in update time one for one.
template <int REMAP, bool KFULL>
__global__ void __launch_bounds__(SLAM_UPDATE_THREADS, 7) k_fs2_update_fast(PoseSoA s, MapRing m, size_t P, ObsSet o, int K,
                                                                         float4 Rm, const NoiseSrc noise, GaussArrival ga,
                                                                         const int *__restrict__ anc, PeerRows pr,
                                                                         long long n_global,
                                                                         int *__restrict__ list, int *__restrict__ count) {
    constexpr int D = SLAM_RING_DEPTH, T = SLAM_UPDATE_THREADS;
    float4 *const ring_cov = reinterpret_cast<float4 *>(slam_smem) + threadIdx.x;                // slot s at [s * T]
    float2 *const ring_mean = reinterpret_cast<float2 *>(reinterpret_cast<float4 *>(slam_smem) + D * T) + threadIdx.x;
    float4 *sp2;
    uint2 *srow;
    float2 *sz;
    stage_obs_fast<(REMAP != 0)>(slam_smem + (sizeof(float4) + sizeof(float2)) * D * T, o, K, sp2, srow, sz);
    const unsigned blk = REMAP ? spread_block(blockIdx.x, gridDim.x, pr.spread, SLAM_SPREAD_CHUNK_SLOTS / T) : blockIdx.x;
    const size_t p = (size_t) blk * blockDim.x + threadIdx.x;
    if (p >= P) return;
    // row r of this particle's column sits at base + r * pitch (bytes)
    const unsigned pitch_m = (unsigned) (P * sizeof(float2)), pitch_c = (unsigned) (P * sizeof(float4));
    const char *const mean_b = reinterpret_cast<const char *>(m.mean + p);
    const char *const cov_b = reinterpret_cast<const char *>(m.cov + p);
    // a stale row (REMAP) is read at the ancestor's column, on whichever rank holds it
    const char *smean_b = mean_b, *scov_b = cov_b;
    if (REMAP) {
        const float2 *smean;
        const float4 *scov;
        remap_source(pr, remap_column(anc, p, n_global), P, smean, scov);
        smean_b = reinterpret_cast<const char *>(smean);
        scov_b = reinterpret_cast<const char *>(scov);
    }
    // row i goes in flight into ring slot `slot`: its own column, or (stale) the ancestor's, then through L1 as well
    auto fetch = [&](int i, int slot) {
        if (REMAP == 2) {
            const unsigned r = srow[i].x;
            cp_async_8(ring_mean + slot * T, row_addr(smean_b, r, pitch_m));
            cp_async_16_ca(ring_cov + slot * T, row_addr(scov_b, r, pitch_c));
        } else if (REMAP == 1) {
            const uint2 rr = srow[i];
            if (rr.x != rr.y) {
                cp_async_8(ring_mean + slot * T, row_addr(smean_b, rr.x, pitch_m));
                cp_async_16_ca(ring_cov + slot * T, row_addr(scov_b, rr.x, pitch_c));
            } else {
                cp_async_8(ring_mean + slot * T, row_addr(mean_b, rr.x, pitch_m));
                cp_async_16(ring_cov + slot * T, row_addr(cov_b, rr.x, pitch_c));
            }
        } else {
            const unsigned r = srow[i].x;
            cp_async_8(ring_mean + slot * T, row_addr(mean_b, r, pitch_m));
            cp_async_16(ring_cov + slot * T, row_addr(cov_b, r, pitch_c));
        }
    };
    // the first D-1 rows go in flight before anything else is looked at
#pragma unroll
    for (int j = 0; j < D - 1; j++) {
        if (KFULL || j < K) fetch(j, j);
        cp_async_commit();
    }
    const float R[4] = {Rm.x, Rm.y, Rm.z, Rm.w};
    float x[3], x0[3], C0[9], L0[6];
    load_pose(s, P, p, x0, C0);
    ProposalInfo pi;
    if (!proposal_begin_fast(C0, L0, pi)) {
        list[atomicAdd(count, 1)] = (int) p;
        cp_async_wait<0>();
        return;
    }
#pragma unroll
    for (int j = 0; j < 3; j++) x[j] = x0[j];

    // pass 1: proposal (read only), sequential over the observations (fastslam2.cpp:314-350); row i sits in slot i % D
    float xf[2], Pf[4];
    for (int i0 = 0; i0 < K; i0 += D) {
        const bool more = i0 + D < K;  // KFULL: rows i0 + D .. i0 + 2D - 1 exist
#pragma unroll
        for (int sl = 0; sl < D; sl++) {
            const int i = i0 + sl;
            if (KFULL || i < K) {
                const int j = i + D - 1;
                if (KFULL ? (sl == 0 || more) : (j < K)) fetch(j, (sl + D - 1) % D);
                cp_async_commit();
                cp_async_wait<D - 1>();  // one group per row: all but the newest D-1 have landed, row i among them
                lm_load_at(ring_mean + sl * T, ring_cov + sl * T, xf, Pf);
                const float2 zi = sz[i];
                const float z[2] = {zi.x, zi.y};
                proposal_step_fast(x, pi, R, xf, Pf, z);
            }
        }
    }
    float g3[3];
    gauss_arrived_load(ga, noise, P, p, g3);
    float xs[3];
    const float ratio = proposal_finish_fast(x0, L0, x, pi, g3, xs);
    float sth, cth;  // the sampled pose is fixed from here on: its sine and cosine once, not an arctangent per landmark
    sincosf(xs[2], &sth, &cth);

    // pass 2: likelihood at the sampled pose + landmark EKF, written back in place.  The observations are independent
    // here, so they are walked BACKWARDS: the last D rows of pass 1 are still in the ring, and the rows read most
    // recently are the ones most likely still in L2 (a forward walk is the LRU worst case).
    float like = 1.0f;
    for (int i0 = ((K - 1) / D) * D; i0 >= 0; i0 -= D) {
        const bool more = i0 >= D;  // rows i0 - D .. i0 - 1 exist
#pragma unroll
        for (int sl = D - 1; sl >= 0; sl--) {
            const int i = i0 + sl;
            if (KFULL || i < K) {
                cp_async_wait<D - 1>();
                lm_load_at(ring_mean + sl * T, ring_cov + sl * T, xf, Pf);
                const float4 ob = sp2[i];  // z_range, sin z_b, cos z_b, destination row
                if (more) fetch(i - D, sl);  // the slot just read takes row i - D
                cp_async_commit();
                like = like * likelihood_and_feature_update_fast(xs, sth, cth, R, xf, Pf, ob.x, ob.y, ob.z);
                const unsigned rd = __float_as_uint(ob.w);
                __stcs(reinterpret_cast<float2 *>(const_cast<char *>(row_addr(mean_b, rd, pitch_m))), make_float2(xf[0], xf[1]));
                __stcs(reinterpret_cast<float4 *>(const_cast<char *>(row_addr(cov_b, rd, pitch_c))),
                       make_float4(Pf[0], Pf[2], Pf[1], Pf[3]));
            }
        }
    }
    s.w[p] = s.w[p] * like * ratio;
    const float Z[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    store_pose(s, P, p, xs, Z);
}

// REMAP as in k_fs2_update_fast: a stale observed row is read at the ancestor's column (on whichever rank holds it) and
// written to a fresh physical row, so the re-columning of the observed rows costs no extra pass.
template <bool REMAP>
__global__ void __launch_bounds__(256) k_fs1_update_fast(PoseSoA s, MapRing m, size_t P, ObsSet o, int K, float4 Rm,
                                                         const int *__restrict__ anc, PeerRows pr, long long n_global) {
    float2 *sz, *ssc;
    unsigned long long *soff, *doff;
    stage_obs_fs2<REMAP>(slam_smem, o, K, P, sz, soff, doff, ssc);
    const unsigned blk = REMAP ? spread_block(blockIdx.x, gridDim.x, pr.spread, SLAM_SPREAD_CHUNK_SLOTS / 256) : blockIdx.x;
    const size_t p = (size_t) blk * blockDim.x + threadIdx.x;
    if (p >= P) return;
    const float R[4] = {Rm.x, Rm.y, Rm.z, Rm.w};
    float x[3];
    load_pose(s, P, p, x, nullptr);
    float sth, cth;
    sincosf(x[2], &sth, &cth);
    float2 *const mean_p = m.mean + p;
    float4 *const cov_p = m.cov + p;
    const float2 *smean = mean_p;
    const float4 *scov = cov_p;
    if (REMAP) remap_source(pr, remap_column(anc, p, n_global), P, smean, scov);
    float wt = 1.0f;
    float xf[2], Pf[4], nxf[2], nPf[4];
    {
        const bool stale = REMAP && soff[0] != doff[0];
        lm_load_at((stale ? smean : mean_p) + soff[0], (stale ? scov : cov_p) + soff[0], nxf, nPf);
    }
    for (int i = 0; i < K; i++) {
        xf[0] = nxf[0]; xf[1] = nxf[1];
#pragma unroll
        for (int j = 0; j < 4; j++) Pf[j] = nPf[j];
        const int n = min(i + 1, K - 1);  // the last iteration re-reads its own row (no branch around the prefetch)
        const unsigned long long noff = soff[n];
        const bool stale = REMAP && noff != doff[n];
        lm_load_at((stale ? smean : mean_p) + noff, (stale ? scov : cov_p) + noff, nxf, nPf);
        const float2 sc = ssc[i];
        wt = wt * likelihood_and_feature_update_fast(x, sth, cth, R, xf, Pf, sz[i].x, sc.x, sc.y);
        const unsigned long long d = doff[i];
        __stcs(mean_p + d, make_float2(xf[0], xf[1]));
        __stcs(cov_p + d, make_float4(Pf[0], Pf[2], Pf[1], Pf[3]));
    }
    s.w[p] = s.w[p] * wt;
}

// K3'  FastSLAM1 per-particle update: computeWeight (fastslam1.cpp:91-118) + featureUpdate, one pass.
__global__ void __launch_bounds__(256) k_fs1_update(PoseSoA s, MapRing m, size_t P, ObsSet o, int K, float4 Rm) {
    float2 *sz;
    int *srow;
    stage_obs(o, K, sz, srow);
    const size_t p = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P) return;
    const float R[4] = {Rm.x, Rm.y, Rm.z, Rm.w};
    float x[3];
    load_pose(s, P, p, x, nullptr);
    float wt = 1.0f;
    float xf[2], Pf[4], nxf[2], nPf[4];
    lm_load(m, P, srow[0], p, nxf, nPf);
    for (int i = 0; i < K; i++) {
        xf[0] = nxf[0]; xf[1] = nxf[1];
#pragma unroll
        for (int j = 0; j < 4; j++) Pf[j] = nPf[j];
        if (i + 1 < K) lm_load(m, P, srow[i + 1], p, nxf, nPf);
        const float z[2] = {sz[i].x, sz[i].y};
        fs1_weight_and_feature_update(x, R, xf, Pf, z, &wt);
        lm_store(m, P, srow[i], p, xf, Pf);
    }
    s.w[p] = s.w[p] * wt;
}

// ---------------------------------------------------------------------------------------------------------
// K4  addFeature (core.cpp:479-509) for Kn new observations at PHYSICAL rows o.row[i].  sample_first implements
// FastSLAM2's "no known landmark this step" branch (fastslam2.cpp:36-42): xv ~ N(xv, Pv), Pv = 0 first.
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_add_feature(PoseSoA s, MapRing m, size_t P, ObsSet o, int Kn, float4 Rm,
                                                     int sample_first, const NoiseSrc noise) {
    float2 *sz;
    int *srow;
    stage_obs(o, Kn, sz, srow);
    const size_t p = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P) return;
    const float R[4] = {Rm.x, Rm.y, Rm.z, Rm.w};
    float x[3];
    if (sample_first) {
        float C[9], xs[3];
        load_pose(s, P, p, x, C);
        float g3[3];
        noise_gauss3(noise, P, p, g3);
        multivariate_gauss<3>(x, C, g3, xs);
        const float Z[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
        store_pose(s, P, p, xs, Z);
        x[0] = xs[0]; x[1] = xs[1]; x[2] = xs[2];
    } else {
        load_pose(s, P, p, x, nullptr);
    }
    for (int i = 0; i < Kn; i++) {
        float xf[2], Pf[4];
        add_feature_one(x, sz[i].x, sz[i].y, R, xf, Pf);
        lm_store(m, P, srow[i], p, xf, Pf);
    }
}

// ---------------------------------------------------------------------------------------------------------
// K2  batched computeJacobians over the resident particle set: out[(i*P+p)*16 + 0..15] in the accelerator
// record order zp[2], Hf[4], Hv[6], Sf[4] (core.cpp:624-664).  64 B written per particle-landmark.
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void jacobian_record(const float *x, const float *R, const float *xf, const float *Pf,
                                                float4 *out) {
    float zp[2], Hf[4], Hv[6], Sf[4];
    jacobian_one(x, R, xf, Pf, zp, Hf, Sf);
    hv_from_hf(Hf, Hv);
    out[0] = make_float4(zp[0], zp[1], Hf[0], Hf[1]);
    out[1] = make_float4(Hf[2], Hf[3], Hv[0], Hv[1]);
    out[2] = make_float4(Hv[2], Hv[3], Hv[4], Hv[5]);
    out[3] = make_float4(Sf[0], Sf[1], Sf[2], Sf[3]);
}

__global__ void __launch_bounds__(256) k_jacobians_batch(PoseSoA s, MapRing m, size_t P, ObsSet o, int K, float4 Rm,
                                                         float4 *__restrict__ out) {
    float2 *sz;
    int *srow;
    stage_obs(o, K, sz, srow);
    const size_t p = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P) return;
    const float R[4] = {Rm.x, Rm.y, Rm.z, Rm.w};
    float x[3];
    load_pose(s, P, p, x, nullptr);
    for (int i = 0; i < K; i++) {
        float xf[2], Pf[4];
        lm_load(m, P, srow[i], p, xf, Pf);
        jacobian_record(x, R, xf, Pf, out + ((size_t) i * P + p) * 4);
    }
}

// The literal accelerator record for ONE particle (core.cpp:586-664): thread i handles landmark i.
__global__ void k_jacobians_record(const float *__restrict__ in, int n, float4 *__restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float x[3] = {in[0], in[1], in[2]};
    const float R[4] = {in[3], in[5], in[4], in[6]};
    const float *r = in + 7 + 6 * i;
    const float xf[2] = {r[0], r[1]};
    const float Pf[4] = {r[2], r[4], r[3], r[5]};
    jacobian_record(x, R, xf, Pf, out + (size_t) i * 4);
}

// ---------------------------------------------------------------------------------------------------------
// K5/K6  weights -> CDF -> keep[]  (resampleParticles core.cpp:718-734, stratifiedResample 780-807)
//
// BLOCKED order (SLAMGPU_SCAN_BLOCKED): the fixed-shape inclusive scan of oracle/slam_oracle.c blocked_scan():
//   a block = SLAM_SCAN_BLOCK consecutive elements, one per thread; each warp scans its 32 elements with the
//   Hillis-Steele ladder (x += shfl_up(x, d), d = 1..16), the 32 warp totals are scanned by the same ladder, and
//   element = warp_exclusive + local.  Block totals are scanned recursively by the same shape.  Every add is one
//   fp32 add, so the result does not depend on how many devices hold the blocks.
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ float ladder32(float v, int lane) {
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const float t = __shfl_up_sync(0xffffffffu, v, d);
        if (lane >= d) v = __fadd_rn(v, t);
    }
    return v;
}
__device__ __forceinline__ float ladder32_max(float v, int lane) {
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const float t = __shfl_up_sync(0xffffffffu, v, d);
        if (lane >= d) v = fmaxf(v, t);
    }
    return v;
}

// Inclusive blocked scan of one 1024-element block held one element per thread.  sh: 32 floats of shared memory.
__device__ __forceinline__ float block_scan_1024(float v, float *sh) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const float loc = ladder32(v, lane);
    __syncthreads();  // sh may still be read by a previous call
    if (lane == 31) sh[wid] = loc;
    __syncthreads();
    if (wid == 0) sh[lane] = ladder32(sh[lane], lane);
    __syncthreads();
    return (wid > 0) ? __fadd_rn(sh[wid - 1], loc) : loc;
}
__device__ __forceinline__ float block_maxscan_1024(float v, float *sh) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const float loc = ladder32_max(v, lane);
    __syncthreads();
    if (lane == 31) sh[wid] = loc;
    __syncthreads();
    if (wid == 0) sh[lane] = ladder32_max(sh[lane], lane);
    __syncthreads();
    return (wid > 0) ? fmaxf(sh[wid - 1], loc) : loc;
}

// ---- the blocked weights pass: Sigma w -> normalise -> in-block CDF -> block totals -> nEff, block maxima --------------
//
// One pass structure shared by the cooperative kernel (k_weights_fused: ONE launch, two grid-wide barriers, the
// cross-rank exchanges done by CTA 0 inside the barriers) and by the split launches (k_weights_a / _c / _d with
// k_xchg in between; SLAMGPU_WEIGHTS_SPLIT=1, same device functions, same bits):
//   A  per 1024-block scan of w                         -> tot_w[b]                       | exchange tot_w
//   B  every CTA: ws = blocked total of tot_w (the recursion's upper levels, read only)
//   C  wn = w / ws (one IEEE divide, stored back), in-block scans of wn and wn^2, running maximum of the in-block
//      CDF                                              -> cdf_loc, mloc_loc, tot_c, tot_s, lmax | exchange those
//   D  CTA 0: Tc = blocked scan of tot_c (in place), bm[b] = prefix max of fl(Tc[b-1] + lmax[b]);
//      CTA 1: nEff = 1 / blocked total of tot_s, and the resample decision of core.cpp:739
// The final CDF fl(Tc[b-1] + cdf_loc[i]) is never stored: rounding is monotone, so the running maximum of the final
// CDF inside a block is fl(Tc[b-1] + mloc_loc[i]) and the block's maximum is fl(Tc[b-1] + lmax[b]) — the search adds
// the block prefix to the in-block values it probes.  What a peer has to see of this device is therefore complete
// after the SECOND exchange: the third exchange (and its barrier) of the earlier version is gone.
struct WeightsArgs {
    float *w;
    size_t n;                              // particles on this device
    float *tot_w, *tot_c, *tot_s, *lmax;   // [nbg] each, indexed by GLOBAL block (the slab every rank holds whole)
    float *bm;                             // [nbg] out: prefix maxima of the block maxima of the final CDF
    float *cdf_loc, *mloc_loc;             // [n] out: in-block CDF of the normalised weights, its running maximum
    int b0;                                // first global block held here
    int nb, nbg;                           // local / global 1024-blocks
    Scalars *sc;
    float neff_threshold;                  // (float) NEFFECTIVE
    int do_resample;
    float uniform_w;                       // 1 / N over the GLOBAL particle count
    int keep_nan;                          // 1: the reference's behaviour on a zero weight sum (NaN weights for good)
};

// A weight sum that is zero or not finite: every particle's likelihood underflowed (a loop closure the whole cloud
// disagrees with, 100 particles and a 1-degree bearing sigma: seen on BASELINE.json config 0) or something overflowed.
// The reference divides by it (core.cpp:727-729): every weight becomes NaN, nEff is NaN, `nEff < NEFFECTIVE` is false
// for ever after and the filter never resamples again.  Unless asked for exactly that (slamgpu_config.keep_nan_weights)
// the library carries the particle set on UNWEIGHTED instead: w = 1/N, nEff = N, no resample this step, and counts the
// event (SLAMGPU_STAT_WEIGHT_RESETS).  Every CTA and every rank sees the same sum, so all take the same branch.
__device__ __forceinline__ bool weights_degenerate(const WeightsArgs &a, float ws) {
    return !a.keep_nan && !(ws > 0.0f && ws <= 3.0e38f);
}

__device__ __forceinline__ void weights_phase_a(const WeightsArgs &a, int lb, float *sh) {
    const size_t i = (size_t) lb * SLAM_SCAN_BLOCK + threadIdx.x;
    const float v = (i < a.n) ? a.w[i] : 0.0f;
    const float sc = block_scan_1024(v, sh);
    const size_t last = min(a.n, ((size_t) lb + 1) * SLAM_SCAN_BLOCK) - 1;
    if (i == last) a.tot_w[a.b0 + lb] = sc;
}

// Normalise (core.cpp:727-729 / 784-786: one quotient serves both), in-block CDF, totals of wn and wn^2.
// pow(w,2) is a double product stored as float == one correctly rounded fp32 multiply.
__device__ __forceinline__ void weights_phase_c(const WeightsArgs &a, int lb, float ws, float *sh) {
    const size_t i = (size_t) lb * SLAM_SCAN_BLOCK + threadIdx.x;
    float wn = 0.0f;
    if (i < a.n) {
        wn = weights_degenerate(a, ws) ? a.uniform_w : __fdiv_rn(a.w[i], ws);
        a.w[i] = wn;
    }
    const float sq = __fmul_rn(wn, wn);
    const float sc = block_scan_1024(wn, sh);
    const float ss = block_scan_1024(sq, sh);
    const float mx = block_maxscan_1024((i < a.n) ? sc : -INFINITY, sh);
    if (i < a.n) {
        a.cdf_loc[i] = sc;
        a.mloc_loc[i] = mx;
    }
    const size_t last = min(a.n, ((size_t) lb + 1) * SLAM_SCAN_BLOCK) - 1;
    if (i == last) {
        a.tot_c[a.b0 + lb] = sc;
        a.tot_s[a.b0 + lb] = ss;
        a.lmax[a.b0 + lb] = mx;
    }
}

// The upper levels of the blocked scan over a totals array x[n], n <= 1024 * 1024, by ONE CTA of 1024 threads:
// chunks of 1024 are scanned one after the other, the chunk totals are scanned as one block (the recursion's last
// level) and added back.  The slab arrays are written by other CTAs / other devices: read through L2 (ld.cg).
// upper_total: read only; returns the blocked SUM of the level below = scanned[n-2] + x[n-1] to every thread.
__device__ __forceinline__ float upper_total(const float *x, int n, float *sh, float *chunk_tot, float *bcast) {
    __syncthreads();  // bcast / chunk_tot may still be read from a previous call
    if (n == 1) {
        if (threadIdx.x == 0) *bcast = __ldcg(x);
        __syncthreads();
        return *bcast;
    }
    const int e = n - 2, ce = e / SLAM_SCAN_BLOCK, oe = e % SLAM_SCAN_BLOCK;
    for (int c = 0; c <= ce; c++) {
        const int i = c * SLAM_SCAN_BLOCK + threadIdx.x;
        const float v = (i < n) ? __ldcg(x + i) : 0.0f;
        const float sc = block_scan_1024(v, sh);
        const int last = min(n, (c + 1) * SLAM_SCAN_BLOCK) - 1;
        if (i == last) chunk_tot[c] = sc;
        if (c == ce && (int) threadIdx.x == oe) *bcast = sc;
    }
    __syncthreads();
    if (ce > 0) {
        const float v = ((int) threadIdx.x < ce) ? chunk_tot[threadIdx.x] : 0.0f;  // chunks 0 .. ce-1 precede element e
        const float sc = block_scan_1024(v, sh);
        if ((int) threadIdx.x == ce - 1) *bcast = __fadd_rn(sc, *bcast);
        __syncthreads();
    }
    return __fadd_rn(*bcast, __ldcg(x + n - 1));
}
// upper_scan_inplace: x[i] <- blocked inclusive scan (add, or max); returns nothing.
__device__ __forceinline__ void upper_scan_inplace(float *x, int n, bool use_max, float *sh, float *chunk_tot) {
    const int nchunk = (n + SLAM_SCAN_BLOCK - 1) / SLAM_SCAN_BLOCK;
    __syncthreads();
    for (int c = 0; c < nchunk; c++) {
        const int i = c * SLAM_SCAN_BLOCK + threadIdx.x;
        const float v = (i < n) ? __ldcg(x + i) : (use_max ? -INFINITY : 0.0f);
        const float sc = use_max ? block_maxscan_1024(v, sh) : block_scan_1024(v, sh);
        if (i < n) __stcg(x + i, sc);
        const int last = min(n, (c + 1) * SLAM_SCAN_BLOCK) - 1;
        if (i == last) chunk_tot[c] = sc;
    }
    __syncthreads();
    if (nchunk > 1) {
        const float v = ((int) threadIdx.x < nchunk) ? chunk_tot[threadIdx.x] : (use_max ? -INFINITY : 0.0f);
        const float sc = use_max ? block_maxscan_1024(v, sh) : block_scan_1024(v, sh);
        __syncthreads();
        if ((int) threadIdx.x < nchunk) chunk_tot[threadIdx.x] = sc;
        __syncthreads();
        for (int c = 1; c < nchunk; c++) {
            const int i = c * SLAM_SCAN_BLOCK + threadIdx.x;
            if (i < n) {
                const float xi = __ldcg(x + i);
                __stcg(x + i, use_max ? fmaxf(chunk_tot[c - 1], xi) : __fadd_rn(chunk_tot[c - 1], xi));
            }
        }
    }
    __syncthreads();
}

// Phase D.  which == 0: the CDF side (Tc in place, csum, bm); which == 1: nEff and the decision.
__device__ __forceinline__ void weights_phase_d(const WeightsArgs &a, int which, float ws, float *sh, float *chunk_tot,
                                                float *bcast) {
    const int n = a.nbg;
    if (which == 0) {
        const float csum = upper_total(a.tot_c, n, sh, chunk_tot, bcast);  // before the scan overwrites the totals
        upper_scan_inplace(a.tot_c, n, false, sh, chunk_tot);
        for (int b = threadIdx.x; b < n; b += blockDim.x) {
            const float l = __ldcg(a.lmax + b);
            __stcg(a.bm + b, b > 0 ? __fadd_rn(__ldcg(a.tot_c + b - 1), l) : l);
        }
        upper_scan_inplace(a.bm, n, true, sh, chunk_tot);
        if (threadIdx.x == 0) {
            a.sc->ws = ws;
            a.sc->csum = csum;
            if (weights_degenerate(a, ws)) a.sc->weight_resets++;
        }
    } else {
        const float s2 = upper_total(a.tot_s, n, sh, chunk_tot, bcast);
        if (threadIdx.x == 0) {
            const float neff = __fdiv_rn(1.0f, s2);  // nEff = 1 / sum(w^2), core.cpp:788-791
            a.sc->s2 = s2;
            a.sc->neff = neff;
            a.sc->did = (a.do_resample && neff < a.neff_threshold) ? 1 : 0;  // core.cpp:739
        }
    }
}

// Grid-wide barrier of the cooperative launch, optionally with an exchange in the middle: every CTA arrives, CTA 0
// waits for all arrivals, runs the exchange (peer stores + wait for the peers), then releases the grid.
__device__ __forceinline__ unsigned gridbar_load(const unsigned *p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void grid_barrier(const GridBar &g, int k, const XchgArgs *x) {
    const unsigned G = gridDim.x;
    const unsigned arrived = g.base + (unsigned) k * (G + 1u) + G, released = arrived + 1u;
    __syncthreads();
    if (blockIdx.x == 0) {
        if (threadIdx.x == 0) {
            __threadfence();
            atomicAdd(g.ctr, 1u);
            while ((int) (gridbar_load(g.ctr) - arrived) < 0) {
            }
            __threadfence();
        }
        __syncthreads();
        if (x) xchg_body(*x);
        if (threadIdx.x == 0) {
            __threadfence();
            atomicAdd(g.ctr, 1u);
        }
        __syncthreads();
    } else {
        if (threadIdx.x == 0) {
            __threadfence();
            atomicAdd(g.ctr, 1u);
            while ((int) (gridbar_load(g.ctr) - released) < 0) {
            }
            __threadfence();
        }
        __syncthreads();
    }
}
#define SLAM_GRIDBAR_TICKS(G) (2u * ((G) + 1u))  // what one k_weights_fused launch adds to the counter

__global__ void __launch_bounds__(SLAM_SCAN_BLOCK, 1) k_weights_fused(const WeightsArgs a, const GridBar g,
                                                                      const XchgArgs x1, const XchgArgs x2) {
    __shared__ float sh[32];
    __shared__ float chunk_tot[SLAM_SCAN_BLOCK];
    __shared__ float bcast;
    for (int lb = blockIdx.x; lb < a.nb; lb += gridDim.x) weights_phase_a(a, lb, sh);
    grid_barrier(g, 0, x1.world > 1 ? &x1 : nullptr);
    const float ws = upper_total(a.tot_w, a.nbg, sh, chunk_tot, &bcast);
    for (int lb = blockIdx.x; lb < a.nb; lb += gridDim.x) weights_phase_c(a, lb, ws, sh);
    grid_barrier(g, 1, x2.world > 1 ? &x2 : nullptr);
    const int second = gridDim.x > 1 ? 1 : 0;
    if (blockIdx.x == 0) weights_phase_d(a, 0, ws, sh, chunk_tot, &bcast);
    if ((int) blockIdx.x == second) weights_phase_d(a, 1, ws, sh, chunk_tot, &bcast);
}
// The same pass as three launches (the exchanges are k_xchg launches in between).
__global__ void __launch_bounds__(SLAM_SCAN_BLOCK) k_weights_a(const WeightsArgs a) {
    __shared__ float sh[32];
    weights_phase_a(a, blockIdx.x, sh);
}
__global__ void __launch_bounds__(SLAM_SCAN_BLOCK) k_weights_c(const WeightsArgs a) {
    __shared__ float sh[32];
    __shared__ float chunk_tot[SLAM_SCAN_BLOCK];
    __shared__ float bcast;
    const float ws = upper_total(a.tot_w, a.nbg, sh, chunk_tot, &bcast);
    weights_phase_c(a, blockIdx.x, ws, sh);
}
__global__ void __launch_bounds__(SLAM_SCAN_BLOCK) k_weights_d(const WeightsArgs a) {
    __shared__ float sh[32];
    __shared__ float chunk_tot[SLAM_SCAN_BLOCK];
    __shared__ float bcast;
    const float ws = upper_total(a.tot_w, a.nbg, sh, chunk_tot, &bcast);
    weights_phase_d(a, blockIdx.x, ws, sh, chunk_tot, &bcast);
}

// Step F: select[] (stratifiedRandom core.cpp:751-769; di precomputed by the reference's own fp32 accumulation
// on the host) and keep[] (the merge at core.cpp:800-806) as a search.
//
// G(c) = first global index i with cdf[i] > select[c] (two-level: block maxima prefix BM, then the in-block running
// maximum).  The sequential merge gives keep[c] = first i >= keep[c-1] with cdf[i] > select[c]; with
// K[c] = max(G[c], G[c-1]) this equals K[c] whenever, for every c, either G[c] >= K[c-1], or K[c] == K[c-1] and
// cdf[K[c-1]] > select[c] (induction on c).  The kernel checks exactly that and raises *anomaly otherwise, in
// which case the host runs the exact sequential merge (k_keep_sequential).  keep = -1 where the reference leaves -1.
struct SearchArgs {
    const float *di;       // [n_local] strata midpoints for this device's global slots
    const double *unif;    // [n_local] (2-slot halo in front), or nullptr: the device generator (slam_rng.cuh)
    unsigned long long seed;
    unsigned draw;
    const float *cdf[SLAM_MAX_RANKS];   // [n_global]-indexed per rank; rank r = index / n_local holds that block: the
    const float *mloc[SLAM_MAX_RANKS];  // IN-BLOCK CDF and its running maximum (k_weights_*), to which the block prefix
                                        // Tc[b-1] is added on the fly.  When sharded these are the peers' buffers
                                        // themselves (CUDA IPC over NVLink): a slot's search lands on its own rank
                                        // unless the weights are badly skewed
    int world;
    const float *Tc;       // [nb_global] blocked scan of the block totals of the CDF
    const float *BM;       // [nb_global] prefix maxima of the block maxima
    float *sel;            // [n_local] out
    int *keep;             // [n_local] out (global indices)
    int *anomaly;          // out flag
    int force_anomaly;     // test hook (SLAMGPU_FORCE_ANOMALY): report an unproven slot regardless
    size_t n_local, n_global, slot_offset;
    int nb_global;
    float k, khalf;        // 1/N and k/2 as the reference forms them
};

// c_local may be -1 / -2: the two slots before this device's first one (di and unif carry a 2-slot halo in front).
__device__ __forceinline__ float select_value(const SearchArgs &a, long long c_local) {
    // *diter + unifRand()*k - k/2 evaluated in double, stored as float (core.cpp:767)
    const double u = a.unif ? a.unif[c_local]
                            : rng_unif(a.seed, (unsigned long long) ((long long) a.slot_offset + c_local), a.draw);
    const double t = __dadd_rn((double) a.di[c_local], __dmul_rn(u, (double) a.k));
    return __double2float_rn(__dsub_rn(t, (double) a.khalf));
}
// The final CDF / running maximum at global index i of block blk: block prefix + in-block value, one fp32 add exactly
// as the blocked scan forms it (block 0 has no prefix).
__device__ __forceinline__ float with_prefix(const SearchArgs &a, long long blk, float v) {
    return blk > 0 ? __fadd_rn(a.Tc[blk - 1], v) : v;
}

__device__ __forceinline__ long long first_exceed(const SearchArgs &a, float s) {
    // first block whose prefix-max exceeds s
    int lo = 0, hi = a.nb_global;  // answer in [lo, hi]; hi == nb_global means none
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (a.BM[mid] > s) hi = mid; else lo = mid + 1;
    }
    if (lo == a.nb_global) return (long long) a.n_global;  // sentinel: never assigned
    const size_t b0 = (size_t) lo * SLAM_SCAN_BLOCK;
    const int len = (int) min((size_t) SLAM_SCAN_BLOCK, a.n_global - b0);
    // a 1024-block never straddles two ranks (n_local is a multiple of 1024 when sharded)
    const float *mloc = a.mloc[a.world > 1 ? (int) (b0 / a.n_local) : 0];
    int l = 0, h = len - 1;  // the block's last running max exceeds s, so an answer exists
    while (l < h) {
        const int mid = (l + h) >> 1;
        if (with_prefix(a, lo, mloc[b0 + mid]) > s) h = mid; else l = mid + 1;
    }
    return (long long) (b0 + l);
}

// The monotone key the two-level search of first_exceed() walks, as one function of the global index: the running
// maximum of the CDF = max(prefix maximum of the earlier blocks, in-block running maximum).
__device__ __forceinline__ float running_max_at(const SearchArgs &a, long long i) {
    const long long blk = i / SLAM_SCAN_BLOCK;
    const float m = with_prefix(a, blk, a.mloc[a.world > 1 ? (int) ((size_t) i / a.n_local) : 0][i]);
    return blk > 0 ? fmaxf(a.BM[blk - 1], m) : m;
}

// first_exceed() by a whole warp: 32 probes per round trip instead of one (two rounds over <= 1024 block maxima, two
// over the 1024 running maxima of the block) — the dependent-load chain of the CTA leaders below is what bounds the
// kernel.  Same answer: both walk the same monotone keys.  All lanes must call it; all return the result.
template <typename Key>
__device__ __forceinline__ int warp_first_true(int lo, int hi, Key key) {
    // answer in [lo, hi]; key(i) is false below the answer and true from it on; hi itself is known (or defined) true
    const int lane = threadIdx.x & 31;
    while (lo < hi) {
        const int span = hi - lo, step = (span + 31) >> 5;
        const int idx = min(lo + (lane + 1) * step - 1, hi - 1);
        const unsigned b = __ballot_sync(0xffffffffu, key(idx));
        if (b == 0) {
            lo = hi;
        } else {
            const int f = __ffs(b) - 1;
            const int top = min(lo + (f + 1) * step - 1, hi - 1);
            lo = (f == 0) ? lo : min(lo + f * step - 1, hi - 1) + 1;
            hi = top;
        }
    }
    return lo;
}
__device__ __forceinline__ long long warp_first_exceed(const SearchArgs &a, float s) {
    const float *BM = a.BM;
    const int blk = warp_first_true(0, a.nb_global, [&](int i) { return BM[i] > s; });
    if (blk == a.nb_global) return (long long) a.n_global;
    const size_t b0 = (size_t) blk * SLAM_SCAN_BLOCK;
    const int len = (int) min((size_t) SLAM_SCAN_BLOCK, a.n_global - b0);
    const float *mloc = a.mloc[a.world > 1 ? (int) (b0 / a.n_local) : 0] + b0;
    return (long long) b0 + warp_first_true(0, len - 1, [&](int i) { return with_prefix(a, blk, mloc[i]) > s; });
}

// select[] is non-decreasing over a CTA's 256 slots, so every G of the CTA lies between the G of its first and of its
// last slot: two threads find those with the full two-level search, the CTA stages the running maxima of that window
// (a few hundred entries when the weights are not degenerate) in shared memory and every thread searches there —
// ~20 dependent L2 round trips per slot become ~10 shared-memory steps.  A window beyond SLAM_SEARCH_WINDOW falls back
// to the per-thread search.  Same G either way: the first index whose running maximum exceeds select[c].
#define SLAM_SEARCH_WINDOW 2048
__global__ void __launch_bounds__(256) k_search_keep(SearchArgs a) {
    __shared__ long long g_sh[256 + 2];
    __shared__ long long win[2];
    __shared__ float swin[2];
    __shared__ float W[SLAM_SEARCH_WINDOW];
    const size_t c = (size_t) blockIdx.x * blockDim.x + threadIdx.x;  // local slot
    const size_t c0 = (size_t) blockIdx.x * blockDim.x;
    const long long SENT = (long long) a.n_global;
    float s = 0.0f;
    long long g = -1;
    if (c < a.n_local) {
        s = select_value(a, (long long) c);
        a.sel[c] = s;
    }
    const int n_here = (int) min((size_t) blockDim.x, a.n_local - c0);  // live slots of this CTA (>= 1)
    {
        // warps 0 and 1: G of the two slots before this CTA's first one (the halo; none before global slot 0).  Slots
        // of the previous device are recomputed from its di/unif: across a device boundary the host passes di/unif
        // with a 2-slot halo in front (see capi), so local index c0 - 2 + t is valid whenever the global slot exists.
        // warps 2 and 3: G of the CTA's first and last slot = the window every other G lies in.
        const int wid = threadIdx.x >> 5;
        if (wid < 2) {
            const long long gc = (long long) (a.slot_offset + c0) - 2 + wid;
            long long gh = -1;
            if (gc >= 0) {
                const long long cl = (long long) c0 - 2 + wid;  // may be -1 / -2: halo stored before index 0
                gh = warp_first_exceed(a, select_value(a, cl));
            }
            if ((threadIdx.x & 31) == 0) g_sh[wid] = gh;
        } else if (wid < 4) {
            const float sw = select_value(a, (long long) (wid == 2 ? c0 : c0 + n_here - 1));
            const long long gw = warp_first_exceed(a, sw);
            if ((threadIdx.x & 31) == 0) {
                swin[wid - 2] = sw;
                win[wid - 2] = gw;
            }
        }
    }
    __syncthreads();
    const long long w0 = win[0], w1 = min(win[1], SENT - 1);  // G of the last slot may be "none" (SENT)
    const long long wlen = w1 - w0 + 1;
    if (w0 < SENT && wlen <= SLAM_SEARCH_WINDOW) {
        for (long long j = threadIdx.x; j < wlen; j += blockDim.x) W[j] = running_max_at(a, w0 + j);
        __syncthreads();
        if (c < a.n_local && (s < swin[0] || s > swin[1])) {
            // select[] dipped by an ulp across a stratum boundary (u == 1.0 next to u == 0.0): outside the window's proof
            g = first_exceed(a, s);
        } else if (c < a.n_local) {
            int lo = 0, hi = (int) wlen;  // first j with W[j] > s; hi == wlen: none inside the window
            while (lo < hi) {
                const int mid = (lo + hi) >> 1;
                if (W[mid] > s) hi = mid; else lo = mid + 1;
            }
            // none inside the window can only mean "none at all": the window ends at the last slot's G or at the
            // last particle
            g = (lo < (int) wlen) ? w0 + lo : SENT;
        }
    } else if (c < a.n_local) {
        g = (w0 >= SENT) ? SENT : first_exceed(a, s);
    }
    g_sh[threadIdx.x + 2] = g;
    __syncthreads();
    if (c >= a.n_local) return;
    const long long g1 = g_sh[threadIdx.x + 1], g2 = g_sh[threadIdx.x];  // G[c-1], G[c-2]  (-1 if absent)
    const long long kc = max(g, g1);
    const long long kprev = max(g1, g2);  // K[c-1]; -1 when c is global slot 0
    bool ok;
    if (g >= kprev) ok = true;
    else
        ok = (kc == kprev) &&
             (kprev == SENT || with_prefix(a, kprev / SLAM_SCAN_BLOCK,
                                           a.cdf[a.world > 1 ? (int) ((size_t) kprev / a.n_local) : 0][kprev]) > s);
    if (!ok || a.force_anomaly) atomicExch(a.anomaly, 1);
    a.keep[c] = (kc == SENT) ? -1 : (int) kc;
}

// The exact sequential merge of core.cpp:800-806 (fallback when k_search_keep reports an anomaly, and the tail of
// REFERENCE mode).  One thread; sel/cdf/keep global.
__global__ void k_keep_sequential(const float *__restrict__ sel, const float *__restrict__ cdf, int *__restrict__ keep,
                                  size_t n) {
    if (blockIdx.x != 0 || threadIdx.x != 0) return;
    for (size_t i = 0; i < n; i++) keep[i] = -1;
    size_t ctr = 0;
    for (size_t i = 0; i < n; i++) {
        const float ci = cdf[i];
        while (ctr < n && sel[ctr] < ci) {
            keep[ctr] = (int) i;
            ctr++;
        }
    }
}

// The final CDF of this device's slots (block prefix + in-block value), for the sequential merge.
__global__ void __launch_bounds__(256) k_cdf_final(const float *__restrict__ cdf_loc, const float *__restrict__ Tc,
                                                   long long slot0, size_t n, float *__restrict__ out) {
    const size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const long long blk = (slot0 + (long long) i) / SLAM_SCAN_BLOCK;
    out[i] = blk > 0 ? __fadd_rn(Tc[blk - 1], cdf_loc[i]) : cdf_loc[i];
}

// The sequential merge, sharded: this rank's slots only, continuing from the CDF position the previous rank ended at
// (slot c takes the first index >= the previous slot's index whose CDF exceeds select[c] — the same index may serve
// again — exactly the while loop of core.cpp:802-806).  The CDF is read wherever it lives (rank = index / n_local).
struct SeqArgs {
    const float *cdf[SLAM_MAX_RANKS];  // [n_global]-indexed per rank, final values in the owner's block
    const float *sel;                  // [n_local]
    int *keep;                         // [n_local]
    size_t n_local, n_global;
    const long long *start;            // nullptr: from index 0
    long long *end;
};
__global__ void k_keep_sequential_part(SeqArgs q) {
    if (blockIdx.x != 0 || threadIdx.x != 0) return;
    long long i = q.start ? *q.start : 0;
    const long long n = (long long) q.n_global;
    for (size_t c = 0; c < q.n_local; c++) {
        const float s = q.sel[c];
        while (i < n && !(s < q.cdf[(size_t) i / q.n_local][i])) i++;
        q.keep[c] = (i < n) ? (int) i : -1;
    }
    *q.end = i;
}

// unifRand() from raw rand() values (slamgpu_update_i32): one IEEE double division, as core.cpp:776.
__global__ void __launch_bounds__(256) k_ints_to_unif(const int *__restrict__ r, double *__restrict__ u, size_t n) {
    const size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) u[i] = __ddiv_rn((double) r[i], (double) SLAM_RAND_MAX);
}
// Throughput mode in the one-thread reference order: the generator's uniforms written out.
__global__ void __launch_bounds__(256) k_fill_unif(double *__restrict__ u, size_t n, unsigned long long seed,
                                                   unsigned long long slot0, unsigned draw) {
    const size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) u[i] = rng_unif(seed, slot0 + i, draw);
}

// REFERENCE order (SLAMGPU_SCAN_REFERENCE): one thread reproduces Eigen's SSE redux (Redux.h:193-248 with the SSE2
// predux of PacketMath.h:406-410) for both sums and the sequential fp32 running sum of cumulativeSum
// (core.cpp:813-824).  Meant for the reference's own particle counts (100 .. ~10^4).
__device__ float eigen_sum_seq(const float *v, long n) {
    const long ps = 4;
    const long size2 = (n / (2 * ps)) * (2 * ps);
    const long size1 = (n / ps) * ps;
    float res;
    if (size1) {
        float p0[4], p1[4];
        for (int j = 0; j < 4; j++) p0[j] = v[j];
        if (size1 > ps) {
            for (int j = 0; j < 4; j++) p1[j] = v[ps + j];
            for (long idx = 2 * ps; idx < size2; idx += 2 * ps)
                for (int j = 0; j < 4; j++) {
                    p0[j] = __fadd_rn(p0[j], v[idx + j]);
                    p1[j] = __fadd_rn(p1[j], v[idx + ps + j]);
                }
            for (int j = 0; j < 4; j++) p0[j] = __fadd_rn(p0[j], p1[j]);
            if (size1 > size2)
                for (int j = 0; j < 4; j++) p0[j] = __fadd_rn(p0[j], v[size2 + j]);
        }
        res = __fadd_rn(__fadd_rn(p0[0], p0[2]), __fadd_rn(p0[1], p0[3]));
        for (long idx = size1; idx < n; idx++) res = __fadd_rn(res, v[idx]);
    } else {
        res = v[0];
        for (long idx = 1; idx < n; idx++) res = __fadd_rn(res, v[idx]);
    }
    return res;
}

__global__ void k_reference_weights(float *__restrict__ w, size_t n, const float *__restrict__ di,
                                    const double *__restrict__ unif, float k, float khalf, float *__restrict__ sq,
                                    float *__restrict__ cdf, float *__restrict__ sel, Scalars *__restrict__ sc,
                                    float neff_threshold, int do_resample, int keep_nan) {
    if (blockIdx.x != 0 || threadIdx.x != 0) return;
    const float ws = eigen_sum_seq(w, (long) n);
    const bool degenerate = !keep_nan && !(ws > 0.0f && ws <= 3.0e38f);  // see weights_degenerate()
    if (degenerate) sc->weight_resets++;
    for (size_t i = 0; i < n; i++) {
        const float wn = degenerate ? (float) (1.0 / (double) n) : __fdiv_rn(w[i], ws);
        w[i] = wn;
        sq[i] = __fmul_rn(wn, wn);
    }
    const float s2 = eigen_sum_seq(sq, (long) n);
    const float neff = __fdiv_rn(1.0f, s2);
    sc->ws = ws;
    sc->s2 = s2;
    sc->neff = neff;
    sc->did = (do_resample && neff < neff_threshold) ? 1 : 0;  // core.cpp:739
    float run = 0.0f;
    for (size_t i = 0; i < n; i++) {
        run = __fadd_rn(run, w[i]);
        cdf[i] = run;
        const double t = __dadd_rn((double) di[i], __dmul_rn(unif[i], (double) k));
        sel[i] = __double2float_rn(__dsub_rn(t, (double) khalf));
    }
}


// ---------------------------------------------------------------------------------------------------------
// K7  resampling the state: particles[i] = old[keep[i]], w = 1/N (core.cpp:736-748).  keep < 0 (undefined behaviour in
// the reference) clamps to the last particle.
//
// Poses (48 B per particle) are gathered eagerly into the second pose buffer.  The landmark maps (24 B x Nf per
// particle: 25 GB at 2^20 x 1024) are NOT: the reference copies all of them on every resample although one update
// touches only the K observed rows.  Here every resample appends its keep[] to a log of the last H epochs, every row
// remembers the epoch its columns belong to, and a row is re-columned when it is next needed:
//     column p of the up-to-date row = column anc(p) of the stored row,
//     anc = keep_{e+1} o keep_{e+2} o ... o keep_E   (k_chain walks the log; a composition of monotone maps)
// The result is the reference's, bit for bit (a copy commutes with being postponed); the traffic per resample drops
// from 2*Nf rows to the rows actually read plus a trickle of background rows that bounds the log.
// Sources may live on a peer device when particles are sharded (PeerMaps; rank = global index / P).
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_gather_pose(PeerMaps pm, int cur, float *__restrict__ dst_pose,
                                                     float *__restrict__ w, const int *__restrict__ keep, size_t P,
                                                     long long n_global, float new_w, int rank,
                                                     unsigned long long *__restrict__ remote_sources,
                                                     int *__restrict__ log_slot, const int *__restrict__ cond) {
    // cond: the device-side resample decision (Scalars::did).  The kernel is launched BEFORE the host has read nEff back
    // and does nothing when the decision went the other way, so the read-back overlaps it instead of preceding it.
    if (cond && *cond == 0) return;
    const size_t p = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    int remote = 0;
    if (p < P) {
        long long k = keep[p];
        if (log_slot) log_slot[p] = (int) k;  // this epoch's entry of the keep log (one device: no separate copy)
        if (k < 0) k = n_global - 1;
        const int r = (pm.world > 1) ? (int) (k / (long long) P) : 0;
        const size_t sidx = (size_t) (k - (long long) r * (long long) P);
        const float *src = pm.pose[cur][r];
#pragma unroll
        for (int c = 0; c < 12; c++) dst_pose[c * P + p] = src[c * P + sidx];
        w[p] = new_w;
        remote = (pm.world > 1 && r != rank);
    }
    if (pm.world > 1) {  // SLAMGPU_STAT_REMOTE_SOURCES: how many slots took their particle from another device
        const int n = __syncthreads_count(remote);
        if (threadIdx.x == 0 && n) atomicAdd(remote_sources, (unsigned long long) n);
    }
}

// SLAMGPU_DIAG: out[0] += slots whose ancestor lives outside [slot0, slot0 + P); out[1] += slots whose ancestor differs
// from the previous slot's (= distinct ancestors, the map being non-decreasing).
__global__ void __launch_bounds__(256) k_diag_ancestors(const int *__restrict__ anc, size_t P, long long slot0,
                                                        unsigned long long *__restrict__ out) {
    const size_t p = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    int remote = 0, fresh = 0;
    if (p < P) {
        const long long a = anc[p];
        remote = a >= 0 && (a < slot0 || a >= slot0 + (long long) P);
        fresh = p == 0 || anc[p - 1] != anc[p];
    }
    const int nr = __syncthreads_count(remote), nf = __syncthreads_count(fresh);
    if (threadIdx.x == 0) {
        if (nr) atomicAdd(out, (unsigned long long) nr);
        if (nf) atomicAdd(out + 1, (unsigned long long) nf);
    }
}

// anc[p] <- map_{n-1}[ ... map_1[ map_0[ start ] ] ... ], start = anc[p] (continuing a chain) or this device's global
// slot (start_identity).  A hop is one of the logged keep arrays (one resample back) or a COARSE map: the composition
// of SLAM_COARSE_STRIDE consecutive keep arrays, built once when the epoch counter passes a multiple of the stride
// (capi.cu resample_commit).  The host lists the hops, newest first (capi.cu plan_hops): single steps down to the next
// multiple of the stride, coarse hops across the middle, single steps for the rest — at most 2 * (stride - 1) + H /
// stride look-ups per particle for a row that fell H epochs behind, instead of H.  Every map holds n_global ints.
#define SLAM_CHAIN_HOPS 64
#define SLAM_COARSE_STRIDE 16
struct ChainHops {
    const int *map[SLAM_CHAIN_HOPS];
    int n;
};
__global__ void __launch_bounds__(256) k_chain(int *__restrict__ anc, size_t P, long long slot0, long long n_global,
                                               const ChainHops hops, int start_identity) {
    const size_t p = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P) return;
    long long a = start_identity ? slot0 + (long long) p : (long long) anc[p];
    for (int h = 0; h < hops.n; h++) {
        const long long k = hops.map[h][a];
        a = (k < 0) ? n_global - 1 : k;
    }
    anc[p] = (int) a;
}

// dst row [p] = src row [anc[p]] for a list of row moves; grid = (ceil(P/256), ceil(nrows/SLAM_GATHER_ROWS)).  Each
// thread keeps its source column in a register and moves SLAM_GATHER_ROWS rows (independent loads in flight).  anc
// is non-decreasing, so reads are monotone with duplicates served by L1/L2; writes are fully coalesced.
__global__ void __launch_bounds__(256) k_materialise_rows(PeerMaps pm, MapRing m, size_t P,
                                                          const int *__restrict__ anc, long long n_global,
                                                          const RowMoves mv, int nrows) {
    const size_t p = (size_t) spread_block(blockIdx.x, gridDim.x, pm.spread, SLAM_SPREAD_CHUNK_SLOTS / 256) * blockDim.x +
                     threadIdx.x;
    if (p >= P) return;
    long long k = anc[p];
    if (k < 0) k = n_global - 1;
    const int r = (pm.world > 1) ? (int) (k / (long long) P) : 0;
    const size_t sidx = (size_t) (k - (long long) r * (long long) P);
    const float2 *smean = pm.mean[r];
    const float4 *scov = pm.cov[r];
    const int i0 = blockIdx.y * SLAM_GATHER_ROWS;
    float2 a[SLAM_GATHER_ROWS];
    float4 c[SLAM_GATHER_ROWS];
    int dst[SLAM_GATHER_ROWS];
#pragma unroll
    for (int q = 0; q < SLAM_GATHER_ROWS; q++) {
        if (i0 + q < nrows) {
            const RowPair rp = mv.p[i0 + q];
            dst[q] = rp.dst;
            a[q] = smean[(size_t) rp.src * P + sidx];
            c[q] = scov[(size_t) rp.src * P + sidx];
        }
    }
#pragma unroll
    for (int q = 0; q < SLAM_GATHER_ROWS; q++) {
        if (i0 + q < nrows) {
            __stcs(&m.mean[(size_t) dst[q] * P + p], a[q]);
            __stcs(&m.cov[(size_t) dst[q] * P + p], c[q]);
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
// K8  pose estimate (ParticleSLAMWrapper.cpp:56-77): double sums of x and y, heading of the FIRST max-weight
// particle.  Two stages, fixed shape.
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void est_combine(EstPartial &a, const EstPartial &b) {
    a.sx += b.sx;
    a.sy += b.sy;
    if (b.wmax > a.wmax || (b.wmax == a.wmax && b.imax < a.imax)) {
        a.wmax = b.wmax;
        a.imax = b.imax;
    }
}
__device__ __forceinline__ EstPartial est_block_reduce(EstPartial v) {
    __shared__ EstPartial sh[32];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        EstPartial o;
        o.sx = __shfl_down_sync(0xffffffffu, v.sx, d);
        o.sy = __shfl_down_sync(0xffffffffu, v.sy, d);
        o.wmax = __shfl_down_sync(0xffffffffu, v.wmax, d);
        o.imax = __shfl_down_sync(0xffffffffu, v.imax, d);
        est_combine(v, o);
    }
    if (lane == 0) sh[wid] = v;
    __syncthreads();
    const int nw = blockDim.x >> 5;
    if (wid == 0) {
        EstPartial z;
        z.sx = 0; z.sy = 0; z.wmax = -INFINITY; z.imax = 0x7fffffffffffffffLL;
        v = (lane < nw) ? sh[lane] : z;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            EstPartial o;
            o.sx = __shfl_down_sync(0xffffffffu, v.sx, d);
            o.sy = __shfl_down_sync(0xffffffffu, v.sy, d);
            o.wmax = __shfl_down_sync(0xffffffffu, v.wmax, d);
            o.imax = __shfl_down_sync(0xffffffffu, v.imax, d);
            est_combine(v, o);
        }
    }
    return v;
}
// One launch: every CTA reduces its particles to a partial; the CTA that finishes last (ticket counter) combines the
// partials in index order — the same order whichever CTA that is — and resets the ticket for the next call.
// pick_b: nullable device flag (Scalars::did); when set and nonzero the poses are read from xv_b (the buffer the
// conditional pose gather has just filled) instead of xv_a.
__global__ void __launch_bounds__(256) k_estimate(const float *__restrict__ xv_a, const float *__restrict__ xv_b,
                                                  const int *__restrict__ pick_b, const float *__restrict__ w, size_t P,
                                                  long long offset, EstPartial *part, unsigned *ticket,
                                                  EstPartial *__restrict__ out) {
    __shared__ bool last;
    const float *__restrict__ xv = (pick_b && *pick_b) ? xv_b : xv_a;
    EstPartial v;
    v.sx = 0; v.sy = 0; v.wmax = -INFINITY; v.imax = 0x7fffffffffffffffLL;
    const size_t gtid = (size_t) blockIdx.x * blockDim.x + threadIdx.x, gstride = (size_t) gridDim.x * blockDim.x;
    size_t done = 0;
    if ((P & 3) == 0 && (reinterpret_cast<uintptr_t>(xv) & 15) == 0 && (reinterpret_cast<uintptr_t>(w) & 15) == 0) {
        // four particles per thread and trip, the three 16-byte loads independent of each other
        const float4 *x4 = reinterpret_cast<const float4 *>(xv), *y4 = reinterpret_cast<const float4 *>(xv + P),
                     *w4 = reinterpret_cast<const float4 *>(w);
        for (size_t q = gtid; q < P / 4; q += gstride) {
            const float4 a = x4[q], b = y4[q], c = w4[q];
            const float xs[4] = {a.x, a.y, a.z, a.w}, ys[4] = {b.x, b.y, b.z, b.w}, ws[4] = {c.x, c.y, c.z, c.w};
#pragma unroll
            for (int j = 0; j < 4; j++) {
                v.sx += xs[j];
                v.sy += ys[j];
                if (ws[j] > v.wmax) {  // within a thread indices only grow: the first maximum wins by itself
                    v.wmax = ws[j];
                    v.imax = offset + (long long) (4 * q + j);
                }
            }
        }
        done = P;
    }
    for (size_t p = done + gtid; p < P; p += gstride) {
        EstPartial o;
        o.sx = xv[p];
        o.sy = xv[P + p];
        o.wmax = w[p];
        o.imax = offset + (long long) p;
        est_combine(v, o);
    }
    v = est_block_reduce(v);
    if (threadIdx.x == 0) {
        part[blockIdx.x] = v;
        __threadfence();
        last = atomicAdd(ticket, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!last) return;
    __threadfence();
    v.sx = 0; v.sy = 0; v.wmax = -INFINITY; v.imax = 0x7fffffffffffffffLL;
    for (int i = threadIdx.x; i < (int) gridDim.x; i += blockDim.x) {
        const volatile EstPartial *q = part + i;  // written by other CTAs of this launch: no cached / hoisted reads
        EstPartial o;
        o.sx = q->sx; o.sy = q->sy; o.wmax = q->wmax; o.theta = 0.0f; o.imax = q->imax;
        est_combine(v, o);
    }
    v = est_block_reduce(v);
    if (threadIdx.x == 0) {
        const long long loc = v.imax - offset;
        v.theta = (loc >= 0 && loc < (long long) P) ? xv[2 * P + (size_t) loc] : 0.0f;
        *out = v;
        *ticket = 0;
    }
}

// The resample's tail in ONE launch: the conditional pose gather of k_gather_pose and the pose estimate of k_estimate.
// Every thread holds the (x, y, weight) its slot ends up with — the gathered particle's when the device-side decision
// (cond) says resample, its own otherwise — so the estimate's reduction runs on registers instead of re-reading 12 B
// per particle in a second kernel.  grid = ceil(P / 256); part holds one partial per CTA.
__global__ void __launch_bounds__(256) k_gather_pose_estimate(PeerMaps pm, int cur, float *__restrict__ dst_pose,
                                                              float *__restrict__ w, const int *__restrict__ keep,
                                                              size_t P, long long n_global, float new_w, int rank,
                                                              unsigned long long *__restrict__ remote_sources,
                                                              int *__restrict__ log_slot, const int *__restrict__ cond,
                                                              long long offset, EstPartial *part, unsigned *ticket,
                                                              EstPartial *__restrict__ out) {
    __shared__ bool last;
    const bool resample = *cond != 0;
    const size_t p = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    const float *own = pm.pose[cur][pm.world > 1 ? rank : 0];
    int remote = 0;
    EstPartial v;
    v.sx = 0; v.sy = 0; v.wmax = -INFINITY; v.imax = 0x7fffffffffffffffLL;
    if (p < P) {
        if (resample) {
            long long k = keep[p];
            if (log_slot) log_slot[p] = (int) k;
            if (k < 0) k = n_global - 1;
            const int r = (pm.world > 1) ? (int) (k / (long long) P) : 0;
            const size_t sidx = (size_t) (k - (long long) r * (long long) P);
            const float *src = pm.pose[cur][r];
            float val[12];
#pragma unroll
            for (int c = 0; c < 12; c++) val[c] = src[c * P + sidx];
#pragma unroll
            for (int c = 0; c < 12; c++) dst_pose[c * P + p] = val[c];
            w[p] = new_w;
            remote = (pm.world > 1 && r != rank);
            v.sx = val[0];
            v.sy = val[1];
            v.wmax = new_w;
        } else {
            v.sx = own[p];
            v.sy = own[P + p];
            v.wmax = w[p];
        }
        v.imax = offset + (long long) p;
    }
    if (pm.world > 1 && resample) {
        const int n = __syncthreads_count(remote);
        if (threadIdx.x == 0 && n) atomicAdd(remote_sources, (unsigned long long) n);
    }
    v = est_block_reduce(v);
    if (threadIdx.x == 0) {
        part[blockIdx.x] = v;
        __threadfence();
        last = atomicAdd(ticket, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!last) return;
    __threadfence();
    v.sx = 0; v.sy = 0; v.wmax = -INFINITY; v.imax = 0x7fffffffffffffffLL;
    for (int i = threadIdx.x; i < (int) gridDim.x; i += blockDim.x) {
        const volatile EstPartial *q = part + i;  // written by other CTAs of this launch
        EstPartial o;
        o.sx = q->sx; o.sy = q->sy; o.wmax = q->wmax; o.theta = 0.0f; o.imax = q->imax;
        est_combine(v, o);
    }
    v = est_block_reduce(v);
    if (threadIdx.x == 0) {
        const long long loc = v.imax - offset;
        const float *heading = (resample ? dst_pose : own) + 2 * P;  // dst_pose: written by other CTAs of this launch
        v.theta = (loc >= 0 && loc < (long long) P) ? __ldcg(heading + (size_t) loc) : 0.0f;
        *out = v;
        *ticket = 0;
    }
}

// Strided download helper: every stride-th particle's pose and landmark means, packed.
__global__ void __launch_bounds__(256) k_decimate(PoseSoA s, MapRing m, size_t P, size_t stride, size_t n, int Nf,
                                                  const int *__restrict__ phys, float *__restrict__ xv_out,
                                                  float2 *__restrict__ lm_out) {
    const size_t q = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n) return;
    const size_t p = q * stride;
    for (int c = 0; c < 3; c++) xv_out[c * n + q] = s.xv[c * P + p];
    if (lm_out)
        for (int l = 0; l < Nf; l++) {
            lm_out[(size_t) l * n + q] = m.mean[(size_t) phys[l] * P + p];
        }
}

// Poses and weights of an arbitrary list of particles, packed ([3][n] | [9][n] | [n]): parity checks that follow
// lineages through a 2^20-particle run read a few thousand poses per cycle without touching the landmark rows.
__global__ void __launch_bounds__(256) k_pick_poses(PoseSoA s, size_t P, const long long *__restrict__ idx, size_t n,
                                                    float *__restrict__ out) {
    const size_t j = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const size_t p = (size_t) idx[j];
    for (int c = 0; c < 3; c++) out[c * n + j] = s.xv[c * P + p];
    for (int c = 0; c < 9; c++) out[(3 + c) * n + j] = s.Pv[c * P + p];
    out[12 * n + j] = s.w[p];
}

// Synthetic-workload initialisation: replicate a tile of Pt particles over all P (dst[p] = tile[p % Pt]).
__global__ void __launch_bounds__(256) k_tile_planes(const float *__restrict__ tile, size_t Pt, float *__restrict__ dst,
                                                     size_t P, int planes) {
    const size_t p = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P) return;
    for (int c = 0; c < planes; c++) dst[c * P + p] = tile[c * Pt + p % Pt];
}
__global__ void __launch_bounds__(256) k_tile_rows(const float2 *__restrict__ tmean, const float4 *__restrict__ tcov,
                                                   size_t Pt, MapRing m, size_t P, const int *__restrict__ phys) {
    const size_t p = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P) return;
    const int l = blockIdx.y;  // row inside the uploaded chunk
    const int row = phys[l];  // physical row of the l-th row of this chunk
    m.mean[(size_t) row * P + p] = tmean[(size_t) l * Pt + p % Pt];
    m.cov[(size_t) row * P + p] = tcov[(size_t) l * Pt + p % Pt];
}

}  // namespace slam
