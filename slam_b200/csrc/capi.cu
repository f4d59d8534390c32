// capi.cu — the C ABI of include/slamgpu.h: handle, device memory, launch sequencing.
//
// Everything per-particle happens in the kernels of kernels.cuh; this file only owns buffers, stages the small
// per-step host inputs (observation set, controls), sequences launches on one stream and reads back the few
// scalars the caller needs (nEff, the resample decision).  There is no host compute path: every entry point that
// would need one fails with SLAMGPU_ENODEVICE when no CUDA device is present.
#include "../../include/slamgpu.h"

#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "handle.h"
#include "kernels.cuh"

using namespace slam;

static thread_local std::string g_err;
int slam_fail(int code, const std::string &msg) {
    g_err = msg;
    return code;
}
// ---- timing helpers -------------------------------------------------------------------------------------
static int span_begin(slamgpu *h, int which, TimedSpan *sp) {
    sp->which = -1;
    if (!h->timing) return SLAMGPU_OK;
    for (int i = 0; i < 2; i++) {
        cudaEvent_t e;
        if (!h->free_events.empty()) {
            e = h->free_events.back();
            h->free_events.pop_back();
        } else {
            CU(cudaEventCreate(&e));
        }
        (i == 0 ? sp->a : sp->b) = e;
    }
    sp->which = which;
    CU(cudaEventRecord(sp->a, h->stream));
    return SLAMGPU_OK;
}
static int timing_flush(slamgpu *h) {
    if (h->spans.empty()) return SLAMGPU_OK;
    CU(cudaStreamSynchronize(h->stream));
    for (auto &sp : h->spans) {
        float ms = 0;
        CU(cudaEventElapsedTime(&ms, sp.a, sp.b));
        h->t_ms[sp.which] += ms;
        h->t_events[sp.which]++;
        h->free_events.push_back(sp.a);
        h->free_events.push_back(sp.b);
    }
    h->spans.clear();
    return SLAMGPU_OK;
}
static int span_end(slamgpu *h, TimedSpan *sp) {
    if (sp->which < 0) return SLAMGPU_OK;
    CU(cudaEventRecord(sp->b, h->stream));
    h->spans.push_back(*sp);
    if (h->spans.size() > 8192) RET(timing_flush(h));
    return SLAMGPU_OK;
}
#define LAUNCHED(h, which)      \
    do {                        \
        (h)->launches++;        \
        (h)->t_launches[which]++; \
    } while (0)
#define KCHECK() CU(cudaGetLastError())

static inline unsigned grid_for(size_t n, int block) { return (unsigned) ((n + block - 1) / block); }

static int flush_predicts(slamgpu *h);

// ---- library-level -------------------------------------------------------------------------------------
extern "C" const char *slamgpu_last_error(void) { return g_err.c_str(); }
extern "C" const char *slamgpu_version(void) { return "slamgpu 0.2 sm_100a"; }

extern "C" void *slamgpu_host_alloc(uint64_t bytes) {
    void *p = nullptr;
    if (cudaHostAlloc(&p, bytes, cudaHostAllocDefault) != cudaSuccess) {
        g_err = "cudaHostAlloc failed";
        return nullptr;
    }
    return p;
}
extern "C" void slamgpu_host_free(void *p) {
    if (p) cudaFreeHost(p);
}

static void free_all(slamgpu *h) {
    if (!h) return;
    cudaSetDevice(h->cfg.device);
    if (h->log_stream) cudaStreamSynchronize(h->log_stream);  // an all-gather of the keep log may still be in flight
    if (h->comm) comm_destroy(h->comm);
    for (auto &sp : h->spans) {
        cudaEventDestroy(sp.a);
        cudaEventDestroy(sp.b);
    }
    for (auto e : h->free_events) cudaEventDestroy(e);
    for (auto e : h->obs_ev)
        if (e) cudaEventDestroy(e);
    for (auto e : {h->ev_in, h->ev_gauss, h->ev_unif, h->ev_keep, h->ev_log})
        if (e) cudaEventDestroy(e);
    if (h->copy_stream) cudaStreamDestroy(h->copy_stream);
    if (h->log_stream) cudaStreamDestroy(h->log_stream);
    void *dev[] = {h->pose[0], h->pose[1], h->w, h->lm_mean, h->lm_cov, h->gauss, h->unif_base, h->di_base, h->cdf,
                   h->mloc, h->sel, h->keep, h->tot_slab, h->xflags, h->d_sc, h->d_z,
                   h->d_rec_in, h->d_rec_out, h->d_est, h->keeplog, h->anc, h->d_phys, h->fb_list, h->fb_count,
                   h->gauss_flags, h->est_ticket, h->gbar, h->d_tail, h->d_chain, h->dbg_w, h->unif_i, h->coarse};
    for (void *p : dev)
        if (p) cudaFree(p);
    if (h->h_sc) cudaFreeHost(h->h_sc);
    if (h->h_obs) cudaFreeHost(h->h_obs);
    if (h->h_tail) cudaFreeHost(h->h_tail);
    if (h->h_ones) cudaFreeHost(h->h_ones);
    if (h->own_stream) cudaStreamDestroy(h->own_stream);
    delete h;
}

// (Re)allocates the buffers whose size depends on the GLOBAL particle count and fills di for this device's slots.
static int setup_global(slamgpu *h) {
    const size_t Pg = h->Pg;
    h->nb = (h->P + SLAM_SCAN_BLOCK - 1) / SLAM_SCAN_BLOCK;
    h->nbg = (Pg + SLAM_SCAN_BLOCK - 1) / SLAM_SCAN_BLOCK;
    void **bufs[] = {(void **) &h->cdf, (void **) &h->mloc, (void **) &h->tot_slab, (void **) &h->xflags};
    for (void **b : bufs)
        if (*b) {
            cudaFree(*b);
            *b = nullptr;
        }
    CU(cudaMalloc(&h->cdf, sizeof(float) * Pg));
    CU(cudaMalloc(&h->mloc, sizeof(float) * Pg));
    const size_t slab_words = SLAM_BLOB_BASE(h->nbg) + SLAM_XCHG_BLOB_SLOTS * SLAM_MAX_RANKS * SLAM_XCHG_BLOB_WORDS;
    CU(cudaMalloc(&h->tot_slab, sizeof(float) * slab_words));
    CU(cudaMemset(h->tot_slab, 0, sizeof(float) * slab_words));
    h->tot_w = h->tot_slab;
    h->tot_c = h->tot_slab + h->nbg;
    h->tot_s = h->tot_slab + 2 * h->nbg;
    h->lmax = h->tot_slab + 3 * h->nbg;
    h->bm = h->tot_slab + 4 * h->nbg;
    {   // the cooperative weights kernel: one CTA of 1024 threads per SM at most, all co-resident
        h->coop_grid = 0;
        int coop = 0, sms = 0, per_sm = 0;
        cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, h->cfg.device);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, h->cfg.device);
        if (coop && !getenv("SLAMGPU_WEIGHTS_SPLIT") &&
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_weights_fused, SLAM_SCAN_BLOCK, 0) == cudaSuccess &&
            per_sm >= 1)
            h->coop_grid = (int) std::min<size_t>(h->nb, (size_t) sms);
    }
    CU(cudaMalloc(&h->xflags, sizeof(unsigned long long) * SLAM_MAX_RANKS));
    CU(cudaMemset(h->xflags, 0, sizeof(unsigned long long) * SLAM_MAX_RANKS));
    if (h->keeplog) {
        cudaFree(h->keeplog);
        h->keeplog = nullptr;
    }
    CU(cudaMalloc(&h->keeplog, sizeof(int) * (size_t) h->H * Pg));
    // coarse maps: one per SLAM_COARSE_STRIDE epochs of the log (none for the short logs some tests use)
    if (h->coarse) {
        cudaFree(h->coarse);
        h->coarse = nullptr;
    }
    h->coarse_stride = (h->H >= 4 * SLAM_COARSE_STRIDE && !getenv("SLAMGPU_NO_COARSE")) ? SLAM_COARSE_STRIDE : 0;
    h->coarse_slots = h->coarse_stride ? h->H / h->coarse_stride : 0;
    h->coarse_epoch.assign((size_t) h->coarse_slots, -1);
    if (h->coarse_slots) CU(cudaMalloc(&h->coarse, sizeof(int) * (size_t) h->coarse_slots * Pg));
    // stratifiedRandom's deterministic midpoints, by the reference's own fp32 accumulation run exactly N times
    // (core.cpp:751-759; oracle so_stratified_di).  Only this device's slots (+2 halo) are kept.
    std::vector<float> di(h->P + 2, 0.0f);
    const float k = (float) (1.0 / (float) Pg);
    float t = k / 2;
    for (size_t i = 0; i < h->slot0 + h->P; i++) {
        if (i + 2 >= h->slot0) di[i + 2 - h->slot0] = t;
        t = t + k;
    }
    CU(cudaMemcpyAsync(h->di_base, di.data(), sizeof(float) * (h->P + 2), cudaMemcpyHostToDevice, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    return SLAMGPU_OK;
}

static int create_impl(const slamgpu_config *cfg, slamgpu *h) {
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return slam_fail(SLAMGPU_ENODEVICE, "no CUDA device: slamgpu has no CPU path");
    if (cfg->device < 0 || cfg->device >= ndev) return slam_fail(SLAMGPU_EINVAL, "device ordinal out of range");
    CU(cudaSetDevice(cfg->device));
    const size_t P = h->P;
    CU(cudaStreamCreateWithFlags(&h->own_stream, cudaStreamNonBlocking));
    h->stream = h->own_stream;
    for (int i = 0; i < 2; i++) CU(cudaMalloc(&h->pose[i], sizeof(float) * 12 * P));
    CU(cudaMalloc(&h->w, sizeof(float) * P));
    CU(cudaMalloc(&h->lm_mean, sizeof(float2) * (size_t) h->rows_total * P));
    CU(cudaMalloc(&h->lm_cov, sizeof(float4) * (size_t) h->rows_total * P));
    CU(cudaMalloc(&h->anc, sizeof(int) * P));
    CU(cudaMalloc(&h->fb_list, sizeof(int) * P));
    CU(cudaMalloc(&h->fb_count, sizeof(int)));
    CU(cudaMalloc(&h->d_phys, sizeof(int) * h->nf_cap));
    CU(cudaMalloc(&h->gauss, sizeof(float) * 3 * P));
    CU(cudaMalloc(&h->unif_base, sizeof(double) * (P + 2)));
    CU(cudaMalloc(&h->di_base, sizeof(float) * (P + 2)));
    CU(cudaMalloc(&h->sel, sizeof(float) * P));
    CU(cudaMalloc(&h->keep, sizeof(int) * P));
    CU(cudaMalloc(&h->d_sc, sizeof(Scalars)));
    CU(cudaMemset(h->d_sc, 0, sizeof(Scalars)));
    CU(cudaHostAlloc(&h->h_sc, sizeof(Scalars), cudaHostAllocDefault));
    memset(h->h_sc, 0, sizeof(Scalars));
    h->est_blocks = 148 * 4;
    CU(cudaMalloc(&h->d_est, sizeof(EstPartial) * std::max<size_t>(h->est_blocks, grid_for(P, 256))));
    CU(cudaMalloc(&h->est_ticket, sizeof(unsigned)));
    CU(cudaMemset(h->est_ticket, 0, sizeof(unsigned)));
    CU(cudaMalloc(&h->d_tail, sizeof(TailBlob)));
    CU(cudaMemset(h->d_tail, 0, sizeof(TailBlob)));
    CU(cudaMalloc(&h->d_chain, sizeof(TailBlob)));
    CU(cudaMemset(h->d_chain, 0, sizeof(TailBlob)));
    CU(cudaHostAlloc(&h->h_tail, sizeof(TailBlob) * SLAM_MAX_RANKS, cudaHostAllocDefault));
    memset(h->h_tail, 0, sizeof(TailBlob) * SLAM_MAX_RANKS);
    CU(cudaMalloc(&h->gbar, sizeof(unsigned)));
    CU(cudaMemset(h->gbar, 0, sizeof(unsigned)));
    CU(cudaHostAlloc(&h->h_ones, sizeof(unsigned), cudaHostAllocDefault));
    *h->h_ones = 0xffffffffu;
    h->force_anomaly = getenv("SLAMGPU_FORCE_ANOMALY") != nullptr;
    {   // the FastSLAM2 fast kernels stage up to 2048 observations next to their landmark ring: > 48 KB
        const int big = (int) SLAM_FS2_FAST_SMEM(2048);
        CU(cudaFuncSetAttribute(k_fs2_update_fast<0, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
        CU(cudaFuncSetAttribute(k_fs2_update_fast<1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
        CU(cudaFuncSetAttribute(k_fs2_update_fast<2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
        CU(cudaFuncSetAttribute(k_fs2_update_fast<0, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
        CU(cudaFuncSetAttribute(k_fs2_update_fast<1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
        CU(cudaFuncSetAttribute(k_fs2_update_fast<2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
        CU(cudaFuncSetAttribute(k_fs2_update_listed<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
        CU(cudaFuncSetAttribute(k_fs2_update_listed<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
    }
    CU(cudaMalloc(&h->gauss_flags, sizeof(unsigned) * SLAM_GAUSS_CHUNKS));
    CU(cudaMemset(h->gauss_flags, 0, sizeof(unsigned) * SLAM_GAUSS_CHUNKS));
    h->obs_cap = 1024;
    CU(cudaMalloc(&h->d_z, (sizeof(float2) + 2 * sizeof(int)) * h->obs_cap));
    CU(cudaHostAlloc(&h->h_obs, (sizeof(float2) + 2 * sizeof(int)) * h->obs_cap * SLAM_OBS_SLOTS, cudaHostAllocDefault));
    // ParticleSLAMWrapper::initializeParticles (ParticleSLAMWrapper.cpp:17-25): w = 1/N, xv = 0, Pv = 0
    CU(cudaMemsetAsync(h->pose[0], 0, sizeof(float) * 12 * P, h->stream));
    CU(cudaMemsetAsync(h->unif_base, 0, sizeof(double) * (P + 2), h->stream));
    std::vector<float> w0(P, 1.0f / (float) P);
    CU(cudaMemcpyAsync(h->w, w0.data(), sizeof(float) * P, cudaMemcpyHostToDevice, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    return setup_global(h);
}

extern "C" int slamgpu_create(const slamgpu_config *cfg, slamgpu_t **out) {
    if (!cfg || !out) return slam_fail(SLAMGPU_EINVAL, "null argument");
    *out = nullptr;
    // below 2^28 per device: the update kernel addresses a landmark row as base + row * (P * 16 bytes) with a 32-bit
    // pitch (one row of one landmark would already be 4 GB there; 2^30 stays the bound on the GLOBAL count, comm_init)
    if (cfg->n_particles <= 0 || cfg->n_particles >= (1ll << 28)) return slam_fail(SLAMGPU_EINVAL, "n_particles out of range");
    if (cfg->nf_cap <= 0) return slam_fail(SLAMGPU_EINVAL, "nf_cap must be positive");
    if (cfg->method != SLAMGPU_FASTSLAM1 && cfg->method != SLAMGPU_FASTSLAM2)
        return slam_fail(SLAMGPU_EINVAL, "method must be SLAMGPU_FASTSLAM1 or SLAMGPU_FASTSLAM2");
    if (cfg->scan_mode != SLAMGPU_SCAN_REFERENCE && cfg->scan_mode != SLAMGPU_SCAN_BLOCKED)
        return slam_fail(SLAMGPU_EINVAL, "scan_mode must be SLAMGPU_SCAN_REFERENCE or SLAMGPU_SCAN_BLOCKED");
    if (cfg->math_mode != SLAMGPU_MATH_FAST && cfg->math_mode != SLAMGPU_MATH_REFERENCE_ORDER)
        return slam_fail(SLAMGPU_EINVAL, "math_mode must be SLAMGPU_MATH_FAST or SLAMGPU_MATH_REFERENCE_ORDER");
    slamgpu *h = new slamgpu();
    h->cfg = *cfg;
    h->P = h->Pg = (size_t) cfg->n_particles;
    h->nf_cap = cfg->nf_cap;
    h->spare = cfg->spare_rows > 0 ? cfg->spare_rows : std::min(64, cfg->nf_cap);
    h->rows_total = h->nf_cap + h->spare;
    h->phys.resize(h->nf_cap);
    for (int l = 0; l < h->nf_cap; l++) h->phys[l] = l;
    for (int r = 0; r < h->spare; r++) h->free_rows.push_back(h->nf_cap + r);
    h->row_epoch.assign(h->nf_cap, 0);
    h->H = cfg->log_epochs > 0 ? std::max(4, cfg->log_epochs) : 256;
    {   // trickle batch: H / 32 resamples apart (1 for the short logs the tests use); SLAMGPU_TRICKLE_BATCH overrides
        const char *e = getenv("SLAMGPU_TRICKLE_BATCH");
        int B = e ? atoi(e) : h->H / 32;
        h->trickle_batch = std::max(1, std::min(B, (h->H - 3) / 2));
    }
    int r = create_impl(cfg, h);
    if (r != SLAMGPU_OK) {
        std::string keep = g_err;
        free_all(h);
        g_err = keep;
        return r;
    }
    *out = h;
    return SLAMGPU_OK;
}

extern "C" void slamgpu_destroy(slamgpu_t *h) { free_all(h); }

extern "C" int slamgpu_set_stream(slamgpu_t *h, void *s) {
    if (!h) return slam_fail(SLAMGPU_EINVAL, "null handle");
    CU(cudaSetDevice(h->cfg.device));
    RET(flush_predicts(h));
    CU(cudaStreamSynchronize(h->stream));
    h->stream = s ? (cudaStream_t) s : h->own_stream;
    return SLAMGPU_OK;
}
extern "C" void *slamgpu_get_stream(slamgpu_t *h) { return h ? (void *) h->stream : nullptr; }
static int report_device_flags(slamgpu *h);
extern "C" int slamgpu_synchronize(slamgpu_t *h) {
    if (!h) return slam_fail(SLAMGPU_EINVAL, "null handle");
    CU(cudaSetDevice(h->cfg.device));
    RET(flush_predicts(h));
    if (h->world > 1)  // an exchange that timed out anywhere since the last report must not go unnoticed
        CU(cudaMemcpyAsync(h->h_sc, h->d_sc, sizeof(Scalars), cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    return report_device_flags(h);
}
extern "C" int slamgpu_landmark_count(slamgpu_t *h) { return h ? h->Nf : SLAMGPU_EINVAL; }
extern "C" int64_t slamgpu_launch_count(slamgpu_t *h) { return h ? h->launches : 0; }
extern "C" int64_t slamgpu_stat(slamgpu_t *h, int32_t which) {
    if (!h) return 0;
    switch (which) {
        case SLAMGPU_STAT_ROWS_MOVED: return h->stat_rows_moved;
        case SLAMGPU_STAT_RESAMPLES: return h->stat_resamples;
        case SLAMGPU_STAT_CHAIN_STEPS: return h->stat_chain_steps;
        case SLAMGPU_STAT_ROWS_FUSED: return h->stat_rows_fused;
        case SLAMGPU_STAT_REMOTE_SOURCES: return (int64_t) h->h_sc->remote_sources;  // as of the last read-back
        case SLAMGPU_STAT_KEEP_FALLBACKS: return h->stat_keep_fallbacks;
        case SLAMGPU_STAT_WEIGHT_RESETS: return (int64_t) h->h_sc->weight_resets;  // as of the last read-back
    }
    return 0;
}

// ---- cross-rank ordering without a barrier per kernel ------------------------------------------------------------------
// Every rank issues the same kernels in the same order with the same row tables, and a rank's stream passes an
// exchange (comm.cu k_xchg) only after every peer's stream has reached it.  So a kernel launched after exchange number
// X knows that every peer has completed all kernels it launched before ITS exchange X.  A kernel that reads rows of a
// peer (src) needs the peers' last writers of those rows complete; a kernel that overwrites rows (dst) needs the
// peers' last remote readers of its copy complete.  Both hold as soon as ONE exchange lies between the two launches:
// the tags below record the exchange count at the last write / remote read of every physical row and pose buffer, and
// a barrier is issued only when a tag equals the current count (rare: the resample's own three exchanges usually lie
// in between).  One device: nothing to do.
static int xchg(slamgpu *h, const int *arrays, int narr, bool halo, const void *blob = nullptr, int blob_bytes = 0,
                int blob_slot = 0) {
    RET(comm_exchange(h->comm, h, arrays, narr, halo, blob, blob_bytes, blob_slot));
    h->xcount++;
    return SLAMGPU_OK;
}
static int order_rows(slamgpu *h, const int *src, int ns, const int *dst, int nd) {
    if (h->world <= 1) return SLAMGPU_OK;
    if (h->row_written_x.size() != (size_t) h->rows_total) {
        h->row_written_x.assign(h->rows_total, h->xcount);  // unknown history: as if touched just now
        h->row_read_x.assign(h->rows_total, h->xcount);
    }
    bool need = false;
    for (int i = 0; i < ns && !need; i++) need = h->row_written_x[src[i]] >= h->xcount;
    for (int i = 0; i < nd && !need; i++)
        need = h->row_read_x[dst[i]] >= h->xcount;
    if (need) RET(xchg(h, nullptr, 0, false));
    for (int i = 0; i < ns; i++) h->row_read_x[src[i]] = h->xcount;
    for (int i = 0; i < nd; i++) h->row_written_x[dst[i]] = h->xcount;
    return SLAMGPU_OK;
}
// After state was written from the host (uploads): every row and both pose buffers count as written just now.
static void order_reset(slamgpu *h) {
    h->row_written_x.clear();
    h->row_read_x.clear();
    for (int b = 0; b < 2; b++) h->pose_written_x[b] = h->pose_read_x[b] = h->xcount;
}
// Before a kernel that writes pose buffer `wbuf` (and, when rbuf >= 0, reads the peers' pose buffer rbuf).
static int order_pose(slamgpu *h, int wbuf, int rbuf) {
    if (h->world <= 1) return SLAMGPU_OK;
    bool need = h->pose_read_x[wbuf] >= h->xcount;
    if (rbuf >= 0) need = need || h->pose_written_x[rbuf] >= h->xcount;
    if (need) RET(xchg(h, nullptr, 0, false));
    h->pose_written_x[wbuf] = h->xcount;
    if (rbuf >= 0) h->pose_read_x[rbuf] = h->xcount;
    return SLAMGPU_OK;
}

// Where a kernel's standard normals come from: the staged buffer, or (throughput mode) the device generator.
static NoiseSrc noise_src(slamgpu *h, const float *staged, unsigned draw) {
    NoiseSrc n;
    n.gauss = staged;
    n.seed = h->rng_seed;
    n.slot0 = (unsigned long long) h->slot0;
    n.draw = draw;
    n.rng = staged ? 0 : 1;
    return n;
}

// ---- deferred predict ticks ---------------------------------------------------------------------------------------
static int flush_predicts(slamgpu *h) {
    if (h->predict_queue.empty()) return SLAMGPU_OK;
    h->est_valid = false;
    PredictBatch b;
    b.n = (int) h->predict_queue.size();
    for (int t = 0; t < b.n; t++) b.a[t] = h->predict_queue[t];
    h->predict_queue.clear();
    RET(order_pose(h, h->cur, -1));
    TimedSpan sp;
    RET(span_begin(h, SLAMGPU_T_PREDICT, &sp));
    if (h->cfg.math_mode == SLAMGPU_MATH_FAST) {
        PredictBatchFast f;
        f.n = b.n;
        f.with_cov = b.a[0].method == 2;
        f.use_heading = b.a[0].use_heading;
        for (int t = 0; t < b.n; t++) {
            const PredictArgs &a = b.a[t];
            PredictTick &k = f.t[t];
            k.vdt = a.V * a.dt;
            k.dt = a.dt;
            k.dtheta = a.V * a.dt * sinf(a.G / a.wheel_base);  // the reference's sin(G/WB), fastslam2.cpp:103
            k.G = a.G;
            k.gu20 = a.dt * sinf(a.G) / a.wheel_base;
            k.gu21 = a.V * a.dt * cosf(a.G) / a.wheel_base;
            for (int j = 0; j < 4; j++) k.Q[j] = a.Q[j];
            k.phi_true = a.phi_true;
            k.R_phi = a.sigma_phi * a.sigma_phi;
        }
        k_predict_batch_fast<<<grid_for(h->P, 256), 256, 0, h->stream>>>(h->soa(h->cur), h->P, f);
    } else
        k_predict_batch<<<grid_for(h->P, 256), 256, 0, h->stream>>>(h->soa(h->cur), h->P, b);
    KCHECK();
    LAUNCHED(h, SLAMGPU_T_PREDICT);
    RET(span_end(h, &sp));
    return SLAMGPU_OK;
}

// ---- landmark row pool: deferred resampling of the maps (kernels.cuh K7) ---------------------------------------
// Chunk stride of spread_block() for a grid of nblocks CTAs of `threads` slots (0 = launch order, when the grid does not
// split into whole chunks): about 0.618 x the chunk count, made coprime to it.
static unsigned spread_stride(unsigned nblocks, unsigned threads) {
    const unsigned per = SLAM_SPREAD_CHUNK_SLOTS / threads;
    if (nblocks % per != 0 || nblocks / per < 8) return 0;
    const unsigned nch = nblocks / per;
    unsigned s = (unsigned) (nch * 0.6180339887) | 1u;
    auto gcd = [](unsigned a, unsigned b) {
        while (b) {
            const unsigned t = a % b;
            a = b;
            b = t;
        }
        return a;
    };
    while (gcd(s, nch) != 1) s += 2;
    return s % nch;
}

static int peer_maps(slamgpu *h, PeerMaps *pm) {
    if (h->world > 1) {
        RET(comm_peer_maps(h->comm, h, pm));
        pm->spread = spread_stride(grid_for(h->P, 256), 256);
        return SLAMGPU_OK;
    }
    memset(pm, 0, sizeof(*pm));
    pm->world = 1;
    pm->pose[0][0] = h->pose[0];
    pm->pose[1][0] = h->pose[1];
    pm->mean[0] = h->lm_mean;
    pm->cov[0] = h->lm_cov;
    pm->cdf[0] = h->cdf;
    pm->mloc[0] = h->mloc;
    return SLAMGPU_OK;
}

// anc <- the ancestor map from epoch `from` back to epoch `to` (to < from <= h->epoch), continuing the map already in
// anc (identity == false: anc then maps the current epoch to `from`) or starting at this device's own slots.
// Hops: logged keep arrays, and the coarse maps across whole strides (kernels.cuh k_chain).
static int launch_chain(slamgpu *h, int *anc, long long from, long long to, bool identity) {
    const size_t Pg = h->Pg;
    const int G = h->coarse_stride;
    long long cur = from;
    bool first = true;
    while (cur > to) {
        ChainHops hops;
        hops.n = 0;
        while (cur > to && hops.n < SLAM_CHAIN_HOPS) {
            if (G > 0 && cur % G == 0 && cur - G >= to && h->coarse_epoch[(size_t) ((cur / G) % h->coarse_slots)] == cur) {
                hops.map[hops.n++] = h->coarse + (size_t) ((cur / G) % h->coarse_slots) * Pg;
                cur -= G;
            } else {
                hops.map[hops.n++] = h->keeplog + (size_t) (cur % h->H) * Pg;
                cur -= 1;
            }
        }
        k_chain<<<grid_for(h->P, 256), 256, 0, h->stream>>>(anc, h->P, (long long) h->slot0, (long long) Pg, hops,
                                                            (identity && first) ? 1 : 0);
        KCHECK();
        LAUNCHED(h, SLAMGPU_T_GATHER);
        h->stat_chain_steps += hops.n;
        first = false;
    }
    return SLAMGPU_OK;
}

// Brings the listed logical rows up to the current resample epoch: column p <- column anc(p) of the stored row.
// Rows are grouped by the epoch they belong to (newest first, so that one ancestor chain is extended group by group)
// and moved in chunks into free physical rows; the vacated rows return to the pool.
static int materialise(slamgpu *h, const std::vector<int> &want) {
    std::vector<std::pair<long long, int>> stale;
    for (int l : want)
        if (h->row_epoch[l] < h->epoch) stale.push_back({h->row_epoch[l], l});
    if (stale.empty()) return SLAMGPU_OK;
    std::sort(stale.begin(), stale.end(), [](const std::pair<long long, int> &a, const std::pair<long long, int> &b) {
        return a.first != b.first ? a.first > b.first : a.second < b.second;
    });
    if (h->epoch - stale.back().first >= h->H)
        return slam_fail(SLAMGPU_EINVAL, "internal: a landmark row fell out of the resample log");
    PeerMaps pm;
    RET(peer_maps(h, &pm));
    const size_t P = h->P, Pg = h->Pg;
    TimedSpan sp;
    RET(span_begin(h, SLAMGPU_T_GATHER, &sp));
    bool identity = true;      // h->anc not started yet
    long long cur_e = h->epoch;  // epoch h->anc currently maps back to
    size_t i = 0;
    while (i < stale.size()) {
        const long long e = stale[i].first;
        size_t j = i;
        while (j < stale.size() && stale[j].first == e) j++;
        const int *anc_ptr;
        if (identity && e == h->epoch - 1) {
            anc_ptr = h->keeplog + (size_t) (h->epoch % h->H) * Pg + h->slot0;  // one epoch back: keep[] itself
        } else {
            RET(launch_chain(h, h->anc, cur_e, e, identity));
            identity = false;
            cur_e = e;
            anc_ptr = h->anc;
        }
        size_t k = i;
        while (k < j) {
            const int n = (int) std::min<size_t>(std::min<size_t>(j - k, SLAM_MOVES_PER_LAUNCH), h->free_rows.size());
            if (n <= 0) return slam_fail(SLAMGPU_EINVAL, "internal: no spare landmark rows");
            RowMoves mv;
            std::vector<int> vacated(n);
            for (int q = 0; q < n; q++) {
                const int l = stale[k + q].second;
                mv.p[q].src = h->phys[l];
                mv.p[q].dst = h->free_rows.back();
                h->free_rows.pop_back();
                vacated[q] = h->phys[l];
                h->phys[l] = mv.p[q].dst;
                h->row_epoch[l] = h->epoch;
            }
            {   // sharded: sources are read on the peers, destinations may still be read by them (order_rows)
                std::vector<int> srcs(n), dsts(n);
                for (int q = 0; q < n; q++) {
                    srcs[q] = mv.p[q].src;
                    dsts[q] = mv.p[q].dst;
                }
                RET(order_rows(h, srcs.data(), n, dsts.data(), n));
            }
            dim3 grid(grid_for(P, 256), (n + SLAM_GATHER_ROWS - 1) / SLAM_GATHER_ROWS);
            k_materialise_rows<<<grid, 256, 0, h->stream>>>(pm, h->ring(), P, anc_ptr, (long long) Pg, mv, n);
            KCHECK();
            LAUNCHED(h, SLAMGPU_T_GATHER);
            h->stat_rows_moved += n;
            for (int q = 0; q < n; q++) h->free_rows.insert(h->free_rows.begin(), vacated[q]);
            k += n;
        }
        i = j;
    }
    RET(span_end(h, &sp));
    return SLAMGPU_OK;
}

static int materialise_all(slamgpu *h) {
    std::vector<int> all(h->Nf);
    for (int l = 0; l < h->Nf; l++) all[l] = l;
    return materialise(h, all);
}

// ---- state transfer ---------------------------------------------------------------------------------------
extern "C" int slamgpu_upload_particles(slamgpu_t *h, const float *xv, const float *Pv, const float *w,
                                        const float *lm_mean, const float *lm_cov, int32_t Nf) {
    if (!h) return slam_fail(SLAMGPU_EINVAL, "null handle");
    if (Nf > h->nf_cap) return slam_fail(SLAMGPU_ECAP, "Nf exceeds nf_cap");
    if (Nf < 0) Nf = (lm_mean || lm_cov) ? h->Nf : -1;  // negative: keep the current landmark count
    CU(cudaSetDevice(h->cfg.device));
    RET(flush_predicts(h));
    order_reset(h);
    h->est_valid = false;
    const size_t P = h->P;
    PoseSoA s = h->soa(h->cur);
    if (xv) CU(cudaMemcpyAsync(s.xv, xv, sizeof(float) * 3 * P, cudaMemcpyHostToDevice, h->stream));
    if (Pv) CU(cudaMemcpyAsync(s.Pv, Pv, sizeof(float) * 9 * P, cudaMemcpyHostToDevice, h->stream));
    if (w) CU(cudaMemcpyAsync(s.w, w, sizeof(float) * P, cudaMemcpyHostToDevice, h->stream));
    if ((lm_mean != nullptr) != (lm_cov != nullptr)) RET(materialise_all(h));  // half a row keeps its other half
    for (int l = 0; l < Nf; l++) {
        const size_t row = (size_t) h->phys_row(l);
        if (lm_mean && lm_cov) h->row_epoch[l] = h->epoch;
        if (lm_mean)
            CU(cudaMemcpyAsync(h->lm_mean + row * P, lm_mean + (size_t) l * P * 2, sizeof(float2) * P,
                               cudaMemcpyHostToDevice, h->stream));
        if (lm_cov)
            CU(cudaMemcpyAsync(h->lm_cov + row * P, lm_cov + (size_t) l * P * 4, sizeof(float4) * P,
                               cudaMemcpyHostToDevice, h->stream));
    }
    if (Nf >= 0) h->Nf = Nf;
    CU(cudaStreamSynchronize(h->stream));
    return SLAMGPU_OK;
}

extern "C" int slamgpu_upload_particles_tiled(slamgpu_t *h, int64_t Pt, const float *xv, const float *Pv,
                                              const float *w, const float *lm_mean, const float *lm_cov, int32_t Nf) {
    if (!h || Pt <= 0 || !xv || !Pv || !w) return slam_fail(SLAMGPU_EINVAL, "bad argument");
    if (Nf < 0 || Nf > h->nf_cap) return slam_fail(SLAMGPU_ECAP, "Nf exceeds nf_cap");
    if (Nf > 0 && (!lm_mean || !lm_cov)) return slam_fail(SLAMGPU_EINVAL, "landmark tile missing");
    CU(cudaSetDevice(h->cfg.device));
    RET(flush_predicts(h));
    order_reset(h);
    h->est_valid = false;
    const size_t P = h->P, T = (size_t) Pt;
    const int chunk = 64;  // landmark rows staged per copy
    float *dt = nullptr;
    CU(cudaMalloc(&dt, sizeof(float) * std::max((size_t) 13 * T, (size_t) chunk * T * 6)));
    int rc = SLAMGPU_OK;
    do {
        cudaError_t e = cudaMemcpyAsync(dt, xv, sizeof(float) * 3 * T, cudaMemcpyHostToDevice, h->stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(dt + 3 * T, Pv, sizeof(float) * 9 * T, cudaMemcpyHostToDevice, h->stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(dt + 12 * T, w, sizeof(float) * T, cudaMemcpyHostToDevice, h->stream);
        if (e != cudaSuccess) { rc = SLAMGPU_ECUDA; break; }
        k_tile_planes<<<grid_for(P, 256), 256, 0, h->stream>>>(dt, T, h->pose[h->cur], P, 12);
        k_tile_planes<<<grid_for(P, 256), 256, 0, h->stream>>>(dt + 12 * T, T, h->w, P, 1);
        h->launches += 2;
        if (cudaStreamSynchronize(h->stream) != cudaSuccess) { rc = SLAMGPU_ECUDA; break; }
        float2 *tm = (float2 *) dt;
        float4 *tc = (float4 *) (dt + (size_t) chunk * T * 2);
        for (int l0 = 0; l0 < Nf && rc == SLAMGPU_OK; l0 += chunk) {
            const int n = std::min(chunk, Nf - l0);
            e = cudaMemcpyAsync(tm, lm_mean + (size_t) l0 * T * 2, sizeof(float2) * n * T, cudaMemcpyHostToDevice, h->stream);
            if (e == cudaSuccess)
                e = cudaMemcpyAsync(tc, lm_cov + (size_t) l0 * T * 4, sizeof(float4) * n * T, cudaMemcpyHostToDevice, h->stream);
            if (e != cudaSuccess) { rc = SLAMGPU_ECUDA; break; }
            e = cudaMemcpyAsync(h->d_phys, h->phys.data() + l0, sizeof(int) * n, cudaMemcpyHostToDevice, h->stream);
            if (e != cudaSuccess) { rc = SLAMGPU_ECUDA; break; }
            k_tile_rows<<<dim3(grid_for(P, 256), n), 256, 0, h->stream>>>(tm, tc, T, h->ring(), P, h->d_phys);
            for (int l = l0; l < l0 + n; l++) h->row_epoch[l] = h->epoch;
            h->launches++;
            if (cudaStreamSynchronize(h->stream) != cudaSuccess) { rc = SLAMGPU_ECUDA; break; }
        }
    } while (0);
    cudaFree(dt);
    if (rc != SLAMGPU_OK) return slam_fail(rc, std::string("upload_particles_tiled: ") + cudaGetErrorString(cudaGetLastError()));
    h->Nf = Nf;
    return SLAMGPU_OK;
}

extern "C" int slamgpu_download_particles(slamgpu_t *h, float *xv, float *Pv, float *w, float *lm_mean,
                                          float *lm_cov, int32_t *Nf) {
    if (!h) return slam_fail(SLAMGPU_EINVAL, "null handle");
    CU(cudaSetDevice(h->cfg.device));
    RET(flush_predicts(h));
    const size_t P = h->P;
    if (lm_mean || lm_cov) RET(materialise_all(h));
    PoseSoA s = h->soa(h->cur);
    if (xv) CU(cudaMemcpyAsync(xv, s.xv, sizeof(float) * 3 * P, cudaMemcpyDeviceToHost, h->stream));
    if (Pv) CU(cudaMemcpyAsync(Pv, s.Pv, sizeof(float) * 9 * P, cudaMemcpyDeviceToHost, h->stream));
    if (w) CU(cudaMemcpyAsync(w, s.w, sizeof(float) * P, cudaMemcpyDeviceToHost, h->stream));
    for (int l = 0; l < h->Nf; l++) {
        const size_t row = (size_t) h->phys_row(l);
        if (lm_mean)
            CU(cudaMemcpyAsync(lm_mean + (size_t) l * P * 2, h->lm_mean + row * P, sizeof(float2) * P,
                               cudaMemcpyDeviceToHost, h->stream));
        if (lm_cov)
            CU(cudaMemcpyAsync(lm_cov + (size_t) l * P * 4, h->lm_cov + row * P, sizeof(float4) * P,
                               cudaMemcpyDeviceToHost, h->stream));
    }
    if (Nf) *Nf = h->Nf;
    CU(cudaStreamSynchronize(h->stream));
    return SLAMGPU_OK;
}

extern "C" int slamgpu_stage_noise(slamgpu_t *h, const float *gauss, int32_t gauss_rows, const double *unif) {
    if (!h) return slam_fail(SLAMGPU_EINVAL, "null handle");
    CU(cudaSetDevice(h->cfg.device));
    if (gauss) {
        if (gauss_rows != 2 && gauss_rows != 3) return slam_fail(SLAMGPU_EINVAL, "gauss_rows must be 2 or 3");
        CU(cudaMemcpyAsync(h->gauss, gauss, sizeof(float) * gauss_rows * h->P, cudaMemcpyHostToDevice, h->stream));
        h->gauss_staged = true;
        h->gauss_rows_staged = gauss_rows;
    }
    if (unif) {
        CU(cudaMemcpyAsync(h->unif(), unif, sizeof(double) * h->P, cudaMemcpyHostToDevice, h->stream));
        h->unif_staged = true;
    }
    CU(cudaStreamSynchronize(h->stream));
    return SLAMGPU_OK;
}

// ---- predict ----------------------------------------------------------------------------------------------
extern "C" int slamgpu_predict(slamgpu_t *h, float V, float G, const float Q[4], float dt, float phi_true,
                               const float *gauss) {
    if (!h || !Q) return slam_fail(SLAMGPU_EINVAL, "null argument");
    CU(cudaSetDevice(h->cfg.device));
    const float *dg = nullptr;
    if (gauss) {
        CU(cudaMemcpyAsync(h->gauss, gauss, sizeof(float) * 2 * h->P, cudaMemcpyHostToDevice, h->stream));
        dg = h->gauss;
    } else if (h->gauss_staged && h->gauss_rows_staged == 2) {
        dg = h->gauss;
    }
    PredictArgs a;
    a.V = V; a.G = G; a.dt = dt; a.phi_true = phi_true;
    for (int j = 0; j < 4; j++) a.Q[j] = Q[j];
    a.wheel_base = h->cfg.wheel_base;
    a.sigma_phi = h->cfg.sigma_phi;
    a.method = h->cfg.method;
    a.use_heading = h->cfg.use_heading;
    h->est_valid = false;
    const bool rng_noise = !dg && h->rng_on && h->rng_predict_noise;  // throughput mode: control noise from the generator
    if (!dg && !rng_noise) {  // no per-particle noise: queue the tick, apply the run of ticks in one pass over the poses
        h->predict_queue.push_back(a);
        if (h->predict_queue.size() >= SLAM_PREDICT_QUEUE) RET(flush_predicts(h));
        return SLAMGPU_OK;
    }
    RET(flush_predicts(h));
    RET(order_pose(h, h->cur, -1));
    TimedSpan sp;
    RET(span_begin(h, SLAMGPU_T_PREDICT, &sp));
    k_predict<<<grid_for(h->P, 256), 256, 0, h->stream>>>(h->soa(h->cur), h->P, a,
                                                          noise_src(h, dg, rng_noise ? h->draw_control++ : 0));
    KCHECK();
    LAUNCHED(h, SLAMGPU_T_PREDICT);
    RET(span_end(h, &sp));
    return SLAMGPU_OK;
}

// ---- update, per-particle part -----------------------------------------------------------------------------
// One pinned buffer, one copy: [K] float2 z | [K] int row | [K] int row_dst (row_dst == row unless the update kernel
// re-columns the row on the fly, kernels.cuh K3 REMAP).
#define OBS_BYTES(cap) ((sizeof(float2) + 2 * sizeof(int)) * (size_t) (cap))
static int stage_obs_host(slamgpu *h, const float *z, const int *rows, const int *rows_dst, int K, ObsSet *o) {
    if (K > h->obs_cap) {
        int cap = std::max(K, 2 * h->obs_cap);
        CU(cudaStreamSynchronize(h->stream));
        cudaFree(h->d_z); h->d_z = nullptr;
        cudaFreeHost(h->h_obs); h->h_obs = nullptr;
        CU(cudaMalloc(&h->d_z, OBS_BYTES(cap)));
        CU(cudaHostAlloc(&h->h_obs, OBS_BYTES(cap) * SLAM_OBS_SLOTS, cudaHostAllocDefault));
        h->obs_cap = cap;
    }
    // pinned staging slots are rotated; a slot is rewritten only after the copy that last read it has run
    const int slot = h->obs_slot++ % SLAM_OBS_SLOTS;
    if (!h->obs_ev[slot]) CU(cudaEventCreateWithFlags(&h->obs_ev[slot], cudaEventDisableTiming));
    else CU(cudaEventSynchronize(h->obs_ev[slot]));
    float2 *hz = (float2 *) ((char *) h->h_obs + OBS_BYTES(h->obs_cap) * slot);
    int *hr = (int *) (hz + K);
    for (int i = 0; i < K; i++) {
        hz[i] = make_float2(z[2 * i], z[2 * i + 1]);
        hr[i] = rows[i];
        hr[K + i] = rows_dst ? rows_dst[i] : rows[i];
    }
    CU(cudaMemcpyAsync(h->d_z, hz, OBS_BYTES(K), cudaMemcpyHostToDevice, h->stream));
    CU(cudaEventRecord(h->obs_ev[slot], h->stream));
    o->z = h->d_z;
    o->row = (const int *) (h->d_z + K);
    o->row_dst = o->row + K;
    return SLAMGPU_OK;
}

// Fused re-columning (kernels.cuh K3 REMAP): when every stale observed row belongs to ONE earlier resample epoch (the
// steady state: the rows observed last step were brought up to date then) the FastSLAM2 update kernel itself reads
// them through the ancestor map and writes the updated landmarks into fresh physical rows.  Fills src / dst physical
// rows per observation; *anc == nullptr means "nothing to fuse" (rows are current, or the general path ran).
#define SLAM_REMAP_MIN_ROWS 3
static int plan_fused_remap(slamgpu *h, const int *idf, int K, std::vector<int> &src, std::vector<int> &dst,
                            const int **anc, std::vector<int> &vacated) {
    *anc = nullptr;
    // the epoch most stale observed rows share is fused; rows of any other epoch (newcomers to the observed set)
    // catch up first through the separate gather kernels
    std::vector<long long> eps;
    for (int i = 0; i < K; i++)
        if (h->row_epoch[idf[i]] < h->epoch) eps.push_back(h->row_epoch[idf[i]]);
    long long e = -1;
    int n_stale = 0;
    if (!eps.empty()) {
        std::sort(eps.begin(), eps.end());
        for (size_t i = 0; i < eps.size();) {
            size_t j = i;
            while (j < eps.size() && eps[j] == eps[i]) j++;
            if ((int) (j - i) >= n_stale) {  // ties: the newer epoch (shorter ancestor chain)
                n_stale = (int) (j - i);
                e = eps[i];
            }
            i = j;
        }
        std::vector<int> others;
        // Fusing pays when many observed rows are stale (right after a resample: all K of them).  For a newcomer or two
        // (a landmark coming into view on a step that follows no resample) the REMAP instantiation of the update kernel
        // would cost more than it saves: it carries ~25 more instructions per particle-landmark for ALL K rows (~30 us at
        // cfg 4), a separate re-column pass ~12 us per row.
        static const int min_rows = getenv("SLAMGPU_REMAP_MIN_ROWS") ? atoi(getenv("SLAMGPU_REMAP_MIN_ROWS")) : SLAM_REMAP_MIN_ROWS;
        const bool fuse = (int) h->free_rows.size() >= n_stale && n_stale >= min_rows;
        for (int i = 0; i < K; i++) {
            const long long re = h->row_epoch[idf[i]];
            if (re < h->epoch && (!fuse || re != e)) others.push_back(idf[i]);
        }
        if (!others.empty()) RET(materialise(h, others));
        if (!fuse) n_stale = 0;
    }
    for (int i = 0; i < K; i++) src[i] = dst[i] = h->phys_row(idf[i]);
    if (n_stale == 0) return SLAMGPU_OK;
    if (h->epoch - e >= h->H) return slam_fail(SLAMGPU_EINVAL, "internal: a landmark row fell out of the resample log");
    if (e == h->epoch - 1) {
        *anc = h->keeplog + (size_t) (h->epoch % h->H) * h->Pg + h->slot0;  // one epoch back: keep[] itself
    } else {
        TimedSpan sp;
        RET(span_begin(h, SLAMGPU_T_GATHER, &sp));
        RET(launch_chain(h, h->anc, h->epoch, e, true));
        RET(span_end(h, &sp));
        *anc = h->anc;
    }
    for (int i = 0; i < K; i++) {
        const int l = idf[i];
        if (h->row_epoch[l] >= h->epoch) continue;
        dst[i] = h->free_rows.back();
        h->free_rows.pop_back();
        vacated.push_back(src[i]);
        h->phys[l] = dst[i];
        h->row_epoch[l] = h->epoch;
    }
    h->stat_rows_moved += n_stale;
    h->stat_rows_fused += n_stale;
    if (getenv("SLAMGPU_DIAG")) {  // lineage statistics of the ancestor map about to be used (debug aid, synchronises)
        unsigned long long *d_cnt, h_cnt[2];
        CU(cudaMalloc(&d_cnt, 16));
        CU(cudaMemsetAsync(d_cnt, 0, 16, h->stream));
        k_diag_ancestors<<<grid_for(h->P, 256), 256, 0, h->stream>>>(*anc, h->P, (long long) h->slot0, d_cnt);
        CU(cudaMemcpyAsync(h_cnt, d_cnt, 16, cudaMemcpyDeviceToHost, h->stream));
        CU(cudaStreamSynchronize(h->stream));
        cudaFree(d_cnt);
        fprintf(stderr, "[slamgpu diag] rank %d epoch %lld lag %lld rows %d remote %.4f distinct %.4f\n", h->rank,
                h->epoch, h->epoch - e, n_stale, (double) h_cnt[0] / h->P, (double) h_cnt[1] / h->P);
    }
    return SLAMGPU_OK;
}

extern "C" int slamgpu_update_particles(slamgpu_t *h, const float *zf, const int32_t *idf, int32_t K, const float *zn,
                                        int32_t Kn, const float R[4], const float *gauss) {
    if (!h || !R) return slam_fail(SLAMGPU_EINVAL, "null argument");
    if (K < 0 || Kn < 0 || (K > 0 && (!zf || !idf)) || (Kn > 0 && !zn)) return slam_fail(SLAMGPU_EINVAL, "bad observation set");
    if (h->Nf + Kn > h->nf_cap) return slam_fail(SLAMGPU_ECAP, "landmark capacity (nf_cap) exceeded");
    if (K > 2048 || Kn > 2048) return slam_fail(SLAMGPU_EINVAL, "at most 2048 observations per update (shared-memory staging)");
    CU(cudaSetDevice(h->cfg.device));
    RET(flush_predicts(h));
    std::vector<int> rows(std::max(K, Kn));
    if (K > 0) {
        std::vector<int> seen(idf, idf + K);
        std::sort(seen.begin(), seen.end());
        for (int i = 0; i < K; i++) {
            if (idf[i] < 0 || idf[i] >= h->Nf) return slam_fail(SLAMGPU_EINVAL, "idf entry outside [0, Nf)");
            if (i > 0 && seen[i] == seen[i - 1])
                return slam_fail(SLAMGPU_EINVAL, "idf holds a landmark twice (known association never does)");
        }
    }
    const bool fs2 = h->cfg.method == SLAMGPU_FASTSLAM2;
    const float *dg = nullptr;
    bool gauss_in_kernel = false;
    h->est_valid = false;
    if (fs2 && (K > 0 || Kn > 0)) {
        dg = h->gauss;
        if (gauss) {
            CU(cudaMemcpyAsync(h->gauss, gauss, sizeof(float) * 3 * h->P, cudaMemcpyHostToDevice, h->stream));
        } else if (h->gauss_inflight) {  // slamgpu_update put the copy on the copy stream: it overlaps the predict pass
            // ... and, chunk by chunk, the fast FastSLAM2 kernel itself, which waits for a chunk's flag where it
            // first needs the noise; every other consumer waits for the whole buffer
            gauss_in_kernel = h->gauss_chunked && K > 0 && h->cfg.math_mode == SLAMGPU_MATH_FAST;
            if (!gauss_in_kernel) {
                CU(cudaStreamWaitEvent(h->stream, h->ev_gauss, 0));
                h->gauss_inflight = false;
            }
        } else if (!(h->gauss_staged && h->gauss_rows_staged == 3)) {
            if (!h->rng_on)
                return slam_fail(SLAMGPU_EINVAL,
                                 "FastSLAM2 update needs gauss[3][P] (or slamgpu_stage_noise / slamgpu_seed first)");
            dg = nullptr;  // throughput mode
        }
    }
    // throughput mode (dg == nullptr): the pose sample's normals come from the device generator, one draw number per update
    const NoiseSrc noise = noise_src(h, dg, (fs2 && !dg && (K > 0 || Kn > 0)) ? h->draw_pose++ : 0);
    const float4 Rm = make_float4(R[0], R[1], R[2], R[3]);
    const size_t P = h->P;
    if (K > 0) {
        const bool fast = h->cfg.math_mode == SLAMGPU_MATH_FAST;
        std::vector<int> dst(K), vacated;
        const int *anc = nullptr;
        if (fast) {
            for (int i = 0; i < K; i++) rows[i] = dst[i] = h->phys_row(idf[i]);
            RET(plan_fused_remap(h, idf, K, rows, dst, &anc, vacated));
        } else {
            RET(materialise(h, std::vector<int>(idf, idf + K)));  // observed rows catch up with the resamples since
            for (int i = 0; i < K; i++) rows[i] = dst[i] = h->phys_row(idf[i]);
        }
        ObsSet o;
        RET(stage_obs_host(h, zf, rows.data(), dst.data(), K, &o));
        const size_t smem = (sizeof(float2) + sizeof(int)) * K;
        TimedSpan sp;
        RET(span_begin(h, SLAMGPU_T_UPDATE, &sp));
        PeerRows pr;
        memset(&pr, 0, sizeof(pr));
        pr.world = 1;
        pr.mean[0] = h->lm_mean;
        pr.cov[0] = h->lm_cov;
        const long long Pgl = (long long) h->Pg;
        if (fast && anc && h->world > 1) {
            // stale rows are read from the rank that holds the ancestor: every rank's earlier writes must have
            // landed, and the fresh rows written here must no longer be read by anyone (they were vacated by an
            // earlier kernel of some rank): one stream-ordered barrier covers both
            PeerMaps pm;
            RET(peer_maps(h, &pm));
            for (int r = 0; r < h->world; r++) {
                pr.mean[r] = pm.mean[r];
                pr.cov[r] = pm.cov[r];
            }
            pr.world = h->world;
            pr.spread = spread_stride(grid_for(P, fs2 ? SLAM_UPDATE_THREADS : 256), fs2 ? SLAM_UPDATE_THREADS : 256);
        }
        if (h->world > 1) {
            // rows read at the ancestor's column on a peer (src != dst) / rows written (all): see order_rows
            std::vector<int> srcs;
            for (int i = 0; i < K; i++)
                if (rows[i] != dst[i]) srcs.push_back(rows[i]);
            RET(order_rows(h, srcs.data(), (int) srcs.size(), dst.data(), K));
            RET(order_pose(h, h->cur, -1));
        }
        if (fs2 && fast) {
            const unsigned grid = grid_for(P, SLAM_UPDATE_THREADS), grid_l = std::min<unsigned>(grid, 148u * 4u);
            const size_t sm_f = SLAM_FS2_FAST_SMEM(K), sm_l = SLAM_OBS_FS2_SMEM(K);
            // no ragged tail in the unrolled walk over the observations
            const bool kfull = K >= SLAM_RING_DEPTH && K % SLAM_RING_DEPTH == 0;
            CU(cudaMemsetAsync(h->fb_count, 0, sizeof(int), h->stream));
            GaussArrival ga = {nullptr, 0, nullptr};
            if (gauss_in_kernel) {
                ga.flags = h->gauss_flags;
                ga.shift = h->gauss_shift();
                ga.late = &h->d_sc->gauss_late;
            }
            if (anc) {
                bool all_stale = true;  // the step after a resample: every observed row is read at the ancestor's column
                for (int i = 0; i < K; i++) all_stale = all_stale && rows[i] != dst[i];
                auto *kern = all_stale ? (kfull ? k_fs2_update_fast<2, true> : k_fs2_update_fast<2, false>)
                                       : (kfull ? k_fs2_update_fast<1, true> : k_fs2_update_fast<1, false>);
                kern<<<grid, SLAM_UPDATE_THREADS, sm_f, h->stream>>>(h->soa(h->cur), h->ring(), P, o, K, Rm, noise, ga, anc, pr,
                                                                     Pgl, h->fb_list, h->fb_count);
                KCHECK();
                LAUNCHED(h, SLAMGPU_T_UPDATE);
                if (gauss_in_kernel) {  // the fallback-list kernel (and anything later) reads the whole buffer
                    CU(cudaStreamWaitEvent(h->stream, h->ev_gauss, 0));
                    h->gauss_inflight = false;
                }
                k_fs2_update_listed<true><<<grid_l, 128, sm_l, h->stream>>>(h->soa(h->cur), h->ring(), P, o, K, Rm, noise,
                                                                            anc, pr, Pgl, h->fb_list, h->fb_count);
            } else {
                auto *kern = kfull ? k_fs2_update_fast<0, true> : k_fs2_update_fast<0, false>;
                kern<<<grid, SLAM_UPDATE_THREADS, sm_f, h->stream>>>(h->soa(h->cur), h->ring(), P, o, K, Rm, noise, ga, nullptr,
                                                                     pr, Pgl, h->fb_list, h->fb_count);
                KCHECK();
                LAUNCHED(h, SLAMGPU_T_UPDATE);
                if (gauss_in_kernel) {
                    CU(cudaStreamWaitEvent(h->stream, h->ev_gauss, 0));
                    h->gauss_inflight = false;
                }
                k_fs2_update_listed<false><<<grid_l, 128, sm_l, h->stream>>>(h->soa(h->cur), h->ring(), P, o, K, Rm, noise,
                                                                             nullptr, pr, Pgl, h->fb_list, h->fb_count);
            }
        } else if (fs2)
            k_fs2_update<<<grid_for(P, 128), 128, smem, h->stream>>>(h->soa(h->cur), h->ring(), P, o, K, Rm, noise);
        else if (fast && anc)
            k_fs1_update_fast<true><<<grid_for(P, 256), 256, SLAM_OBS_FS2_SMEM(K), h->stream>>>(
                h->soa(h->cur), h->ring(), P, o, K, Rm, anc, pr, Pgl);
        else if (fast)
            k_fs1_update_fast<false><<<grid_for(P, 256), 256, SLAM_OBS_FS2_SMEM(K), h->stream>>>(
                h->soa(h->cur), h->ring(), P, o, K, Rm, nullptr, pr, Pgl);
        else
            k_fs1_update<<<grid_for(P, 256), 256, smem, h->stream>>>(h->soa(h->cur), h->ring(), P, o, K, Rm);
        if (fast)
            for (int v : vacated) h->free_rows.insert(h->free_rows.begin(), v);
        KCHECK();
        LAUNCHED(h, SLAMGPU_T_UPDATE);
        RET(span_end(h, &sp));
    }
    if (Kn > 0) {
        for (int i = 0; i < Kn; i++) {
            rows[i] = h->phys_row(h->Nf + i);
            h->row_epoch[h->Nf + i] = h->epoch;
        }
        RET(order_rows(h, nullptr, 0, rows.data(), Kn));
        RET(order_pose(h, h->cur, -1));
        ObsSet o;
        RET(stage_obs_host(h, zn, rows.data(), nullptr, Kn, &o));
        const size_t smem = (sizeof(float2) + sizeof(int)) * Kn;
        TimedSpan sp;
        RET(span_begin(h, SLAMGPU_T_ADDFEAT, &sp));
        k_add_feature<<<grid_for(P, 256), 256, smem, h->stream>>>(h->soa(h->cur), h->ring(), P, o, Kn, Rm,
                                                                   (fs2 && K == 0) ? 1 : 0, noise);
        KCHECK();
        LAUNCHED(h, SLAMGPU_T_ADDFEAT);
        RET(span_end(h, &sp));
        h->Nf += Kn;
    }
    return SLAMGPU_OK;
}

// ---- resample ---------------------------------------------------------------------------------------------
// Where the resample uniforms of this call come from: the staged / uploaded double buffer, or the device generator.
struct UnifChoice {
    bool rng;
    unsigned draw;
};

static int resample_indices(slamgpu *h, const UnifChoice &uc) {
    const size_t P = h->P, Pg = h->Pg;
    const float k = (float) (1.0 / (float) Pg);
    const float khalf = k / 2;
    Scalars *d = h->d_sc;
    CU(cudaMemsetAsync(&h->d_tail->anomaly, 0, sizeof(int), h->stream));
    if (h->cfg.scan_mode == SLAMGPU_SCAN_REFERENCE) {
        if (h->world > 1) return slam_fail(SLAMGPU_EINVAL, "SLAMGPU_SCAN_REFERENCE is a one-device order");
        if (uc.rng) {  // the one-thread reference order reads its uniforms from the buffer
            k_fill_unif<<<grid_for(P, 256), 256, 0, h->stream>>>(h->unif(), P, h->rng_seed, (unsigned long long) h->slot0,
                                                                 uc.draw);
            KCHECK();
            LAUNCHED(h, SLAMGPU_T_WEIGHTS);
        }
        k_reference_weights<<<1, 1, 0, h->stream>>>(h->w, P, h->di(), h->unif(), k, khalf, h->mloc, h->cdf, h->sel, d,
                                                    (float) h->cfg.n_effective, h->cfg.do_resample,
                                                    h->cfg.keep_nan_weights);
        KCHECK();
        LAUNCHED(h, SLAMGPU_T_WEIGHTS);
        k_keep_sequential<<<1, 1, 0, h->stream>>>(h->sel, h->cdf, h->keep, P);
        KCHECK();
        LAUNCHED(h, SLAMGPU_T_WEIGHTS);
        return SLAMGPU_OK;
    }
    WeightsArgs wa;
    wa.w = h->w;
    wa.n = P;
    wa.tot_w = h->tot_w;
    wa.tot_c = h->tot_c;
    wa.tot_s = h->tot_s;
    wa.lmax = h->lmax;
    wa.bm = h->bm;
    wa.cdf_loc = h->cdf + h->slot0;
    wa.mloc_loc = h->mloc + h->slot0;
    wa.b0 = (int) (h->slot0 / SLAM_SCAN_BLOCK);  // first global block held here (P is a multiple of 1024 when sharded)
    wa.nb = (int) h->nb;
    wa.nbg = (int) h->nbg;
    wa.sc = d;
    wa.neff_threshold = (float) h->cfg.n_effective;
    wa.do_resample = h->cfg.do_resample;
    wa.uniform_w = 1.0f / (float) Pg;
    wa.keep_nan = h->cfg.keep_nan_weights;
    const int arr1[1] = {0}, arr2[3] = {1, 2, 3};
    const bool halo = !uc.rng;  // the generator can produce any slot's uniform anywhere: nothing to send
    if (h->world > 1 && h->log_pending) {
        // The all-gather of the previous keep[] (NCCL, own stream) has had a whole update kernel to finish; it must not
        // still be running next to the cooperative launch below.  A cooperative grid starts only when every CTA fits,
        // an NCCL kernel holding SM resources can delay that — and if this rank's NCCL kernel waits for a peer whose
        // SMs are full of a weights kernel spinning in its exchange for THIS rank's weights kernel, nobody moves.
        CU(cudaStreamWaitEvent(h->stream, h->ev_log, 0));
        h->log_pending = false;
    }
    if (h->coop_grid > 0) {
        // ONE launch: both grid-wide barriers and (sharded) both exchanges happen inside the kernel
        XchgArgs x1, x2;
        memset(&x1, 0, sizeof(x1));
        memset(&x2, 0, sizeof(x2));
        x1.world = x2.world = 1;
        if (h->world > 1) {
            RET(comm_fill_xchg(h->comm, h, arr1, 1, halo, nullptr, 0, 0, &x1));
            RET(comm_fill_xchg(h->comm, h, arr2, 3, false, nullptr, 0, 0, &x2));
            h->xcount += 2;
        }
        GridBar gb = {h->gbar, h->gbar_base};
        void *args[] = {(void *) &wa, (void *) &gb, (void *) &x1, (void *) &x2};
        CU(cudaLaunchCooperativeKernel((const void *) k_weights_fused, dim3(h->coop_grid), dim3(SLAM_SCAN_BLOCK), args, 0,
                                       h->stream));
        h->gbar_base += SLAM_GRIDBAR_TICKS((unsigned) h->coop_grid);
        LAUNCHED(h, SLAMGPU_T_WEIGHTS);
    } else {
        k_weights_a<<<wa.nb, SLAM_SCAN_BLOCK, 0, h->stream>>>(wa);
        KCHECK();
        LAUNCHED(h, SLAMGPU_T_WEIGHTS);
        if (h->world > 1) RET(xchg(h, arr1, 1, halo));
        k_weights_c<<<wa.nb, SLAM_SCAN_BLOCK, 0, h->stream>>>(wa);
        KCHECK();
        LAUNCHED(h, SLAMGPU_T_WEIGHTS);
        if (h->world > 1) RET(xchg(h, arr2, 3, false));
        k_weights_d<<<2, SLAM_SCAN_BLOCK, 0, h->stream>>>(wa);
        KCHECK();
        LAUNCHED(h, SLAMGPU_T_WEIGHTS);
    }
    // select + search.  The peers' in-block CDFs are read in place through the peer mappings; they were complete
    // before the second exchange, and a rank overwrites them only after the next update's first exchange.
    PeerMaps pm;
    RET(peer_maps(h, &pm));
    SearchArgs a;
    a.di = h->di();
    a.unif = uc.rng ? nullptr : h->unif();
    a.seed = h->rng_seed;
    a.draw = uc.draw;
    for (int r = 0; r < SLAM_MAX_RANKS; r++) {
        a.cdf[r] = pm.cdf[r];
        a.mloc[r] = pm.mloc[r];
    }
    a.world = h->world;
    a.Tc = h->tot_c;
    a.BM = h->bm;
    a.sel = h->sel;
    a.keep = h->keep;
    a.anomaly = &h->d_tail->anomaly;
    a.force_anomaly = h->force_anomaly ? 1 : 0;
    a.n_local = P;
    a.n_global = Pg;
    a.slot_offset = h->slot0;
    a.nb_global = (int) h->nbg;
    a.k = k;
    a.khalf = khalf;
    k_search_keep<<<grid_for(P, 256), 256, 0, h->stream>>>(a);
    KCHECK();
    LAUNCHED(h, SLAMGPU_T_WEIGHTS);
    return SLAMGPU_OK;
}

// The pose half of "particles[i] = oldParticles[keep[i]]" (core.cpp:739-748) into the second pose buffer.  cond: the
// device-side decision (nullptr = unconditional).  The buffers are flipped by resample_commit once the host knows.
static int launch_gather_pose(slamgpu *h, const int *cond) {
    const size_t P = h->P, Pg = h->Pg;
    const float new_w = 1.0f / (float) Pg;
    PeerMaps pm;
    RET(peer_maps(h, &pm));
    // sharded: every rank's update kernels must be done before a peer reads its poses, and the buffer written here
    // was the source of the previous resample: both are normally ordered by the exchanges of resample_indices
    RET(order_pose(h, h->cur ^ 1, h->cur));
    // one device: the kernel also writes this epoch's entry of the keep log.  Slot (epoch + 1) % H holds an epoch no
    // row can still need (materialise keeps every row within H - 1 epochs), so writing it before the decision is known
    // is harmless when the decision turns out to be "no resample".
    int *slot = h->keeplog + (size_t) ((h->epoch + 1) % h->H) * Pg;
    TimedSpan sg;
    RET(span_begin(h, SLAMGPU_T_GATHER, &sg));
    k_gather_pose<<<grid_for(P, 256), 256, 0, h->stream>>>(pm, h->cur, h->pose[h->cur ^ 1], h->w, h->keep, P,
                                                           (long long) Pg, new_w, h->rank, &h->d_sc->remote_sources,
                                                           h->world > 1 ? nullptr : slot, cond);
    KCHECK();
    LAUNCHED(h, SLAMGPU_T_GATHER);
    RET(span_end(h, &sg));
    return SLAMGPU_OK;
}

// The speculative tail of a resample in one launch: conditional pose gather + pose estimate (k_gather_pose_estimate).
static int launch_gather_estimate(slamgpu *h, const int *cond) {
    const size_t P = h->P, Pg = h->Pg;
    const float new_w = 1.0f / (float) Pg;
    PeerMaps pm;
    RET(peer_maps(h, &pm));
    RET(order_pose(h, h->cur ^ 1, h->cur));  // see launch_gather_pose
    int *slot = h->keeplog + (size_t) ((h->epoch + 1) % h->H) * Pg;
    TimedSpan sg;
    RET(span_begin(h, SLAMGPU_T_GATHER, &sg));
    k_gather_pose_estimate<<<grid_for(P, 256), 256, 0, h->stream>>>(
        pm, h->cur, h->pose[h->cur ^ 1], h->w, h->keep, P, (long long) Pg, new_w, h->rank, &h->d_sc->remote_sources,
        h->world > 1 ? nullptr : slot, cond, (long long) h->slot0, h->d_est, h->est_ticket, &h->d_tail->est);
    KCHECK();
    LAUNCHED(h, SLAMGPU_T_GATHER);
    RET(span_end(h, &sg));
    return SLAMGPU_OK;
}

// K8 into the tail record.  pick: device flag selecting the freshly gathered pose buffer (nullptr: the current one).
static int launch_estimate(slamgpu *h, const int *pick) {
    TimedSpan sp;
    RET(span_begin(h, SLAMGPU_T_ESTIMATE, &sp));
    k_estimate<<<h->est_blocks, 256, 0, h->stream>>>(h->pose[h->cur], h->pose[h->cur ^ 1], pick, h->w, h->P,
                                                     (long long) h->slot0, h->d_est, h->est_ticket, &h->d_tail->est);
    KCHECK();
    LAUNCHED(h, SLAMGPU_T_ESTIMATE);
    RET(span_end(h, &sp));
    return SLAMGPU_OK;
}

// The tail record of every rank into pinned memory (sharded: one exchange pushes it into every rank's slab first —
// that exchange is also how the ranks agree on the anomaly flag).
static int fetch_tail(slamgpu *h) {
    static_assert(sizeof(TailBlob) == 4 * SLAM_XCHG_BLOB_WORDS, "tail record size");
    if (h->world > 1) {
        // two landing areas by parity: a peer can be at most one exchange ahead
        const int slot = (int) (h->tail_count++ & 1u);
        RET(xchg(h, nullptr, 0, false, h->d_tail, (int) sizeof(TailBlob), slot));
        const float *area = h->tot_slab + SLAM_BLOB_BASE(h->nbg) + (size_t) slot * SLAM_MAX_RANKS * SLAM_XCHG_BLOB_WORDS;
        CU(cudaMemcpyAsync(h->h_tail, area, sizeof(TailBlob) * h->world, cudaMemcpyDeviceToHost, h->stream));
    } else {
        CU(cudaMemcpyAsync(h->h_tail, h->d_tail, sizeof(TailBlob), cudaMemcpyDeviceToHost, h->stream));
    }
    CU(cudaMemcpyAsync(h->h_sc, h->d_sc, sizeof(Scalars), cudaMemcpyDeviceToHost, h->stream));
    return SLAMGPU_OK;
}

// Errors the device raised since the last report (sticky flags in Scalars): checked at every host synchronisation
// that follows an exchange or an in-kernel wait, cleared only here.
static int report_device_flags(slamgpu *h) {
    if (h->h_sc->comm_error) {
        cudaMemsetAsync(&h->d_sc->comm_error, 0, sizeof(int), h->stream);
        h->h_sc->comm_error = 0;
        return slam_fail(SLAMGPU_ECOMM, "a peer device never arrived at an exchange (sharded resample)");
    }
    if (h->h_sc->gauss_late) {
        cudaMemsetAsync(&h->d_sc->gauss_late, 0, sizeof(int), h->stream);
        h->h_sc->gauss_late = 0;
        return slam_fail(SLAMGPU_ECUDA, "the noise upload never arrived behind the update kernel");
    }
    return SLAMGPU_OK;
}

// Rank-order combine of the tail records: identical result on every rank.
static void combine_estimate(slamgpu *h) {
    double sx = h->h_tail[0].est.sx, sy = h->h_tail[0].est.sy, th = h->h_tail[0].est.theta;
    float best = h->h_tail[0].est.wmax;
    long long bi = h->h_tail[0].est.imax;
    for (int r = 1; r < h->world; r++) {
        const EstPartial &e = h->h_tail[r].est;
        sx += e.sx;
        sy += e.sy;
        if (e.wmax > best || (e.wmax == best && e.imax < bi)) {
            best = e.wmax;
            bi = e.imax;
            th = e.theta;
        }
    }
    h->est_x = sx / (double) h->Pg;
    h->est_y = sy / (double) h->Pg;
    h->est_t = th;
    h->est_valid = true;
}

// keep[] by the exact sequential merge (core.cpp:800-806) when some rank's search met a slot it could not prove.
// One device: one thread over the final CDF.  Sharded: the merge is inherently sequential, so the ranks take turns
// in rank order — rank r continues from the CDF position rank r-1 ended at, which travels in the third landing area.
static int keep_fallback(slamgpu *h) {
    const size_t P = h->P, Pg = h->Pg;
    PeerMaps pm;
    RET(peer_maps(h, &pm));
    // the final CDF of this device's slots into its mloc block (the search is done with the running maxima)
    k_cdf_final<<<grid_for(P, 256), 256, 0, h->stream>>>(h->cdf + h->slot0, h->tot_c, (long long) h->slot0, P,
                                                         h->mloc + h->slot0);
    KCHECK();
    LAUNCHED(h, SLAMGPU_T_WEIGHTS);
    if (h->world <= 1) {
        k_keep_sequential<<<1, 1, 0, h->stream>>>(h->sel, h->mloc, h->keep, P);
        KCHECK();
        LAUNCHED(h, SLAMGPU_T_WEIGHTS);
        return SLAMGPU_OK;
    }
    RET(xchg(h, nullptr, 0, false));  // every rank's final CDF is in place
    SeqArgs q;
    for (int r = 0; r < SLAM_MAX_RANKS; r++) q.cdf[r] = pm.mloc[r];
    q.sel = h->sel;
    q.keep = h->keep;
    q.n_local = P;
    q.n_global = Pg;
    const long long *area = (const long long *) (h->tot_slab + SLAM_BLOB_BASE(h->nbg) +
                                                 (size_t) 2 * SLAM_MAX_RANKS * SLAM_XCHG_BLOB_WORDS);
    for (int r = 0; r < h->world; r++) {
        if (r == h->rank) {
            q.start = r > 0 ? area + (size_t) (r - 1) * (SLAM_XCHG_BLOB_WORDS / 2) : nullptr;
            q.end = (long long *) h->d_chain;
            k_keep_sequential_part<<<1, 1, 0, h->stream>>>(q);
            KCHECK();
            LAUNCHED(h, SLAMGPU_T_WEIGHTS);
        }
        RET(xchg(h, nullptr, 0, false, h->d_chain, (int) sizeof(TailBlob), 2));
    }
    return SLAMGPU_OK;
}

// What follows a resample that did happen, once the host knows: the pose buffers flip, keep[] joins the log, and the
// background trickle re-columns the stalest landmark rows (the maps themselves are not copied, kernels.cuh K7).
static int resample_commit(slamgpu *h) {
    const size_t P = h->P, Pg = h->Pg;
    if (h->log_pending) {  // the previous epoch's log must be whole before ancestor chains step through it
        CU(cudaStreamWaitEvent(h->stream, h->ev_log, 0));
        h->log_pending = false;
    }
    h->epoch++;
    h->stat_resamples++;
    h->cur ^= 1;
    int *slot = h->keeplog + (size_t) (h->epoch % h->H) * Pg;
    if (h->world > 1) CU(cudaMemcpyAsync(slot + h->slot0, h->keep, sizeof(int) * P, cudaMemcpyDeviceToDevice, h->stream));
    // a stride of the log has just been completed: compose it into one coarse map for this device's slots (the
    // chain's first hop reads this device's part of the newest keep[], the others the whole earlier arrays)
    int *coarse_new = nullptr;
    if (h->coarse_stride && h->epoch % h->coarse_stride == 0 && h->epoch >= h->coarse_stride) {
        const size_t cs = (size_t) ((h->epoch / h->coarse_stride) % h->coarse_slots);
        coarse_new = h->coarse + cs * Pg;
        h->coarse_epoch[cs] = -1;  // (the hops of this very chain must all be single keeps)
        const int G = h->coarse_stride;
        h->coarse_stride = 0;
        int rc = launch_chain(h, coarse_new + h->slot0, h->epoch, h->epoch - G, true);
        h->coarse_stride = G;
        RET(rc);
        h->coarse_epoch[cs] = h->epoch;
    }
    if (h->world > 1) {
        // The peers' parts of this epoch's keep[] are only read by ancestor chains that START at a later epoch (a
        // chain's first step looks up its own slot, i.e. this device's part): the all-gather runs on its own stream
        // behind the coming predict / update and is waited for at the next resample.
        if (!h->log_stream) {
            CU(cudaStreamCreateWithFlags(&h->log_stream, cudaStreamNonBlocking));
            CU(cudaEventCreateWithFlags(&h->ev_keep, cudaEventDisableTiming));
            CU(cudaEventCreateWithFlags(&h->ev_log, cudaEventDisableTiming));
        }
        CU(cudaEventRecord(h->ev_keep, h->stream));
        CU(cudaStreamWaitEvent(h->log_stream, h->ev_keep, 0));
        RET(comm_allgather_i32(h->comm, slot, P, h->log_stream));
        if (coarse_new) RET(comm_allgather_i32(h->comm, coarse_new, P, h->log_stream));
        CU(cudaEventRecord(h->ev_log, h->log_stream));
        h->log_pending = true;
    }
    // Background trickle: every row has to be re-columned before its backlog outgrows the log (H epochs).  The rows
    // cost 48 B x P each, and bringing a row forward from epoch e costs one ancestor chain of (epoch - e) look-ups per
    // particle — per GROUP of rows that share e, not per row.  So the trickle works in batches: every B-th resample
    // it takes the stalest B x ceil(Nf / (H - 2 - B)) rows at once.  Rows that were re-columned by one batch share an
    // epoch and come up together again ~H resamples later: ONE long chain serves the whole batch (with B = 1 every
    // resample paid a chain of ~H steps for ~Nf / H rows).  No row waits longer than H - 2 epochs: all Nf rows pass
    // in (H - 2 - B) / B batches, B resamples apart.
    const int B = h->trickle_batch;
    std::vector<std::pair<long long, int>> order;
    for (int l = 0; l < h->Nf; l++)
        if (h->row_epoch[l] < h->epoch) order.push_back({h->row_epoch[l], l});
    std::sort(order.begin(), order.end());
    const bool batch_now = (h->epoch % B) == 0;
    const int per = batch_now ? (int) (((long long) h->Nf * B + (h->H - 3 - B)) / (h->H - 2 - B)) : 0;
    std::vector<int> pick;
    for (size_t i = 0; i < order.size(); i++)
        if ((int) i < per || h->epoch - order[i].first >= h->H - 1) pick.push_back(order[i].second);
    return materialise(h, pick);
}

extern "C" int slamgpu_resample(slamgpu_t *h, const double *unif, int32_t *keep_out, float *neff,
                                int32_t *did_resample) {
    if (!h) return slam_fail(SLAMGPU_EINVAL, "null handle");
    CU(cudaSetDevice(h->cfg.device));
    RET(flush_predicts(h));
    UnifChoice uc = {false, 0};
    if (unif) {
        CU(cudaMemcpyAsync(h->unif(), unif, sizeof(double) * h->P, cudaMemcpyHostToDevice, h->stream));
    } else if (h->unif_inflight) {  // copied by slamgpu_update on the copy stream while the update kernel ran
        CU(cudaStreamWaitEvent(h->stream, h->ev_unif, 0));
        h->unif_inflight = false;
        if (h->unif_is_ints) {  // raw rand() values: unifRand() = rand() / double(RAND_MAX), core.cpp:775-777
            k_ints_to_unif<<<grid_for(h->P, 256), 256, 0, h->stream>>>(h->unif_i, h->unif(), h->P);
            KCHECK();
            LAUNCHED(h, SLAMGPU_T_WEIGHTS);
            h->unif_is_ints = false;
        }
    } else if (!h->unif_staged) {
        if (!h->rng_on)
            return slam_fail(SLAMGPU_EINVAL, "resample needs unif[P] (or slamgpu_stage_noise / slamgpu_seed first)");
        uc.rng = true;
        uc.draw = h->draw_resample++;
    }
    h->est_valid = false;
    const size_t P = h->P;
    if (h->dbg_capture) {
        if (!h->dbg_w) CU(cudaMalloc(&h->dbg_w, sizeof(float) * 2 * P));
        CU(cudaMemcpyAsync(h->dbg_w, h->w, sizeof(float) * P, cudaMemcpyDeviceToDevice, h->stream));
    }
    TimedSpan sp;
    RET(span_begin(h, SLAMGPU_T_WEIGHTS, &sp));
    RET(resample_indices(h, uc));
    RET(span_end(h, &sp));
    if (h->dbg_capture) CU(cudaMemcpyAsync(h->dbg_w + P, h->w, sizeof(float) * P, cudaMemcpyDeviceToDevice, h->stream));
    // The tail is enqueued BEFORE the host knows nEff: the pose gather and the pose estimate read the decision the
    // weights kernel left on the device (Scalars::did), so the one read-back of the step overlaps them instead of
    // sitting between two launches, and slamgpu_estimate right after an update costs neither a launch nor a second
    // synchronisation.
    const int *cond = h->cfg.do_resample ? &h->d_sc->did : nullptr;
    if (cond) RET(launch_gather_estimate(h, cond));
    else RET(launch_estimate(h, nullptr));
    RET(fetch_tail(h));
    CU(cudaStreamSynchronize(h->stream));
    h->gauss_chunked = false;
    RET(report_device_flags(h));
    const float ne = h->h_sc->neff;
    const bool did = h->cfg.do_resample && (ne < (float) h->cfg.n_effective);  // core.cpp:739
    if (did != (h->h_sc->did != 0)) return slam_fail(SLAMGPU_ECUDA, "internal: host and device disagree on the resample decision");
    bool anomaly = false;
    for (int r = 0; r < h->world; r++) anomaly = anomaly || h->h_tail[r].anomaly != 0;
    if (anomaly) {
        // knife-edge ordering the search could not prove (on some rank: the flags were exchanged with the tail records,
        // so every rank takes this branch together): exact sequential merge, then the gather again with the exact keep[]
        RET(keep_fallback(h));
        if (did) RET(launch_gather_pose(h, nullptr));
        h->stat_keep_fallbacks++;
    } else {
        combine_estimate(h);
    }
    if (keep_out) {
        CU(cudaMemcpyAsync(keep_out, h->keep, sizeof(int) * h->P, cudaMemcpyDeviceToHost, h->stream));
        CU(cudaStreamSynchronize(h->stream));
    }
    if (did) RET(resample_commit(h));
    if (neff) *neff = ne;
    if (did_resample) *did_resample = did ? 1 : 0;
    return SLAMGPU_OK;
}

static int update_impl(slamgpu_t *h, const float *zf, const int32_t *idf, int32_t K, const float *zn, int32_t Kn,
                       const float R[4], const float *gauss, const void *unif, bool unif_ints, int32_t *did_resample,
                       float *neff) {
    if (!h) return slam_fail(SLAMGPU_EINVAL, "null handle");
    // Host noise goes up on a second stream: the 12 B/particle of Gaussians overlap the predict pass, the resample
    // uniforms (8 B/particle as doubles, 4 B as raw rand() ints) overlap the whole update kernel (they are first read
    // by the search that follows it).
    const bool want_gauss = gauss && h->cfg.method == SLAMGPU_FASTSLAM2 && (K > 0 || Kn > 0);
    if (want_gauss || unif) {
        CU(cudaSetDevice(h->cfg.device));
        if (!h->copy_stream) {
            CU(cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
            CU(cudaEventCreateWithFlags(&h->ev_in, cudaEventDisableTiming));
            CU(cudaEventCreateWithFlags(&h->ev_gauss, cudaEventDisableTiming));
            CU(cudaEventCreateWithFlags(&h->ev_unif, cudaEventDisableTiming));
        }
        const bool chunked = want_gauss && K > 0 && h->cfg.math_mode == SLAMGPU_MATH_FAST && h->P >= (1u << 16);
        if (chunked) CU(cudaMemsetAsync(h->gauss_flags, 0, sizeof(unsigned) * SLAM_GAUSS_CHUNKS, h->stream));
        CU(cudaEventRecord(h->ev_in, h->stream));  // earlier readers of the two buffers are behind this point
        CU(cudaStreamWaitEvent(h->copy_stream, h->ev_in, 0));
        if (chunked) RET(flush_predicts(h));  // the queued predict pass runs while the chunks are being enqueued
        if (want_gauss) {
            if (chunked) {
                // column chunks in particle order, each followed by its arrival flag: the update kernel's CTAs run in
                // the same order and wait per chunk (kernels.cuh gauss_arrived_load), so only the first chunk's
                // transfer is exposed instead of all 12 B x P.  The flag is a 4-byte copy from a pinned constant: the
                // copy engine writes it, in stream order behind the chunk — no SM slot is needed while every SM is full
                // of waiting update CTAs (a cudaMemsetAsync is a kernel and would have to squeeze in between them).
                const size_t C = (size_t) 1 << h->gauss_shift();
                for (size_t c0 = 0, c = 0; c0 < h->P; c0 += C, c++) {
                    const size_t n = std::min(C, h->P - c0);
                    CU(cudaMemcpy2DAsync(h->gauss + c0, sizeof(float) * h->P, gauss + c0, sizeof(float) * h->P,
                                         sizeof(float) * n, 3, cudaMemcpyHostToDevice, h->copy_stream));
                    CU(cudaMemcpyAsync(h->gauss_flags + c, h->h_ones, sizeof(unsigned), cudaMemcpyHostToDevice,
                                       h->copy_stream));
                }
            } else {
                CU(cudaMemcpyAsync(h->gauss, gauss, sizeof(float) * 3 * h->P, cudaMemcpyHostToDevice, h->copy_stream));
            }
            CU(cudaEventRecord(h->ev_gauss, h->copy_stream));
            h->gauss_inflight = true;
            h->gauss_chunked = chunked;
            h->gauss_staged = false;
        }
        if (unif) {
            if (unif_ints) {
                if (!h->unif_i) CU(cudaMalloc(&h->unif_i, sizeof(int) * h->P));
                CU(cudaMemcpyAsync(h->unif_i, unif, sizeof(int) * h->P, cudaMemcpyHostToDevice, h->copy_stream));
            } else {
                CU(cudaMemcpyAsync(h->unif(), unif, sizeof(double) * h->P, cudaMemcpyHostToDevice, h->copy_stream));
            }
            CU(cudaEventRecord(h->ev_unif, h->copy_stream));
            h->unif_inflight = true;
            h->unif_is_ints = unif_ints;
            h->unif_staged = false;
        }
    }
    int rc = slamgpu_update_particles(h, zf, idf, K, zn, Kn, R, nullptr);
    if (rc == SLAMGPU_OK) rc = slamgpu_resample(h, nullptr, nullptr, neff, did_resample);
    if (rc != SLAMGPU_OK) {  // never leave a copy racing with the caller's buffers
        if (h->copy_stream) cudaStreamSynchronize(h->copy_stream);
        h->gauss_inflight = h->unif_inflight = false;
    }
    return rc;
}

extern "C" int slamgpu_update(slamgpu_t *h, const float *zf, const int32_t *idf, int32_t K, const float *zn,
                              int32_t Kn, const float R[4], const float *gauss, const double *unif,
                              int32_t *did_resample, float *neff) {
    return update_impl(h, zf, idf, K, zn, Kn, R, gauss, unif, false, did_resample, neff);
}
extern "C" int slamgpu_update_i32(slamgpu_t *h, const float *zf, const int32_t *idf, int32_t K, const float *zn,
                                  int32_t Kn, const float R[4], const float *gauss, const int32_t *rand_ints,
                                  int32_t *did_resample, float *neff) {
    return update_impl(h, zf, idf, K, zn, Kn, R, gauss, rand_ints, true, did_resample, neff);
}

extern "C" int slamgpu_seed(slamgpu_t *h, uint64_t seed, int32_t predict_noise) {
    if (!h) return slam_fail(SLAMGPU_EINVAL, "null handle");
    h->rng_on = true;
    h->rng_predict_noise = predict_noise != 0;
    h->rng_seed = seed;
    h->draw_pose = h->draw_resample = h->draw_control = 0;
    h->gauss_staged = h->unif_staged = false;  // staged host noise would take precedence
    return SLAMGPU_OK;
}

// ---- computeJacobians ---------------------------------------------------------------------------------------
extern "C" int slamgpu_compute_jacobians(slamgpu_t *h, const float *in, int32_t n, float *out) {
    if (!in || !out || n < 0) return slam_fail(SLAMGPU_EINVAL, "bad argument");
    if (n == 0) return SLAMGPU_OK;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return slam_fail(SLAMGPU_ENODEVICE, "no CUDA device: slamgpu has no CPU path");
    CU(cudaSetDevice(h ? h->cfg.device : 0));
    cudaStream_t st = h ? h->stream : (cudaStream_t) 0;
    float *din = nullptr, *dout = nullptr;
    const size_t nin = 7 + 6 * (size_t) n, nout = 16 * (size_t) n;
    if (h) {
        if (h->rec_cap < (size_t) n) {
            CU(cudaStreamSynchronize(st));
            if (h->d_rec_in) cudaFree(h->d_rec_in);
            if (h->d_rec_out) cudaFree(h->d_rec_out);
            h->d_rec_in = h->d_rec_out = nullptr;
            h->rec_cap = 0;
            const size_t cap = std::max((size_t) n, (size_t) 4096);
            CU(cudaMalloc(&h->d_rec_in, sizeof(float) * (7 + 6 * cap)));
            CU(cudaMalloc(&h->d_rec_out, sizeof(float) * 16 * cap));
            h->rec_cap = cap;
        }
        din = h->d_rec_in;
        dout = h->d_rec_out;
    } else {
        CU(cudaMalloc(&din, sizeof(float) * nin));
        CU(cudaMalloc(&dout, sizeof(float) * nout));
    }
    int rc = SLAMGPU_OK;
    do {
        if (cudaMemcpyAsync(din, in, sizeof(float) * nin, cudaMemcpyHostToDevice, st) != cudaSuccess) { rc = SLAMGPU_ECUDA; break; }
        k_jacobians_record<<<grid_for(n, 128), 128, 0, st>>>(din, n, (float4 *) dout);
        if (h) h->launches++;
        if (cudaMemcpyAsync(out, dout, sizeof(float) * nout, cudaMemcpyDeviceToHost, st) != cudaSuccess) { rc = SLAMGPU_ECUDA; break; }
        if (cudaStreamSynchronize(st) != cudaSuccess) { rc = SLAMGPU_ECUDA; break; }
    } while (0);
    if (!h) {
        cudaFree(din);
        cudaFree(dout);
    }
    if (rc != SLAMGPU_OK) return slam_fail(rc, std::string("compute_jacobians: ") + cudaGetErrorString(cudaGetLastError()));
    return SLAMGPU_OK;
}

extern "C" int slamgpu_compute_jacobians_batch(slamgpu_t *h, const int32_t *idf, int32_t K, const float R[4],
                                               float *out) {
    if (!h || !idf || !R || !out || K <= 0) return slam_fail(SLAMGPU_EINVAL, "bad argument");
    CU(cudaSetDevice(h->cfg.device));
    RET(flush_predicts(h));
    std::vector<int> rows(K);
    std::vector<float> z(2 * (size_t) K, 0.0f);
    for (int i = 0; i < K; i++) {
        if (idf[i] < 0 || idf[i] >= h->Nf) return slam_fail(SLAMGPU_EINVAL, "idf entry outside [0, Nf)");
    }
    RET(materialise(h, std::vector<int>(idf, idf + K)));
    for (int i = 0; i < K; i++) rows[i] = h->phys_row(idf[i]);
    ObsSet o;
    RET(stage_obs_host(h, z.data(), rows.data(), nullptr, K, &o));
    float4 *dout = nullptr;
    const size_t bytes = sizeof(float) * 16 * (size_t) K * h->P;
    CU(cudaMalloc(&dout, bytes));
    k_jacobians_batch<<<grid_for(h->P, 256), 256, (sizeof(float2) + sizeof(int)) * K, h->stream>>>(
        h->soa(h->cur), h->ring(), h->P, o, K, make_float4(R[0], R[1], R[2], R[3]), dout);
    h->launches++;
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(out, dout, bytes, cudaMemcpyDeviceToHost, h->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
    cudaFree(dout);
    if (e != cudaSuccess) return slam_fail(SLAMGPU_ECUDA, std::string("compute_jacobians_batch: ") + cudaGetErrorString(e));
    return SLAMGPU_OK;
}

// ---- estimate / decimated download ------------------------------------------------------------------------------
extern "C" int slamgpu_estimate(slamgpu_t *h, double *x, double *y, double *t) {
    if (!h) return slam_fail(SLAMGPU_EINVAL, "null handle");
    CU(cudaSetDevice(h->cfg.device));
    RET(flush_predicts(h));
    if (!h->est_valid) {  // (right after an update the resample's tail has already produced it: no launch, no wait)
        RET(launch_estimate(h, nullptr));
        RET(fetch_tail(h));
        CU(cudaStreamSynchronize(h->stream));
        RET(report_device_flags(h));
        combine_estimate(h);
    }
    if (x) *x = h->est_x;
    if (y) *y = h->est_y;
    if (t) *t = h->est_t;
    return SLAMGPU_OK;
}

extern "C" int slamgpu_download_poses_at(slamgpu_t *h, const int64_t *idx, int64_t n, float *xv, float *Pv, float *w) {
    if (!h || !idx || n <= 0) return slam_fail(SLAMGPU_EINVAL, "bad argument");
    for (int64_t j = 0; j < n; j++)
        if (idx[j] < 0 || (size_t) idx[j] >= h->P) return slam_fail(SLAMGPU_EINVAL, "particle index outside [0, P)");
    CU(cudaSetDevice(h->cfg.device));
    RET(flush_predicts(h));
    const size_t N = (size_t) n;
    long long *d_idx = nullptr;
    float *d_out = nullptr;
    CU(cudaMalloc(&d_idx, sizeof(long long) * N));
    cudaError_t e = cudaMalloc(&d_out, sizeof(float) * 13 * N);
    std::vector<float> out(13 * N);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_idx, idx, sizeof(long long) * N, cudaMemcpyHostToDevice, h->stream);
    if (e == cudaSuccess) {
        k_pick_poses<<<grid_for(N, 256), 256, 0, h->stream>>>(h->soa(h->cur), h->P, d_idx, N, d_out);
        h->launches++;
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(out.data(), d_out, sizeof(float) * 13 * N, cudaMemcpyDeviceToHost, h->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
    cudaFree(d_idx);
    if (d_out) cudaFree(d_out);
    if (e != cudaSuccess) return slam_fail(SLAMGPU_ECUDA, std::string("download_poses_at: ") + cudaGetErrorString(e));
    if (xv) memcpy(xv, out.data(), sizeof(float) * 3 * N);
    if (Pv) memcpy(Pv, out.data() + 3 * N, sizeof(float) * 9 * N);
    if (w) memcpy(w, out.data() + 12 * N, sizeof(float) * N);
    return SLAMGPU_OK;
}

// ---- parity instrumentation -----------------------------------------------------------------------------------------
extern "C" int slamgpu_debug_capture(slamgpu_t *h, int32_t on) {
    if (!h) return slam_fail(SLAMGPU_EINVAL, "null handle");
    h->dbg_capture = on != 0;
    return SLAMGPU_OK;
}
extern "C" int slamgpu_debug_read(slamgpu_t *h, int32_t which, void *out) {
    if (!h || !out) return slam_fail(SLAMGPU_EINVAL, "null argument");
    CU(cudaSetDevice(h->cfg.device));
    const size_t P = h->P;
    const void *src = nullptr;
    size_t bytes = sizeof(float) * P;
    switch (which) {
        case SLAMGPU_DEBUG_W_BEFORE: src = h->dbg_w; break;
        case SLAMGPU_DEBUG_W_NORMALISED: src = h->dbg_w ? h->dbg_w + P : nullptr; break;
        case SLAMGPU_DEBUG_KEEP: src = h->keep; bytes = sizeof(int) * P; break;
        case SLAMGPU_DEBUG_SELECT: src = h->sel; break;
        default: return slam_fail(SLAMGPU_EINVAL, "unknown debug item");
    }
    if (!src) return slam_fail(SLAMGPU_EINVAL, "nothing captured (slamgpu_debug_capture, then a resample)");
    CU(cudaMemcpyAsync(out, src, bytes, cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    return SLAMGPU_OK;
}

extern "C" int slamgpu_download_slice(slamgpu_t *h, int64_t p0, int64_t n, float *xv, float *Pv, float *w,
                                      float *lm_mean, float *lm_cov) {
    if (!h || p0 < 0 || n <= 0 || (size_t) (p0 + n) > h->P) return slam_fail(SLAMGPU_EINVAL, "slice outside [0, P)");
    CU(cudaSetDevice(h->cfg.device));
    RET(flush_predicts(h));
    const size_t P = h->P, N = (size_t) n;
    if (lm_mean || lm_cov) RET(materialise_all(h));
    PoseSoA s = h->soa(h->cur);
    const cudaMemcpyKind D2H = cudaMemcpyDeviceToHost;
    if (xv) CU(cudaMemcpy2DAsync(xv, sizeof(float) * N, s.xv + p0, sizeof(float) * P, sizeof(float) * N, 3, D2H, h->stream));
    if (Pv) CU(cudaMemcpy2DAsync(Pv, sizeof(float) * N, s.Pv + p0, sizeof(float) * P, sizeof(float) * N, 9, D2H, h->stream));
    if (w) CU(cudaMemcpyAsync(w, s.w + p0, sizeof(float) * N, D2H, h->stream));
    for (int l = 0; l < h->Nf; l++) {
        const size_t row = (size_t) h->phys_row(l);
        if (lm_mean)
            CU(cudaMemcpyAsync(lm_mean + (size_t) l * N * 2, h->lm_mean + row * P + p0, sizeof(float2) * N, D2H, h->stream));
        if (lm_cov)
            CU(cudaMemcpyAsync(lm_cov + (size_t) l * N * 4, h->lm_cov + row * P + p0, sizeof(float4) * N, D2H, h->stream));
    }
    CU(cudaStreamSynchronize(h->stream));
    return SLAMGPU_OK;
}

extern "C" int slamgpu_download_decimated(slamgpu_t *h, int64_t stride, float *xv_out, float *lm_out) {
    if (!h || stride <= 0 || !xv_out) return slam_fail(SLAMGPU_EINVAL, "bad argument");
    CU(cudaSetDevice(h->cfg.device));
    RET(flush_predicts(h));
    const size_t n = (h->P + (size_t) stride - 1) / (size_t) stride;
    float *dx = nullptr;
    float2 *dl = nullptr;
    CU(cudaMalloc(&dx, sizeof(float) * 3 * n));
    if (lm_out && h->Nf > 0) {
        cudaError_t e = cudaMalloc(&dl, sizeof(float2) * (size_t) h->Nf * n);
        if (e != cudaSuccess) {
            cudaFree(dx);
            return slam_fail(SLAMGPU_ENOMEM, "decimated landmark buffer");
        }
    }
    if (dl) {
        RET(materialise_all(h));
        CU(cudaMemcpyAsync(h->d_phys, h->phys.data(), sizeof(int) * h->Nf, cudaMemcpyHostToDevice, h->stream));
    }
    k_decimate<<<grid_for(n, 256), 256, 0, h->stream>>>(h->soa(h->cur), h->ring(), h->P, (size_t) stride, n, h->Nf,
                                                        h->d_phys, dx, dl);
    h->launches++;
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(xv_out, dx, sizeof(float) * 3 * n, cudaMemcpyDeviceToHost, h->stream);
    if (e == cudaSuccess && dl)
        e = cudaMemcpyAsync(lm_out, dl, sizeof(float2) * (size_t) h->Nf * n, cudaMemcpyDeviceToHost, h->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
    cudaFree(dx);
    if (dl) cudaFree(dl);
    if (e != cudaSuccess) return slam_fail(SLAMGPU_ECUDA, std::string("download_decimated: ") + cudaGetErrorString(e));
    return SLAMGPU_OK;
}

// ---- timing ---------------------------------------------------------------------------------------------
extern "C" int slamgpu_timing_enable(slamgpu_t *h, int32_t on) {
    if (!h) return slam_fail(SLAMGPU_EINVAL, "null handle");
    h->timing = on != 0;
    return SLAMGPU_OK;
}
extern "C" int slamgpu_timing_read(slamgpu_t *h, int32_t which, double *total_ms, int64_t *events, int64_t *launches) {
    if (!h || which < 0 || which >= SLAMGPU_T_COUNT) return slam_fail(SLAMGPU_EINVAL, "bad argument");
    CU(cudaSetDevice(h->cfg.device));
    RET(flush_predicts(h));
    RET(timing_flush(h));
    if (total_ms) *total_ms = h->t_ms[which];
    if (events) *events = h->t_events[which];
    if (launches) *launches = h->t_launches[which];
    return SLAMGPU_OK;
}
extern "C" int slamgpu_timing_reset(slamgpu_t *h) {
    if (!h) return slam_fail(SLAMGPU_EINVAL, "null handle");
    CU(cudaSetDevice(h->cfg.device));
    RET(timing_flush(h));
    for (int i = 0; i < SLAMGPU_T_COUNT; i++) {
        h->t_ms[i] = 0;
        h->t_events[i] = 0;
        h->t_launches[i] = 0;
    }
    return SLAMGPU_OK;
}

// ---- sharding ---------------------------------------------------------------------------------------------
extern "C" int slamgpu_comm_unique_id(void *id128) {
    if (!id128) return slam_fail(SLAMGPU_EINVAL, "null argument");
    std::string err;
    if (comm_unique_id(id128, &err) != 0) return slam_fail(SLAMGPU_ECOMM, err);
    return SLAMGPU_OK;
}

extern "C" int slamgpu_comm_init(slamgpu_t *h, int32_t rank, int32_t world, const void *id128) {
    if (!h || !id128 || world < 1 || rank < 0 || rank >= world) return slam_fail(SLAMGPU_EINVAL, "bad argument");
    if (world > 1 && (h->P % SLAM_SCAN_BLOCK) != 0)
        return slam_fail(SLAMGPU_EINVAL, "sharded particle counts must be a multiple of 1024 per device");
    if (world > 1 && h->cfg.scan_mode != SLAMGPU_SCAN_BLOCKED)
        return slam_fail(SLAMGPU_EINVAL, "sharding needs SLAMGPU_SCAN_BLOCKED");
    // keep[], the keep log and the ancestor maps are int32 GLOBAL indices and the upper scan level handles at most
    // 1024 x 1024 block totals: the same bound slamgpu_create puts on one handle holds for the whole sharded set
    if ((long long) h->P * (long long) world > (1ll << 30))
        return slam_fail(SLAMGPU_EINVAL, "n_particles x world exceeds 2^30 (int32 global particle indices)");
    CU(cudaSetDevice(h->cfg.device));
    std::string err;
    SlamComm *c = comm_create(rank, world, id128, h->cfg.device, &err);
    if (!c) return slam_fail(SLAMGPU_ECOMM, err);
    h->comm = c;
    h->rank = rank;
    h->world = world;
    h->Pg = h->P * (size_t) world;
    h->slot0 = h->P * (size_t) rank;
    // w = 1/N over the GLOBAL count (ParticleSLAMWrapper.cpp:17-25)
    std::vector<float> w0(h->P, 1.0f / (float) h->Pg);
    CU(cudaMemcpyAsync(h->w, w0.data(), sizeof(float) * h->P, cudaMemcpyHostToDevice, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    return setup_global(h);
}
