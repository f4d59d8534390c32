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
extern "C" const char *slamgpu_version(void) { return "slamgpu 0.1 sm_100a"; }

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
                   h->gauss_flags, h->est_ticket};
    for (void *p : dev)
        if (p) cudaFree(p);
    if (h->h_sc) cudaFreeHost(h->h_sc);
    if (h->h_obs) cudaFreeHost(h->h_obs);
    if (h->h_est) cudaFreeHost(h->h_est);
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
    CU(cudaMalloc(&h->tot_slab, sizeof(float) * (4 * h->nbg + 2 * SLAM_MAX_RANKS * SLAM_XCHG_BLOB_WORDS)));
    h->tot_w = h->tot_slab;
    h->tot_c = h->tot_slab + h->nbg;
    h->tot_s = h->tot_slab + 2 * h->nbg;
    h->bmax = h->tot_slab + 3 * h->nbg;
    CU(cudaMalloc(&h->xflags, sizeof(unsigned long long) * SLAM_MAX_RANKS));
    CU(cudaMemset(h->xflags, 0, sizeof(unsigned long long) * SLAM_MAX_RANKS));
    if (h->keeplog) {
        cudaFree(h->keeplog);
        h->keeplog = nullptr;
    }
    CU(cudaMalloc(&h->keeplog, sizeof(int) * (size_t) h->H * Pg));
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
    CU(cudaMalloc(&h->d_est, sizeof(EstPartial) * (h->est_blocks + 1)));
    CU(cudaMalloc(&h->est_ticket, sizeof(unsigned)));
    CU(cudaMemset(h->est_ticket, 0, sizeof(unsigned)));
    CU(cudaHostAlloc(&h->h_est, sizeof(EstPartial) * SLAM_MAX_RANKS, cudaHostAllocDefault));
    {   // the FastSLAM2 fast kernels stage up to 2048 observations next to their landmark ring: > 48 KB
        const int big = (int) SLAM_FS2_FAST_SMEM(2048);
        CU(cudaFuncSetAttribute(k_fs2_update_fast<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
        CU(cudaFuncSetAttribute(k_fs2_update_fast<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, big));
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
    if (cfg->n_particles <= 0 || cfg->n_particles > (1ll << 30)) return slam_fail(SLAMGPU_EINVAL, "n_particles out of range");
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
    h->H = cfg->log_epochs > 0 ? std::max(4, cfg->log_epochs) : 64;
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
extern "C" int slamgpu_synchronize(slamgpu_t *h) {
    if (!h) return slam_fail(SLAMGPU_EINVAL, "null handle");
    CU(cudaSetDevice(h->cfg.device));
    RET(flush_predicts(h));
    CU(cudaStreamSynchronize(h->stream));
    return SLAMGPU_OK;
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
        case SLAMGPU_STAT_REMOTE_SOURCES: {  // as of the last weights read-back (lags the newest resample by one)
            unsigned long long v;
            memcpy(&v, &h->h_sc->pad[1], sizeof(v));
            return (int64_t) v;
        }
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

// ---- deferred predict ticks ---------------------------------------------------------------------------------------
static int flush_predicts(slamgpu *h) {
    if (h->predict_queue.empty()) return SLAMGPU_OK;
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
            k_chain<<<grid_for(P, 256), 256, 0, h->stream>>>(h->anc, P, (long long) h->slot0, (long long) Pg, h->keeplog,
                                                            h->H, cur_e, e + 1, identity ? 1 : 0);
            KCHECK();
            LAUNCHED(h, SLAMGPU_T_GATHER);
            h->stat_chain_steps += cur_e - e;
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
    if (!dg) {  // no per-particle noise: queue the tick, apply the run of ticks in one pass over the poses
        h->predict_queue.push_back(a);
        if (h->predict_queue.size() >= SLAM_PREDICT_QUEUE) RET(flush_predicts(h));
        return SLAMGPU_OK;
    }
    RET(flush_predicts(h));
    RET(order_pose(h, h->cur, -1));
    TimedSpan sp;
    RET(span_begin(h, SLAMGPU_T_PREDICT, &sp));
    k_predict<<<grid_for(h->P, 256), 256, 0, h->stream>>>(h->soa(h->cur), h->P, a, dg);
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
        const bool fuse = (int) h->free_rows.size() >= n_stale;
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
        k_chain<<<grid_for(h->P, 256), 256, 0, h->stream>>>(h->anc, h->P, (long long) h->slot0, (long long) h->Pg,
                                                            h->keeplog, h->H, h->epoch, e + 1, 1);
        KCHECK();
        LAUNCHED(h, SLAMGPU_T_GATHER);
        RET(span_end(h, &sp));
        h->stat_chain_steps += h->epoch - e;
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
    if (fs2 && (K > 0 || Kn > 0)) {
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
            return slam_fail(SLAMGPU_EINVAL, "FastSLAM2 update needs gauss[3][P] (or slamgpu_stage_noise first)");
        }
        dg = h->gauss;
    }
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
            CU(cudaMemsetAsync(h->fb_count, 0, sizeof(int), h->stream));
            GaussArrival ga = {nullptr, 0, nullptr};
            if (gauss_in_kernel) {
                ga.flags = h->gauss_flags;
                ga.shift = h->gauss_shift();
                ga.late = &h->d_sc->gauss_late;
            }
            if (anc) {
                k_fs2_update_fast<true><<<grid, SLAM_UPDATE_THREADS, sm_f, h->stream>>>(
                    h->soa(h->cur), h->ring(), P, o, K, Rm, dg, ga, anc, pr, Pgl, h->fb_list, h->fb_count);
                KCHECK();
                LAUNCHED(h, SLAMGPU_T_UPDATE);
                if (gauss_in_kernel) {  // the fallback-list kernel (and anything later) reads the whole buffer
                    CU(cudaStreamWaitEvent(h->stream, h->ev_gauss, 0));
                    h->gauss_inflight = false;
                }
                k_fs2_update_listed<true><<<grid_l, 128, sm_l, h->stream>>>(h->soa(h->cur), h->ring(), P, o, K, Rm, dg, anc,
                                                                            pr, Pgl, h->fb_list, h->fb_count);
            } else {
                k_fs2_update_fast<false><<<grid, SLAM_UPDATE_THREADS, sm_f, h->stream>>>(
                    h->soa(h->cur), h->ring(), P, o, K, Rm, dg, ga, nullptr, pr, Pgl, h->fb_list, h->fb_count);
                KCHECK();
                LAUNCHED(h, SLAMGPU_T_UPDATE);
                if (gauss_in_kernel) {
                    CU(cudaStreamWaitEvent(h->stream, h->ev_gauss, 0));
                    h->gauss_inflight = false;
                }
                k_fs2_update_listed<false><<<grid_l, 128, sm_l, h->stream>>>(h->soa(h->cur), h->ring(), P, o, K, Rm, dg,
                                                                             nullptr, pr, Pgl, h->fb_list, h->fb_count);
            }
        } else if (fs2)
            k_fs2_update<<<grid_for(P, 128), 128, smem, h->stream>>>(h->soa(h->cur), h->ring(), P, o, K, Rm, dg);
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
                                                                   (fs2 && K == 0) ? 1 : 0, dg);
        KCHECK();
        LAUNCHED(h, SLAMGPU_T_ADDFEAT);
        RET(span_end(h, &sp));
        h->Nf += Kn;
    }
    return SLAMGPU_OK;
}

// ---- resample ---------------------------------------------------------------------------------------------
static int launch_upper(slamgpu *h, float *a0, float *s0, float *a1, float *s1, int n, int use_max,
                        float *recip1 = nullptr) {
    UpperArgs u;
    memset(&u, 0, sizeof(u));
    u.arr[0] = a0; u.sum_out[0] = s0;
    u.arr[1] = a1; u.sum_out[1] = s1;
    u.recip_out[1] = recip1;
    u.n = n;
    u.use_max = use_max;
    k_scan_upper<<<a1 ? 2 : 1, SLAM_SCAN_BLOCK, 0, h->stream>>>(u);
    KCHECK();
    LAUNCHED(h, SLAMGPU_T_WEIGHTS);
    return SLAMGPU_OK;
}

static int resample_indices(slamgpu *h) {
    const size_t P = h->P, Pg = h->Pg;
    const float k = (float) (1.0 / (float) Pg);
    const float khalf = k / 2;
    Scalars *d = h->d_sc;
    CU(cudaMemsetAsync(&d->anomaly, 0, 2 * sizeof(int), h->stream));  // anomaly and pad[0] = exchange error
    if (h->cfg.scan_mode == SLAMGPU_SCAN_REFERENCE) {
        if (h->world > 1) return slam_fail(SLAMGPU_EINVAL, "SLAMGPU_SCAN_REFERENCE is a one-device order");
        k_reference_weights<<<1, 1, 0, h->stream>>>(h->w, P, h->di(), h->unif(), k, khalf, h->mloc, h->cdf, h->sel,
                                                    &d->neff);
        KCHECK();
        LAUNCHED(h, SLAMGPU_T_WEIGHTS);
        k_keep_sequential<<<1, 1, 0, h->stream>>>(h->sel, h->cdf, h->keep, P);
        KCHECK();
        LAUNCHED(h, SLAMGPU_T_WEIGHTS);
        return SLAMGPU_OK;
    }
    const unsigned nb = (unsigned) h->nb;
    const size_t b0 = h->slot0 / SLAM_SCAN_BLOCK;  // first global block held here (P is a multiple of 1024 when sharded)
    // A: block totals of w
    k_block_totals<<<nb, SLAM_SCAN_BLOCK, 0, h->stream>>>(h->w, P, h->tot_w + b0);
    KCHECK();
    LAUNCHED(h, SLAMGPU_T_WEIGHTS);
    if (h->world > 1) {
        const int arr[1] = {0};
        RET(xchg(h, arr, 1, false));
    }
    // B: recursive levels -> ws
    RET(launch_upper(h, h->tot_w, &d->ws, nullptr, nullptr, (int) h->nbg, 0));
    // C: normalise, local cdf scan, totals
    k_normalise_scan<<<nb, SLAM_SCAN_BLOCK, 0, h->stream>>>(h->w, P, &d->ws, h->cdf + h->slot0, h->tot_c + b0,
                                                            h->tot_s + b0);
    KCHECK();
    LAUNCHED(h, SLAMGPU_T_WEIGHTS);
    if (h->world > 1) {
        const int arr[2] = {1, 2};
        RET(xchg(h, arr, 2, false));
    }
    // D: recursive levels of both -> csum, s2
    RET(launch_upper(h, h->tot_c, &d->csum, h->tot_s, &d->s2, (int) h->nbg, 0, &d->neff));  // nEff = 1 / sum(w^2)
    // E: final cdf + running maxima
    k_cdf_finalise<<<nb, SLAM_SCAN_BLOCK, 0, h->stream>>>(h->cdf + h->slot0, P, h->tot_c, b0, h->mloc + h->slot0,
                                                          h->bmax + b0);
    KCHECK();
    LAUNCHED(h, SLAMGPU_T_WEIGHTS);
    PeerMaps pm;
    RET(peer_maps(h, &pm));
    if (h->world > 1) {
        // the block maxima are exchanged; the CDF and its running maxima stay where they are and are read through the
        // peer mappings.  This exchange is also the stream-ordered barrier that makes every rank's k_cdf_finalise
        // visible before any rank searches (and the next update's first exchange keeps a rank from overwriting its
        // CDF while a peer still searches it).
        const int arr[1] = {3};
        RET(xchg(h, arr, 1, true));
    }
    RET(launch_upper(h, h->bmax, nullptr, nullptr, nullptr, (int) h->nbg, 1));
    // F: select + search
    SearchArgs a;
    a.di = h->di();
    a.unif = h->unif();
    for (int r = 0; r < SLAM_MAX_RANKS; r++) {
        a.cdf[r] = pm.cdf[r];
        a.mloc[r] = pm.mloc[r];
    }
    a.world = h->world;
    a.BM = h->bmax;
    a.sel = h->sel;
    a.keep = h->keep;
    a.anomaly = &d->anomaly;
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

// The "particles[i] = oldParticles[keep[i]]" of core.cpp:739-748: poses now, maps deferred (kernels.cuh K7).
static int resample_state(slamgpu *h) {
    const size_t P = h->P, Pg = h->Pg;
    const float new_w = 1.0f / (float) Pg;
    TimedSpan sg;
    RET(span_begin(h, SLAMGPU_T_GATHER, &sg));
    if (h->log_pending) {  // the previous epoch's log must be whole before ancestor chains step through it
        CU(cudaStreamWaitEvent(h->stream, h->ev_log, 0));
        h->log_pending = false;
    }
    h->epoch++;
    h->stat_resamples++;
    int *slot = h->keeplog + (size_t) (h->epoch % h->H) * Pg;
    if (h->world > 1) {
        CU(cudaMemcpyAsync(slot + h->slot0, h->keep, sizeof(int) * P, cudaMemcpyDeviceToDevice, h->stream));
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
        CU(cudaEventRecord(h->ev_log, h->log_stream));
        h->log_pending = true;
    }
    PeerMaps pm;
    RET(peer_maps(h, &pm));
    // sharded: every rank's update kernels must be done before a peer reads its poses, and the buffer written here
    // was the source of the previous resample: both are normally ordered by the three exchanges of resample_indices
    RET(order_pose(h, h->cur ^ 1, h->cur));
    k_gather_pose<<<grid_for(P, 256), 256, 0, h->stream>>>(pm, h->cur, h->pose[h->cur ^ 1], h->w, h->keep, P,
                                                           (long long) Pg, new_w, h->rank,
                                                           (unsigned long long *) &h->d_sc->pad[1],
                                                           h->world > 1 ? nullptr : slot);
    KCHECK();
    LAUNCHED(h, SLAMGPU_T_GATHER);
    h->cur ^= 1;
    RET(span_end(h, &sg));
    // background trickle: the stalest rows, enough per resample that no row's backlog reaches the log length
    const int per = (h->Nf + (h->H - 3)) / (h->H - 2);
    std::vector<std::pair<long long, int>> order;
    for (int l = 0; l < h->Nf; l++)
        if (h->row_epoch[l] < h->epoch) order.push_back({h->row_epoch[l], l});
    std::sort(order.begin(), order.end());
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
    if (unif) {
        CU(cudaMemcpyAsync(h->unif(), unif, sizeof(double) * h->P, cudaMemcpyHostToDevice, h->stream));
    } else if (h->unif_inflight) {  // copied by slamgpu_update on the copy stream while the update kernel ran
        CU(cudaStreamWaitEvent(h->stream, h->ev_unif, 0));
        h->unif_inflight = false;
    } else if (!h->unif_staged) {
        return slam_fail(SLAMGPU_EINVAL, "resample needs unif[P] (or slamgpu_stage_noise first)");
    }
    TimedSpan sp;
    RET(span_begin(h, SLAMGPU_T_WEIGHTS, &sp));
    RET(resample_indices(h));
    CU(cudaMemcpyAsync(h->h_sc, h->d_sc, sizeof(Scalars), cudaMemcpyDeviceToHost, h->stream));
    RET(span_end(h, &sp));
    CU(cudaStreamSynchronize(h->stream));
    if (h->h_sc->pad[0]) return slam_fail(SLAMGPU_ECOMM, "a peer device never arrived at an exchange (sharded resample)");
    h->gauss_chunked = false;
    if (h->h_sc->gauss_late) {
        cudaMemset(&h->d_sc->gauss_late, 0, sizeof(int));
        return slam_fail(SLAMGPU_ECUDA, "the noise upload never arrived behind the update kernel");
    }
    if (h->h_sc->anomaly) {
        // knife-edge ordering the search could not prove: run the exact sequential merge (core.cpp:800-806)
        if (h->world > 1) return slam_fail(SLAMGPU_ECOMM, "sequential keep[] fallback is not available when sharded");
        k_keep_sequential<<<1, 1, 0, h->stream>>>(h->sel, h->cdf, h->keep, h->P);
        KCHECK();
        LAUNCHED(h, SLAMGPU_T_WEIGHTS);
    }
    const float ne = h->h_sc->neff;
    const bool did = h->cfg.do_resample && (ne < (float) h->cfg.n_effective);  // core.cpp:739
    if (keep_out) {
        CU(cudaMemcpyAsync(keep_out, h->keep, sizeof(int) * h->P, cudaMemcpyDeviceToHost, h->stream));
        CU(cudaStreamSynchronize(h->stream));
    }
    if (did) {
        RET(resample_state(h));
    }
    if (neff) *neff = ne;
    if (did_resample) *did_resample = did ? 1 : 0;
    return SLAMGPU_OK;
}

extern "C" int slamgpu_update(slamgpu_t *h, const float *zf, const int32_t *idf, int32_t K, const float *zn,
                              int32_t Kn, const float R[4], const float *gauss, const double *unif,
                              int32_t *did_resample, float *neff) {
    if (!h) return slam_fail(SLAMGPU_EINVAL, "null handle");
    // Host noise goes up on a second stream: the 12 B/particle of Gaussians overlap the predict pass, the 8 B/particle
    // of resample uniforms overlap the whole update kernel (they are first read by the search that follows it).
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
                // transfer is exposed instead of all 12 B x P
                const size_t C = (size_t) 1 << h->gauss_shift();
                for (size_t c0 = 0, c = 0; c0 < h->P; c0 += C, c++) {
                    const size_t n = std::min(C, h->P - c0);
                    CU(cudaMemcpy2DAsync(h->gauss + c0, sizeof(float) * h->P, gauss + c0, sizeof(float) * h->P,
                                         sizeof(float) * n, 3, cudaMemcpyHostToDevice, h->copy_stream));
                    CU(cudaMemsetAsync(h->gauss_flags + c, 0xff, sizeof(unsigned), h->copy_stream));
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
            CU(cudaMemcpyAsync(h->unif(), unif, sizeof(double) * h->P, cudaMemcpyHostToDevice, h->copy_stream));
            CU(cudaEventRecord(h->ev_unif, h->copy_stream));
            h->unif_inflight = true;
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
    TimedSpan sp;
    RET(span_begin(h, SLAMGPU_T_ESTIMATE, &sp));
    PoseSoA s = h->soa(h->cur);
    k_estimate<<<h->est_blocks, 256, 0, h->stream>>>(s.xv, s.w, h->P, (long long) h->slot0, h->d_est, h->est_ticket,
                                                     h->d_est + h->est_blocks);
    KCHECK();
    LAUNCHED(h, SLAMGPU_T_ESTIMATE);
    static_assert(sizeof(EstPartial) == 4 * SLAM_XCHG_BLOB_WORDS, "estimate record size");
    if (h->world > 1) {
        // every rank's partial lands in every rank's slab (peer stores, one exchange), then one read-back: no NCCL
        // call, no second synchronisation.  Two landing areas by parity: a peer can be at most one exchange ahead.
        const int slot = (int) (h->est_count++ & 1u);
        RET(xchg(h, nullptr, 0, false, h->d_est + h->est_blocks, (int) sizeof(EstPartial), slot));
        const float *area = h->tot_slab + 4 * h->nbg + (size_t) slot * SLAM_MAX_RANKS * SLAM_XCHG_BLOB_WORDS;
        CU(cudaMemcpyAsync(h->h_est, area, sizeof(EstPartial) * h->world, cudaMemcpyDeviceToHost, h->stream));
    } else {
        CU(cudaMemcpyAsync(h->h_est, h->d_est + h->est_blocks, sizeof(EstPartial), cudaMemcpyDeviceToHost, h->stream));
    }
    RET(span_end(h, &sp));
    CU(cudaStreamSynchronize(h->stream));
    double sx = h->h_est[0].sx, sy = h->h_est[0].sy, th = h->h_est[0].theta;
    float best = h->h_est[0].wmax;
    long long bi = h->h_est[0].imax;
    for (int r = 1; r < h->world; r++) {  // rank order: identical result on every rank
        const EstPartial &e = h->h_est[r];
        sx += e.sx;
        sy += e.sy;
        if (e.wmax > best || (e.wmax == best && e.imax < bi)) {
            best = e.wmax;
            bi = e.imax;
            th = e.theta;
        }
    }
    if (x) *x = sx / (double) h->Pg;
    if (y) *y = sy / (double) h->Pg;
    if (t) *t = th;
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
