// handle.h — private: the device-side state behind a slamgpu_t and the error helpers shared by capi.cu / comm.cu.
#pragma once

#include <cuda_runtime.h>

#include <string>
#include <vector>

#include "../../include/slamgpu.h"
#include "types.cuh"

int slam_fail(int code, const std::string &msg);  // records the text for slamgpu_last_error(), returns code

#define CU(call)                                                                                          \
    do {                                                                                                  \
        cudaError_t e_ = (call);                                                                          \
        if (e_ != cudaSuccess)                                                                            \
            return slam_fail(e_ == cudaErrorMemoryAllocation ? SLAMGPU_ENOMEM : SLAMGPU_ECUDA,            \
                             std::string(#call) + ": " + cudaGetErrorString(e_) + " (" __FILE__ ":" +     \
                                 std::to_string(__LINE__) + ")");                                         \
    } while (0)
#define RET(call)                        \
    do {                                 \
        int r_ = (call);                 \
        if (r_ != SLAMGPU_OK) return r_; \
    } while (0)

struct SlamComm;
using slam::EstPartial;
using slam::MapRing;
using slam::PeerMaps;
using slam::PoseSoA;

#define SLAM_OBS_SLOTS 4  // rotating pinned staging slots for the observation set

using slam::Scalars;   // device scalars of the resample + pinned mirror (types.cuh)
using slam::TailBlob;  // pose-estimate partial + anomaly flag a resample's tail leaves for the host

struct TimedSpan {
    cudaEvent_t a, b;
    int which;
};

struct slamgpu {
    slamgpu_config cfg;
    size_t P = 0;         // local particles
    size_t Pg = 0;        // global particles (== P unless sharded)
    size_t slot0 = 0;     // first global particle held here
    int rank = 0, world = 1;
    int nf_cap = 0, Nf = 0, spare = 0, rows_total = 0;
    // landmark row pool (see kernels.cuh K7): logical -> physical table, free physical rows, and the resample epoch
    // each logical row's columns belong to
    std::vector<int> phys, free_rows;
    // Sharded hazard tracking (capi.cu order_rows): the exchange count at which a physical row was last written /
    // last read by a kernel that peers' copies of the same kernel may touch remotely.  One exchange between two
    // kernels orders them across ranks (a rank passes an exchange only after every peer's stream has reached it).
    long long xcount = 0;
    std::vector<long long> row_written_x, row_read_x;
    long long pose_written_x[2] = {-1, -1}, pose_read_x[2] = {-1, -1};
    std::vector<long long> row_epoch;
    long long epoch = 0;         // number of resamples applied to the poses so far
    int H = 0;                   // epochs of keep[] kept in the log
    int trickle_batch = 1;       // resamples between two batches of the background trickle (capi.cu resample_commit)
    int *keeplog = nullptr;      // [H][Pg] global keep arrays, epoch e at slot e % H
    int *coarse = nullptr;       // [coarse_slots][Pg] compositions of coarse_stride consecutive keep arrays (k_chain)
    int coarse_stride = 0, coarse_slots = 0;
    std::vector<long long> coarse_epoch;  // the epoch each coarse slot ends at (-1: empty)
    int *anc = nullptr;          // [P] scratch: composed ancestor map
    int *fb_list = nullptr, *fb_count = nullptr;  // [P], [1]: particles the fast FastSLAM2 kernel hands to the reference-order one
    slam::RowPair *d_pairs = nullptr, *h_pairs = nullptr;  // row-move lists (device / pinned), nf_cap entries
    int *d_phys = nullptr;       // device copy of phys[] for the download helpers
    std::vector<slam::PredictArgs> predict_queue;  // noise-free predict ticks not yet applied (flush_predicts)
    int64_t stat_rows_moved = 0, stat_resamples = 0, stat_chain_steps = 0, stat_rows_fused = 0, stat_keep_fallbacks = 0;
    cudaStream_t stream = nullptr, own_stream = nullptr;
    float *pose[2] = {nullptr, nullptr};  // each: xv[3][P] then Pv[9][P]
    int cur = 0;
    float *w = nullptr;
    float2 *lm_mean = nullptr;
    float4 *lm_cov = nullptr;
    float *gauss = nullptr;      // [3][P] staged normals
    double *unif_base = nullptr; // [2 + P] (2-slot halo in front for the sharded search)
    float *di_base = nullptr;    // [2 + P]
    float *cdf = nullptr, *mloc = nullptr;  // [Pg]
    float *sel = nullptr;        // [P]
    int *keep = nullptr;         // [P]
    float *tot_slab = nullptr;   // [SLAM_SLAB_ARRAYS][nb_global]: the arrays below, one allocation (one peer mapping)
    float *tot_w = nullptr, *tot_c = nullptr, *tot_s = nullptr, *lmax = nullptr, *bm = nullptr;  // [nb_global] each
    unsigned *gbar = nullptr;    // grid-barrier counter of the cooperative weights kernel (only ever grows)
    unsigned gbar_base = 0;      // its value before the next launch
    int coop_grid = 0;           // CTAs of the cooperative launch (0: split launches, SLAMGPU_WEIGHTS_SPLIT / unsupported)
    bool force_anomaly = false;  // SLAMGPU_FORCE_ANOMALY: every search reports an unproven slot (exercises the fallback)
    unsigned long long *xflags = nullptr;  // [SLAM_MAX_RANKS] arrival counters written by the peers (comm.cu k_xchg)
    size_t nb = 0, nbg = 0;      // local / global 1024-blocks
    Scalars *d_sc = nullptr, *h_sc = nullptr;
    float2 *d_z = nullptr;       // staged observation set: [K] z | [K] row | [K] row_dst (capi.cu stage_obs_host)
    cudaEvent_t obs_ev[SLAM_OBS_SLOTS] = {nullptr, nullptr, nullptr, nullptr};
    int obs_slot = 0;
    float *d_rec_in = nullptr, *d_rec_out = nullptr;
    size_t rec_cap = 0;
    int obs_cap = 0;
    void *h_obs = nullptr;       // pinned staging for the observation set
    EstPartial *d_est = nullptr;  // per-CTA partials of k_estimate
    TailBlob *d_tail = nullptr, *h_tail = nullptr;  // device record / pinned [SLAM_MAX_RANKS] (one per rank when sharded)
    TailBlob *d_chain = nullptr;  // record of the sharded keep[] fallback: where this rank's sequential merge ended
    int est_blocks = 0;
    // the estimate a resample's tail has already produced (slamgpu_estimate returns it without a launch); dropped by
    // anything that moves the poses
    bool est_valid = false;
    double est_x = 0, est_y = 0, est_t = 0;
    // throughput mode: noise from the device generator (slam_rng.cuh) instead of host buffers
    bool rng_on = false, rng_predict_noise = false;
    unsigned long long rng_seed = 0;
    unsigned draw_pose = 0, draw_resample = 0, draw_control = 0;
    // parity instrumentation (slamgpu_debug_capture): weights before / after normalisation of the last resample
    bool dbg_capture = false;
    float *dbg_w = nullptr;       // [2][P]
    int *unif_i = nullptr;        // [P] raw rand() ints of slamgpu_update_i32 (converted on the device)
    bool unif_is_ints = false;
    unsigned *h_ones = nullptr;   // pinned constant: arrival flags are written by the copy engine, not by a kernel
    bool gauss_staged = false, unif_staged = false;
    // slamgpu_update's host noise travels on a second stream (capi.cu)
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t ev_in = nullptr, ev_gauss = nullptr, ev_unif = nullptr;
    // sharded: the newest keep[] is all-gathered into the log on its own stream, behind the next predict / update
    cudaStream_t log_stream = nullptr;
    cudaEvent_t ev_keep = nullptr, ev_log = nullptr;
    bool log_pending = false;
    bool gauss_inflight = false, unif_inflight = false;
    bool gauss_chunked = false;       // the in-flight Gaussians carry per-chunk arrival flags (types.cuh GaussArrival)
    unsigned *gauss_flags = nullptr;  // [SLAM_GAUSS_CHUNKS]
    unsigned *est_ticket = nullptr;   // k_estimate's last-CTA counter (self-resetting)
    unsigned tail_count = 0;          // sharded: parity selects one of two landing areas for the peers' tail records
    int gauss_rows_staged = 0;
    // timing
    bool timing = false;
    std::vector<TimedSpan> spans;
    std::vector<cudaEvent_t> free_events;
    double t_ms[SLAMGPU_T_COUNT] = {0};
    int64_t t_events[SLAMGPU_T_COUNT] = {0};
    int64_t t_launches[SLAMGPU_T_COUNT] = {0};
    int64_t launches = 0;
    SlamComm *comm = nullptr;

    PoseSoA soa(int which) const {
        PoseSoA s;
        s.xv = pose[which];
        s.Pv = pose[which] + 3 * P;
        s.w = w;
        return s;
    }
    MapRing ring() const {
        MapRing m;
        m.mean = lm_mean;
        m.cov = lm_cov;
        m.rows = rows_total;
        return m;
    }
    int phys_row(int l) const { return phys[l]; }
    double *unif() const { return unif_base + 2; }
    unsigned gauss_shift() const {  // log2 of the column chunk of a chunked noise upload: ceil(P / chunks), power of two
        unsigned sh = 5;
        while ((((size_t) 1 << sh) * SLAM_GAUSS_CHUNKS) < P) sh++;
        return sh;
    }
    float *di() const { return di_base + 2; }
};


// ---- comm.cu: NCCL / peer-memory exchange for the sharded resample -----------------------------------------
int comm_unique_id(void *id128, std::string *err);
SlamComm *comm_create(int rank, int world, const void *id128, int device, std::string *err);
void comm_destroy(SlamComm *c);
// Maps every peer's pose / landmark buffers (CUDA IPC, once) and returns the table the gather kernels index by rank.
int comm_peer_maps(SlamComm *c, slamgpu *h, PeerMaps *out);
// Small exchange over peer memory instead of NCCL (latency: one 1-CTA kernel): all-gathers this rank's chunk
// [b0, b0 + nb) of the listed block-total arrays (0 tot_w, 1 tot_c, 2 tot_s, 3 lmax), optionally sends the uniform
// halo to the next rank, optionally pushes a small record (blob) into one of the landing areas of every rank, and
// is a stream-ordered barrier over all ranks.  narr == 0 and no blob: barrier only.
// comm_fill_xchg prepares the arguments (and consumes one exchange number) without launching: the cooperative weights
// kernel runs the exchange itself (xchg.cuh); comm_exchange = fill + launch of the stand-alone kernel.
int comm_fill_xchg(SlamComm *c, slamgpu *h, const int *arrays, int narr, bool halo, const void *blob, int blob_bytes,
                   int blob_slot, slam::XchgArgs *out);
int comm_exchange(SlamComm *c, slamgpu *h, const int *arrays, int narr, bool halo, const void *blob = nullptr,
                  int blob_bytes = 0, int blob_slot = 0);
// In-place NCCL all-gather over a global int array (the keep log): this rank's count elements sit at buf + rank*count.
int comm_allgather_i32(SlamComm *c, int *buf, size_t count, cudaStream_t st);
