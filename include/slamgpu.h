/* slamgpu.h — C ABI of the B200 FastSLAM backend (slam_b200/libslamgpu.so).
 *
 * This is the drop-in boundary.  It sits where the reference's FPGA AcceleratorHandler sits
 * (src/backend/AcceleratorHandler.h:10-33: map device, getMemoryPointer/setN/start/isDone) but one
 * level higher: instead of "one particle's Jacobians" the library owns the whole particle set on the
 * device and exposes the FastSLAM1 / FastSLAM2 method entry points
 *     predict(particles, xTrue, V, G, Q, dt)            algorithms/fastslam2.h:20-21, fastslam1.h:29-30
 *     update(particles, zf, zn, idf, z, table, R)       algorithms/fastslam2.h:23-24, fastslam1.h:32-33
 * with `particles` replaced by an opaque handle.  A literal computeJacobians drop-in with the FPGA
 * float-record layout (core.cpp:586-664) is kept as slamgpu_compute_jacobians.
 *
 * Conventions
 *   - every call returns 0 on success, a negative SLAMGPU_E* code otherwise; slamgpu_last_error()
 *     gives the text.  Nothing throws across the ABI and nothing calls exit().
 *   - the library owns all device memory; the caller owns every host buffer it passes; no host
 *     pointer is retained after a call returns (except slamgpu_stage_noise's copy target on device).
 *   - one caller thread per handle (the reference loop is single threaded).
 *   - there is NO CPU fallback: without a CUDA device slamgpu_create fails with SLAMGPU_ENODEVICE.
 *
 * Flat state layout (all fp32, P = particles on this handle, Nf = landmarks held per particle):
 *     xv[c*P+p]            c<3  pose (x, y, theta)                         Particle::_xv   Particle.h:46
 *     Pv[(r*3+c)*P+p]           pose covariance planes, row-major 3x3      Particle::_Pv   Particle.h:47
 *     w[p]                      weight                                     Particle::_w    Particle.h:45
 *     lm_mean[(l*P+p)*2+j]      landmark mean,  l<Nf                       _landmarkXs     Particle.h:48
 *     lm_cov[(l*P+p)*4+j]       landmark covariance, j = column-major linear index (P(0),P(1),P(2),P(3)
 *                               exactly like the accelerator record, core.cpp:613-615)    Particle.h:49
 */
#ifndef SLAMGPU_H
#define SLAMGPU_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SLAMGPU_OK 0
#define SLAMGPU_EINVAL (-1)    /* bad argument */
#define SLAMGPU_ENODEVICE (-2) /* no usable CUDA device: the library has no CPU path */
#define SLAMGPU_ECUDA (-3)     /* a CUDA call failed; see slamgpu_last_error() */
#define SLAMGPU_ENOMEM (-4)    /* device allocation failed */
#define SLAMGPU_ECAP (-5)      /* landmark capacity (nf_cap) exceeded */
#define SLAMGPU_ECOMM (-6)     /* multi-device exchange failed */

#define SLAMGPU_FASTSLAM1 1
#define SLAMGPU_FASTSLAM2 2

/* Order of the floating-point sums behind the resampling CDF.
 * REFERENCE: Eigen's SSE redux order for the weight sums and a sequential fp32 running sum for the CDF —
 *            bit-identical to core.cpp:718-729,780-824; one thread, meant for the reference's own sizes.
 * BLOCKED  : fixed-shape blocked scan (blocks of 1024, 32-lane Hillis-Steele ladders, recursive over
 *            block totals) — the order used at scale; identical bits for any device count. */
#define SLAMGPU_SCAN_REFERENCE 0
#define SLAMGPU_SCAN_BLOCKED 1

/* Arithmetic of the per-particle update kernels (same estimator, same quirks, different operation order):
 * FAST           : information-form proposal, shared Cholesky factors, rsqrt / rcp (slam_math.cuh) — the default.
 * REFERENCE_ORDER: the reference's own sequence of Eigen operations (llt().solve, PartialPivLU inverse, ...),
 *                  IEEE division and square root; ~3.5x more instructions; for parity investigations. */
#define SLAMGPU_MATH_FAST 0
#define SLAMGPU_MATH_REFERENCE_ORDER 1

typedef struct slamgpu slamgpu_t;

typedef struct slamgpu_config {
    int64_t n_particles; /* NPARTICLES held by THIS handle (the local shard when sharded); 1 .. 2^28 - 1 */
    int32_t nf_cap;      /* landmark rows reserved per particle */
    int32_t method;      /* SLAMGPU_FASTSLAM1 / SLAMGPU_FASTSLAM2  (SLAMBackendApplication.cpp:26-29) */
    float wheel_base;    /* WHEELBASE  (fastslam2wrapper.cpp:18-23) */
    float sigma_phi;     /* sigmaT */
    int32_t n_effective; /* NEFFECTIVE: resample when nEff < n_effective (core.cpp:739; int like the reference) */
    int32_t use_heading; /* SWITCH_HEADING_KNOWN */
    int32_t do_resample; /* SWITCH_RESAMPLE */
    int32_t scan_mode;   /* SLAMGPU_SCAN_* */
    int32_t device;      /* CUDA device ordinal */
    int32_t spare_rows;  /* free physical landmark rows for deferred gathers (0 = default: min(64, nf_cap)) */
    int32_t math_mode;   /* SLAMGPU_MATH_* */
    int32_t log_epochs;  /* resamples whose keep[] is retained for deferred map gathers (0 = default 256) */
    int32_t keep_nan_weights; /* 0 (default): a zero / non-finite weight sum (every likelihood underflowed) sets w = 1/N and
                               * skips that resample, counted by SLAMGPU_STAT_WEIGHT_RESETS; 1: the reference's behaviour
                               * (core.cpp:727-729 divides by the sum: NaN weights, nEff NaN, no resample ever again) */
    int32_t reserved[1];
} slamgpu_config;

const char *slamgpu_last_error(void);
/* Library / build identification, e.g. "slamgpu 0.2 sm_100a". */
const char *slamgpu_version(void);

/* Replaces AcceleratorHandler::AcceleratorHandler / ~AcceleratorHandler (AcceleratorHandler.cpp:13-47):
 * particles start as in ParticleSLAMWrapper::initializeParticles (w = 1/N, xv = 0, Pv = 0, no landmarks). */
int slamgpu_create(const slamgpu_config *cfg, slamgpu_t **out);
void slamgpu_destroy(slamgpu_t *h);

/* The library launches everything on one stream.  Passing a caller-owned cudaStream_t (as void*) makes
 * caller-side CUDA events see the work; NULL restores the internal stream. */
int slamgpu_set_stream(slamgpu_t *h, void *cuda_stream);
void *slamgpu_get_stream(slamgpu_t *h);
int slamgpu_synchronize(slamgpu_t *h);

/* Pinned host memory for callers that want true async copies (cudaHostAlloc / cudaFreeHost). */
void *slamgpu_host_alloc(uint64_t bytes);
void slamgpu_host_free(void *p);

/* Whole-state transfer in the flat layout above (vector<Particle> <-> device).  Any pointer may be NULL
 * to skip that field.  Nf rows of lm_* are moved and Nf becomes the landmark count; upload with Nf < 0 leaves the
 * landmark count as it is (e.g. to replace only the weights). */
int slamgpu_upload_particles(slamgpu_t *h, const float *xv, const float *Pv, const float *w, const float *lm_mean,
                             const float *lm_cov, int32_t Nf);
int slamgpu_download_particles(slamgpu_t *h, float *xv, float *Pv, float *w, float *lm_mean, float *lm_cov,
                               int32_t *Nf);
/* Synthetic workloads (the -SYNTH_* configurations): a tile of Pt particles in the flat layout (with P := Pt)
 * is replicated on the device over all P particles, particle p taking tile particle p % Pt. */
int slamgpu_upload_particles_tiled(slamgpu_t *h, int64_t Pt, const float *xv, const float *Pv, const float *w,
                                   const float *lm_mean, const float *lm_cov, int32_t Nf);
int slamgpu_landmark_count(slamgpu_t *h);

/* FastSLAM{1,2}::predict (fastslam2.cpp:51-60,70-125; fastslam1.cpp:37-86): every particle.
 * Q row-major 2x2.  phi_true = xTrue(2) (used when use_heading).  gauss: NULL (no predict noise) or host
 * [2][P] standard normals, row 0 for V, row 1 for G, drawn like nRandMat::randn(2,1) per particle
 * (core.cpp:383-419).  FastSLAM1 passes gauss (addPredictNoise is forced on, fastslam1wrapper.cpp:20). */
int slamgpu_predict(slamgpu_t *h, float V, float G, const float Q[4], float dt, float phi_true,
                    const float *gauss);

/* FastSLAM{1,2}::update (fastslam2.cpp:21-48, fastslam1.cpp:18-35) = per-particle part + resampleParticles.
 *   zf [K][2] range/bearing of known landmarks, idf [K] their (distinct) landmark rows, zn [Kn][2] new ones
 *   (appended at rows Nf..Nf+Kn-1, core.cpp:479-509), R row-major 2x2.
 *   gauss: host [3][P] standard normals for the FastSLAM2 pose sample (multivariateGauss, core.cpp:452-458);
 *          ignored for FastSLAM1.  unif: host [P] doubles = rand()/double(RAND_MAX) (core.cpp:775-777),
 *          consumed on every update like the reference.  Both may be NULL after slamgpu_stage_noise().
 *   did_resample / neff (nullable) report the branch taken at core.cpp:739 and nEff. */
int slamgpu_update(slamgpu_t *h, const float *zf, const int32_t *idf, int32_t K, const float *zn, int32_t Kn,
                   const float R[4], const float *gauss, const double *unif, int32_t *did_resample, float *neff);

/* slamgpu_update with the resample uniforms as the raw rand() values (4 bytes per particle instead of 8): the device
 * forms unifRand() = rand() / double(RAND_MAX) itself, one IEEE division as core.cpp:776 — the same bits. */
int slamgpu_update_i32(slamgpu_t *h, const float *zf, const int32_t *idf, int32_t K, const float *zn, int32_t Kn,
                       const float R[4], const float *gauss, const int32_t *rand_ints, int32_t *did_resample,
                       float *neff);

/* Throughput mode.  The reference draws its noise INSIDE predict / update from rand() (fastslam2.cpp:353 ->
 * core.cpp:452-458, 383-419; core.cpp:767, 775-777).  After slamgpu_seed, calls made with gauss == NULL / unif == NULL
 * (and nothing staged) do the same on the device: a counter-based Philox4x32-10 keyed by `seed`, counter = (global
 * particle index, draw number, stream), fed through the reference's own rand()/RAND_MAX and Box-Muller arithmetic
 * (slam_b200/csrc/slam_rng.cuh; oracle/philox.py is its numpy twin).  The numbers do not depend on the sharding.
 * predict_noise != 0 also draws the control noise of every predict tick (SWITCH_PREDICT_NOISE / FastSLAM1).
 * Host-injected noise (parity mode) keeps working and takes precedence whenever it is passed or staged. */
int slamgpu_seed(slamgpu_t *h, uint64_t seed, int32_t predict_noise);

/* The two halves of slamgpu_update, for step-level parity checks. */
int slamgpu_update_particles(slamgpu_t *h, const float *zf, const int32_t *idf, int32_t K, const float *zn,
                             int32_t Kn, const float R[4], const float *gauss);
/* resampleParticles (core.cpp:718-749).  keep_out: nullable host [P] ints = stratifiedResample's keep[]
 * (global particle indices when sharded; -1 exactly where the reference leaves -1). */
int slamgpu_resample(slamgpu_t *h, const double *unif, int32_t *keep_out, float *neff, int32_t *did_resample);

/* Copies noise for the NEXT predict/update calls into device buffers so that those calls can be issued with
 * gauss == NULL / unif == NULL and no host->device traffic (device-resident benchmarking, device RNG feeds).
 * gauss_rows = 2 (predict) or 3 (FastSLAM2 update); either pointer may be NULL. */
int slamgpu_stage_noise(slamgpu_t *h, const float *gauss, int32_t gauss_rows, const double *unif);

/* computeJacobians, literal drop-in for the accelerator path (core.cpp:586-664), one particle:
 *   in : xv[3], R(0..3) column-major linear, then n x { xf[2], Pf(0..3) column-major linear }   (7+6n floats)
 *   out: n x { zp[2], Hf[4] row-major, Hv[6] row-major, Sf[4] row-major }                        (16n floats)
 * h may be NULL (device 0, default stream). */
int slamgpu_compute_jacobians(slamgpu_t *h, const float *in, int32_t n, float *out);
/* The batched form over the resident particle set: out host [K][P][16] in the record order above. */
int slamgpu_compute_jacobians_batch(slamgpu_t *h, const int32_t *idf, int32_t K, const float R[4], float *out);

/* ParticleSLAMWrapper::computeEstimatedPosition (ParticleSLAMWrapper.cpp:56-77): mean x, mean y (double
 * accumulation) and the heading of the first maximum-weight particle. */
int slamgpu_estimate(slamgpu_t *h, double *x, double *y, double *t);
/* The same estimate WITHOUT a handle and without allocating anything, for callers that cannot call cudaMalloc (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid it).  It has no counterpart in the reference.
 * The kernel's scratch (per-CTA partials + ticket, 33 KB) is a
 * static __device__ array that every GPU loading the library holds until exit; it bounds the call at
 * SLAMGPU_NOALLOC_MAX_PARTICLES, beyond which the call fails with SLAMGPU_EINVAL.  Runs on the CURRENT device.
 *   d_xv  : DEVICE memory, [3][n_particles] floats: the rows x, y, theta
 *   d_w   : DEVICE memory, [n_particles] weights
 *   d_out : DEVICE memory, 4 doubles: mean x, mean y, heading of the first maximum-weight particle, that weight
 *   cuda_stream : a cudaStream_t (NULL = the default stream); the result is ready when the stream reaches that point
 * No cudaMalloc / cudaFree / allocating routine is called (one cudaGetSymbolAddress the first time a device is used,
 * one 4-byte cudaMemsetAsync and one kernel per call).  All calls on a device share the scratch: calls that overlap
 * on the GPU (different streams, no ordering between them) race; serialise them. */
#define SLAMGPU_NOALLOC_MAX_PARTICLES (1 << 22)
int slamgpu_estimate_noalloc(const float *d_xv, const float *d_w, int64_t n_particles, double *d_out, void *cuda_stream);
/* Every stride-th particle's pose (and optionally its first Nf landmark means) for the plot feed
 * (ParticleSLAMWrapper.cpp:34-54 sends all of them; that cannot scale to 1M particles).
 * xv_out [3][n], lm_out [Nf][n][2] (nullable), n = ceil(P/stride). */
int slamgpu_download_decimated(slamgpu_t *h, int64_t stride, float *xv_out, float *lm_out);

/* A contiguous range [p0, p0 + n) of this handle's particles in the flat layout with P := n (any pointer may be
 * NULL): what whole-state parity checks at 2^20 particles x 1024 landmarks use instead of a 25 GB download. */
int slamgpu_download_slice(slamgpu_t *h, int64_t p0, int64_t n, float *xv, float *Pv, float *w, float *lm_mean,
                           float *lm_cov);

/* Poses, pose covariances and weights of the listed particles (local indices), packed with P := n. */
int slamgpu_download_poses_at(slamgpu_t *h, const int64_t *idx, int64_t n, float *xv, float *Pv, float *w);

/* Parity instrumentation.  With capture on, every resample keeps a device copy of the weights as they entered it and
 * as they left the normalisation; slamgpu_debug_read copies an item of the LAST resample to the host ([P] each):
 * the weights before / after normalisation (float), keep[] (int32, global indices) and select[] (float).  The data
 * path is otherwise unchanged (two device-to-device copies of 4 B per particle). */
#define SLAMGPU_DEBUG_W_BEFORE 0
#define SLAMGPU_DEBUG_W_NORMALISED 1
#define SLAMGPU_DEBUG_KEEP 2
#define SLAMGPU_DEBUG_SELECT 3
int slamgpu_debug_capture(slamgpu_t *h, int32_t on);
int slamgpu_debug_read(slamgpu_t *h, int32_t which, void *out);

/* Device-time accounting per kernel family (CUDA events on the launching stream). */
#define SLAMGPU_T_PREDICT 0
#define SLAMGPU_T_UPDATE 1   /* the per-particle update kernel (FastSLAM1 or FastSLAM2) */
#define SLAMGPU_T_ADDFEAT 2
#define SLAMGPU_T_WEIGHTS 3  /* weight sums, normalise, CDF, select, search */
#define SLAMGPU_T_GATHER 4   /* resample gather of poses and landmark maps */
#define SLAMGPU_T_ESTIMATE 5
#define SLAMGPU_T_COUNT 6
int slamgpu_timing_enable(slamgpu_t *h, int32_t on);
/* Sums since the last reset; synchronises the stream.  launches = kernel launches of that family. */
int slamgpu_timing_read(slamgpu_t *h, int32_t which, double *total_ms, int64_t *events, int64_t *launches);
int slamgpu_timing_reset(slamgpu_t *h);
/* Counters since creation: landmark rows re-columned by deferred resampling, resamples applied, ancestor-chain
 * steps walked (see slam_b200/csrc/kernels.cuh, K7). */
#define SLAMGPU_STAT_ROWS_MOVED 0
#define SLAMGPU_STAT_RESAMPLES 1
#define SLAMGPU_STAT_CHAIN_STEPS 2
#define SLAMGPU_STAT_REMOTE_SOURCES 4 /* sharded: resampled slots whose particle came from another device */
#define SLAMGPU_STAT_ROWS_FUSED 3 /* of ROWS_MOVED: re-columned inside the FastSLAM2 update kernel itself (no extra pass) */
#define SLAMGPU_STAT_KEEP_FALLBACKS 5 /* resamples whose keep[] came from the exact sequential merge (core.cpp:800-806) */
#define SLAMGPU_STAT_WEIGHT_RESETS 6 /* resamples that met a zero / non-finite weight sum (see keep_nan_weights) */
int64_t slamgpu_stat(slamgpu_t *h, int32_t which);
/* Total kernel launches issued by this handle since creation. */
int64_t slamgpu_launch_count(slamgpu_t *h);

/* ---- sharding over several devices (one handle per device, one process per device) ------------------
 * Particles are block-partitioned: rank r holds global particles [r*P, (r+1)*P).  predict / update_particles
 * need no exchange.  Resampling exchanges (1) the per-1024-block weight / CDF totals and block maxima in rank order
 * (the blocked scan is then bit-identical to the one-device run), (2) searches the owning device's CDF in place and
 * (3) migrates resampled particles — all three through peer memory (CUDA IPC over NVLink; the devices must be
 * peer-accessible, i.e. one box).  NCCL carries the bootstrap and the log of keep[]; the caller distributes the
 * 128-byte unique id.  SLAMGPU_STAT_REMOTE_SOURCES counts the particles that changed device. */
int slamgpu_comm_unique_id(void *id128);
int slamgpu_comm_init(slamgpu_t *h, int32_t rank, int32_t world, const void *id128);


/* ---- EKF-SLAM entry point (the reference's third -method, algorithms/ekfslam.h:14-56) -----------------------
 * One joint state x = [pose; landmarks] (n = 3 + 2*Nf) with its dense n x n covariance, resident on the device.
 * slamgpu_ekf_sim is EKFSLAM::sim (ekfslam.cpp:18-43): predict, optional compass update (observeHeading), and when
 * `observe` the association (gated nearest neighbour, or the known table), batchUpdate and augment.  Matrices cross
 * the ABI row-major.  Like the reference's wrapper (ekfslamwrapper.cpp:80-83: zf, idf, zn and the association table
 * are passed BY VALUE) the table is read-only here; the association made by the call is reported back. */
typedef struct slamgpu_ekf slamgpu_ekf_t;
typedef struct slamgpu_ekf_config {
    int32_t nf_cap;            /* landmarks the state can grow to */
    int32_t m_cap;             /* most observations per call (0 = 64) */
    float wheel_base;          /* WHEELBASE */
    float sigma_phi;           /* sigmaT */
    float gate_reject;         /* GATE_REJECT  (ekfslam.cpp:172) */
    float gate_augment;        /* GATE_AUGMENT (ekfslam.cpp:184) */
    int32_t use_heading;       /* SWITCH_HEADING_KNOWN */
    int32_t batch_update;      /* SWITCH_BATCH_UPDATE */
    int32_t association_known; /* SWITCH_ASSOCIATION_KNOWN */
    int32_t device;
    int32_t reserved[4];
} slamgpu_ekf_config;
int slamgpu_ekf_create(const slamgpu_ekf_config *cfg, slamgpu_ekf_t **out); /* x = 0, P = zeros(3,3) */
void slamgpu_ekf_destroy(slamgpu_ekf_t *h);
/* z [m][2] range/bearing of the visible landmarks (landmarksRangeBearing), idz [m] their identifiers and table [ntab]
 * (only read when association_known).  Qe / Re / R row-major 2x2 (Re gates and augments, R updates: ekfslam.cpp:34-41).
 * Out (nullable): idf_out [<= m] map slots of the associated observations in call order, their count, and the
 * number of new landmarks appended. */
int slamgpu_ekf_sim(slamgpu_ekf_t *h, float V, float G, const float Qe[4], float dt, float phi, int32_t observe,
                    const float *z, const int32_t *idz, int32_t m, const int32_t *table, int32_t ntab,
                    const float Re[4], const float R[4], int32_t *idf_out, int32_t *n_zf, int32_t *n_zn);
int slamgpu_ekf_size(slamgpu_ekf_t *h); /* n = x.size() */
int slamgpu_ekf_set_state(slamgpu_ekf_t *h, int32_t n, const float *x, const float *P);
int slamgpu_ekf_get_state(slamgpu_ekf_t *h, float *x, float *P); /* x [n], P [n][n]; either may be NULL */
int slamgpu_ekf_get_pose(slamgpu_ekf_t *h, float xv[3], float Pvv[9]); /* per-tick plot feed, ekfslamwrapper.cpp:91-101 */
int64_t slamgpu_ekf_launch_count(slamgpu_ekf_t *h);

#ifdef __cplusplus
}
#endif
#endif /* SLAMGPU_H */
