// types.cuh — plain structs shared by the kernels and the host side of the library.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace slam {

#define SLAM_SCAN_BLOCK 1024  // the blocked-scan shape shared with oracle/slam_oracle.c (SO_SCAN_BLOCK)

struct PoseSoA {  // xv[3][P], Pv[9][P], w[P]
    float *xv;
    float *Pv;
    float *w;
};

struct MapRing {   // landmark rows: a pool of (nf_cap + spare) physical rows; the host keeps the logical -> physical table
    float2 *mean;  // [rows][P]
    float4 *cov;   // [rows][P], column-major linear 2x2 (core.cpp:613-615)
    int rows;
};

#define SLAM_MAX_RANKS 8
struct PeerMaps {  // this device's and (when sharded) the peers' buffers, indexed by rank
    const float *pose[2][SLAM_MAX_RANKS];
    const float2 *mean[SLAM_MAX_RANKS];
    const float4 *cov[SLAM_MAX_RANKS];
    const float *cdf[SLAM_MAX_RANKS];   // each rank's [Pg]-indexed CDF / running-maximum buffers (its own block filled)
    const float *mloc[SLAM_MAX_RANKS];
    float *slab[SLAM_MAX_RANKS];               // block-total arrays [5][nbg]: tot_w | tot_c | tot_s | lmax | bm
    unsigned long long *flags[SLAM_MAX_RANKS]; // [SLAM_MAX_RANKS] arrival counters of the peer-memory exchange
    double *unif[SLAM_MAX_RANKS];              // unif_base (2-slot halo in front)
    int world;
    unsigned spread;  // see PeerRows
};

// One small exchange over peer memory (comm.cu k_xchg): every rank pushes its chunk of up to 3 block-total arrays into
// every peer's copy, optionally its last two resample uniforms into the next rank's halo, then signals and waits.
struct XchgArgs {
    float *slab[SLAM_MAX_RANKS];
    unsigned long long *flags[SLAM_MAX_RANKS];
    double *unif[SLAM_MAX_RANKS];
    int rank, world;
    unsigned long long epoch;
    int narr, arr_off[3];  // element offsets of the arrays inside the slab
    int b0, count;         // this rank's chunk
    int halo;
    unsigned long long P;
    int *error;            // set when a peer never arrived
    // optional small record (the pose-estimate partial): blob_words 32-bit words from blob_src go to word offset
    // blob_off + rank * blob_words of EVERY rank's slab (its own included), so that each rank ends up with all of them
    const unsigned *blob_src;
    int blob_words, blob_off;
};
#define SLAM_XCHG_BLOB_WORDS 12  // sizeof(TailBlob) / 4
#define SLAM_XCHG_BLOB_SLOTS 3   // two landing areas alternate between consecutive tails, the third carries the
                                 // sequential-merge chain of the keep[] fallback
#define SLAM_SLAB_ARRAYS 5       // tot_w | tot_c | tot_s | lmax (exchanged) | bm (local)
// word offset of the landing areas behind the arrays of a slab over nbg blocks (16-byte aligned: records hold doubles)
#define SLAM_BLOB_BASE(nbg) ((SLAM_SLAB_ARRAYS * (size_t) (nbg) + 3) & ~(size_t) 3)

struct PeerRows {  // landmark pools by rank, for kernels that read a particle's ancestor column wherever it lives
    const float2 *mean[SLAM_MAX_RANKS];
    const float4 *cov[SLAM_MAX_RANKS];
    int world;
    unsigned spread;  // 0, or the chunk stride of spread_block(): remote-heavy CTAs interleaved with local ones
};

// Host noise arriving behind a running kernel (slamgpu_update): the [3][P] Gaussians go up in SLAM_GAUSS_CHUNKS column
// chunks on the copy stream, each followed by a write of its flag; a warp of the update kernel waits for the flag of
// the chunk its particles sit in, right before it samples the pose.  flags == nullptr: the buffer is already whole.
#define SLAM_GAUSS_CHUNKS 8
struct GaussArrival {
    const unsigned *flags;  // [SLAM_GAUSS_CHUNKS], nonzero once the chunk has landed
    unsigned shift;         // chunk of particle p = p >> shift
    int *late;              // raised when a flag never came (bounded wait)
};

struct RowPair {  // one row move of a materialisation: physical source row -> physical destination row
    int src, dst;
};
#define SLAM_MOVES_PER_LAUNCH 64
struct RowMoves {  // passed to the kernel by value (512 B): no staging buffer to keep alive
    RowPair p[SLAM_MOVES_PER_LAUNCH];
};

struct PredictArgs {
    float V, G, Q[4], dt, phi_true, wheel_base, sigma_phi;
    int method, use_heading;
};

struct EstPartial {
    double sx, sy;
    float wmax;
    float theta;  // filled by k_estimate's last CTA: heading of particle imax when it lives on this device
    long long imax;
};

// What the tail of a resample leaves for the host (and, sharded, pushes into every rank's slab): this device's part
// of the pose estimate and whether its search met a slot it could not prove (k_search_keep).
struct TailBlob {
    EstPartial est;
    int anomaly;
    int pad[3];
};

// Device scalars of the resample with a pinned mirror on the host.
struct Scalars {
    float ws, s2, csum, neff;
    int did;                            // the decision of core.cpp:739, taken on the device: do_resample && nEff < NEFFECTIVE
    int comm_error;                     // sticky: a peer never arrived at an exchange (cleared only after it was reported)
    unsigned long long remote_sources;  // resampled slots whose particle came from another device (running total)
    int gauss_late;                     // sticky: a chunk of an in-flight noise upload never arrived (GaussArrival)
    int weight_resets;                  // resamples that met a zero / non-finite weight sum and went on unweighted
    int pad[6];
};

// Grid-wide barrier of the cooperative weights kernel: a counter that only ever grows; `base` is its value before the
// launch (the host knows it: every launch adds the same amount).
struct GridBar {
    unsigned *ctr;
    unsigned base;
};

#define SLAM_GATHER_ROWS 4  // landmark rows moved per thread by the resample gather

}  // namespace slam
