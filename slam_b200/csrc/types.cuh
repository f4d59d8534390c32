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
    float *slab[SLAM_MAX_RANKS];               // block-total arrays [4][nbg]: tot_w | tot_c | tot_s | bmax
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
#define SLAM_XCHG_BLOB_WORDS 8  // sizeof(EstPartial) / 4

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

#define SLAM_GATHER_ROWS 4  // landmark rows moved per thread by the resample gather

}  // namespace slam
