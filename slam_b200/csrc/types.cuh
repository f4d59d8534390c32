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

struct MapRing {  // landmark rows
    float2 *mean;  // [ring_rows][P]
    float4 *cov;   // [ring_rows][P], column-major linear 2x2 (core.cpp:613-615)
    int ring_rows;
    int ring_base;
};

struct EstPartial {
    double sx, sy;
    float wmax;
    float theta;  // filled by k_estimate_final: heading of particle imax when it lives on this device
    long long imax;
};

#define SLAM_GATHER_ROWS 4  // landmark rows moved per thread by the resample gather

}  // namespace slam
