// slam_rng.cuh — the device generator of throughput mode ("Philox in-kernel").
//
// The reference draws its noise INSIDE update/predict from rand() (core.cpp:392, 428, 776): per particle 4 uniforms
// for the pose sample (nRandMat::randn(3,1), core.cpp:383-419), then one uniform per particle for stratifiedRandom
// (core.cpp:767, 775-777); with control noise 3 uniforms per particle per predict tick (randn(2,1)).  Parity mode
// passes those draws in from the host (slamgpu_update's gauss / unif).  Throughput mode (slamgpu_seed) generates
// them here: a counter-based Philox4x32-10 keyed by the seed, counter = (global particle index, draw number, stream),
// so the numbers do not depend on how the particles are sharded, on the launch shape or on the kernel that asks.
// Everything downstream of the raw 32-bit words is the reference's own arithmetic: a word becomes a rand()-like int
// in [1, RAND_MAX] (0 would make logf(0) = -inf, a 2^-31 hazard the reference has and this generator avoids), a
// uniform is int / double(RAND_MAX), and the normals are the reference's Box-Muller pairing.
// oracle/philox.py restates the generator in numpy; tests compare the two word for word.
#pragma once

#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define SLAM_RNG_HD __host__ __device__ __forceinline__
#else
#define SLAM_RNG_HD inline
#endif

namespace slam {

#define SLAM_RNG_STREAM_POSE 0u      // 4 words -> the 3 normals of the pose sample
#define SLAM_RNG_STREAM_RESAMPLE 1u  // word 0 -> the resample uniform
#define SLAM_RNG_STREAM_CONTROL 2u   // 3 words -> the 2 normals of the control noise
#define SLAM_RAND_MAX 2147483647

SLAM_RNG_HD uint32_t rng_mulhi(uint32_t a, uint32_t b) {
#if defined(__CUDA_ARCH__)
    return __umulhi(a, b);
#else
    return (uint32_t) (((uint64_t) a * (uint64_t) b) >> 32);
#endif
}

// Philox4x32-10 (Salmon et al., SC'11): ten rounds, the key bumped by the Weyl constants between rounds.
SLAM_RNG_HD void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
#pragma unroll
    for (int r = 0; r < 10; r++) {
        const uint32_t hi0 = rng_mulhi(0xD2511F53u, c[0]), lo0 = 0xD2511F53u * c[0];
        const uint32_t hi1 = rng_mulhi(0xCD9E8D57u, c[2]), lo1 = 0xCD9E8D57u * c[2];
        const uint32_t n0 = hi1 ^ c[1] ^ k0, n2 = hi0 ^ c[3] ^ k1;
        c[0] = n0;
        c[1] = lo1;
        c[2] = n2;
        c[3] = lo0;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
}

SLAM_RNG_HD void rng_words(unsigned long long seed, unsigned long long slot, uint32_t draw, uint32_t stream,
                           uint32_t out[4]) {
    out[0] = (uint32_t) slot;
    out[1] = (uint32_t) (slot >> 32);
    out[2] = draw;
    out[3] = stream;
    philox4x32_10(out, (uint32_t) seed, (uint32_t) (seed >> 32));
}

// A rand()-like draw in [1, RAND_MAX].
SLAM_RNG_HD int rng_int(uint32_t word) {
    const int r = (int) (word >> 1);
    return r < 1 ? 1 : r;
}

// nRandMat::randn(count, 1) on `count + 1` rand() values (core.cpp:383-419), count <= 3.
template <int count>
SLAM_RNG_HD void rng_randn(const int *r, float *out) {
    float amp = 0.0f, angle = 0.0f;
#pragma unroll
    for (int k = 0; k < count; k++) {
        if ((k & 1) == 0) {
            // rand() * 1.0 / RAND_MAX stored as float (core.cpp:392-394).  The double product with the reciprocal is the
            // correctly rounded quotient except when that quotient sits within 2^-52 of a float rounding boundary
            // (a handful of the 2^31 ints, one float ulp): a double multiply instead of a ~30-instruction division
            const float uk = (float) ((double) r[k] * (1.0 / (double) SLAM_RAND_MAX));
            const float uk1 = (float) ((double) r[k + 1] * (1.0 / (double) SLAM_RAND_MAX));
            float square = -2.0f * logf(uk);  // exact doubling: the same float as the reference's double product
            if (square < 0.0f) square = 0.0f;
            amp = sqrtf(square);
            angle = (float) (2.0 * 3.14159265358979323846 * (double) uk1);
            out[k] = amp * sinf(angle);
        } else {
            out[k] = amp * cosf(angle);
        }
    }
}

SLAM_RNG_HD void rng_gauss3(unsigned long long seed, unsigned long long slot, uint32_t draw, float g[3]) {
    uint32_t w[4];
    rng_words(seed, slot, draw, SLAM_RNG_STREAM_POSE, w);
    const int r[4] = {rng_int(w[0]), rng_int(w[1]), rng_int(w[2]), rng_int(w[3])};
    rng_randn<3>(r, g);
}
SLAM_RNG_HD void rng_gauss2(unsigned long long seed, unsigned long long slot, uint32_t draw, float g[2]) {
    uint32_t w[4];
    rng_words(seed, slot, draw, SLAM_RNG_STREAM_CONTROL, w);
    const int r[3] = {rng_int(w[0]), rng_int(w[1]), rng_int(w[2])};
    rng_randn<2>(r, g);
}
// unifRand() = rand() / double(RAND_MAX) (core.cpp:775-777)
SLAM_RNG_HD double rng_unif(unsigned long long seed, unsigned long long slot, uint32_t draw) {
    uint32_t w[4];
    rng_words(seed, slot, draw, SLAM_RNG_STREAM_RESAMPLE, w);
    return (double) rng_int(w[0]) / (double) SLAM_RAND_MAX;
}

// Where a kernel takes its standard normals from: a staged buffer ([rows][P], written by the host) or the generator.
struct NoiseSrc {
    const float *gauss;       // nullptr when rng
    unsigned long long seed;
    unsigned long long slot0; // first global particle held by this device
    uint32_t draw;
    int rng;
};

#if defined(__CUDACC__)
__device__ __forceinline__ void noise_gauss3(const NoiseSrc &n, size_t P, size_t p, float g[3]) {
    if (n.rng) {
        rng_gauss3(n.seed, n.slot0 + p, n.draw, g);
    } else {
        g[0] = n.gauss[p];
        g[1] = n.gauss[P + p];
        g[2] = n.gauss[2 * P + p];
    }
}
__device__ __forceinline__ void noise_gauss2(const NoiseSrc &n, size_t P, size_t p, float g[2]) {
    if (n.rng) {
        rng_gauss2(n.seed, n.slot0 + p, n.draw, g);
    } else {
        g[0] = n.gauss[p];
        g[1] = n.gauss[P + p];
    }
}
#endif

}  // namespace slam
