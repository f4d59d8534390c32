// TEST INFRASTRUCTURE — slam_b200/csrc/slam_rng.cuh (the device generator of throughput mode) compiled for the host,
// so that its words / ints / normals can be diffed against oracle/philox.py on the build container (no GPU there).
#include "../../slam_b200/csrc/slam_rng.cuh"

extern "C" {
void re_philox(const unsigned *ctr, const unsigned *key, unsigned *out) {
    for (int i = 0; i < 4; i++) out[i] = ctr[i];
    slam::philox4x32_10(out, key[0], key[1]);
}
void re_words(unsigned long long seed, unsigned long long slot, unsigned draw, unsigned stream, unsigned *out) {
    slam::rng_words(seed, slot, draw, stream, out);
}
void re_gauss3(unsigned long long seed, long n, unsigned draw, float *out) {  // out [3][n], slots 0..n-1
    for (long p = 0; p < n; p++) {
        float g[3];
        slam::rng_gauss3(seed, (unsigned long long) p, draw, g);
        for (int c = 0; c < 3; c++) out[c * n + p] = g[c];
    }
}
void re_gauss2(unsigned long long seed, long n, unsigned draw, float *out) {
    for (long p = 0; p < n; p++) {
        float g[2];
        slam::rng_gauss2(seed, (unsigned long long) p, draw, g);
        for (int c = 0; c < 2; c++) out[c * n + p] = g[c];
    }
}
void re_unif(unsigned long long seed, long n, unsigned draw, double *out) {
    for (long p = 0; p < n; p++) out[p] = slam::rng_unif(seed, (unsigned long long) p, draw);
}
}
