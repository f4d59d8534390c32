/* TEST INFRASTRUCTURE ONLY — CPU restatement of the reference's FastSLAM hot path.
 *
 * Plain-C restatement, on flat SoA arrays, of the functions SURVEY.md §8(a) lists
 * (reference files cited per function in slam_oracle.c).  It exists so that the
 * CUDA path can be checked at sizes the unmodified reference cannot run
 * (power-of-two / 1M particle counts abort at core.cpp:762; cumulativeSum is O(N^2)).
 *
 * Parity pinning: tests/test_oracle_vs_ref.py compares every function here with the
 * unmodified reference compiled from /root/reference (oracle/_ref/libslamref.so) and
 * tests/golden/ holds vectors produced by that reference (tests/golden/make_golden.py).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library.  The product (slam_b200/libslamgpu.so) never does.
 *
 * Flat layout (same as include/slamgpu.h):
 *   xv[c*P+p] c<3 | Pv[(r*3+c)*P+p] | w[p] | lm_mean[(l*P+p)*2+j] |
 *   lm_cov[(l*P+p)*4+j], j = column-major linear index of the 2x2 covariance.
 */
#ifndef SLAM_ORACLE_H
#define SLAM_ORACLE_H

#ifdef __cplusplus
extern "C" {
#endif

#define SO_RAND_MAX 2147483647

float so_trig_offset(float ang);
/* Box-Muller of core.cpp:383-419 over count+1 raw rand() values; returns values consumed. */
int so_randn(const int *rnd, int count, float *out);
float so_eigen_sum(const float *v, long n);
void so_compute_jacobians(const float *in, int n, float *out);
float so_gauss_evaluate(const float *v, const float *S, int D);
void so_cholesky_update(float *x, float *P, const float *v, const float *R, const float *H);
void so_joseph_update(float *x, float *P, float v, float R);
void so_multivariate_gauss(int n, const float *x, const float *P, const float *g, float *out);

/* gauss: NULL (no predict noise) or [2][P] standard normals (row 0 for V, row 1 for G). */
void so_fs2_predict(long P, float *xv, float *Pv, float V, float G, const float *Q, float dt, float phi_true,
                    float wheel_base, float sigma_phi, int use_heading, const float *gauss);
void so_fs1_predict(long P, float *xv, float *Pv, float V, float G, const float *Q, float dt, float phi_true,
                    float wheel_base, float sigma_phi, int use_heading, const float *gauss);

/* Per-particle part of FastSLAM2::update.  gauss = [3][P] standard normals for the pose sample.
 * lm_* must have room for Nf+Kn rows.  Returns the new landmark count. */
int so_fs2_update_particles(long P, int Nf, float *xv, float *Pv, float *w, float *lm_mean, float *lm_cov,
                            const float *zf, const int *idf, int K, const float *zn, int Kn, const float *R,
                            const float *gauss);
int so_fs1_update_particles(long P, int Nf, float *xv, float *Pv, float *w, float *lm_mean, float *lm_cov,
                            const float *zf, const int *idf, int K, const float *zn, int Kn, const float *R);

/* Number of strata the reference's own loop emits (== N where the reference is valid). */
long so_stratified_count(long N);
/* di[] of stratifiedRandom before dithering, loop run exactly N times. */
void so_stratified_di(long N, float *di);

/* scan_mode 0: reference order (Eigen redux sums, sequential fp32 CDF).
 * scan_mode 1: fixed-shape blocked order (blocks of SO_SCAN_BLOCK, see slam_oracle.c).
 * w is normalised in place; unif = rand()/double(RAND_MAX), one per particle.
 * keep[] may contain -1 exactly where the reference leaves it. */
#define SO_SCAN_BLOCK 1024
void so_resample_indices(long P, float *w, const double *unif, int scan_mode, int *keep, float *neff,
                         float *cdf_out /* nullable */);

/* particles[i] = old[keep[i]] over the whole flat state (keep<0 clamps to P-1), w = 1/P. */
void so_resample_gather(long P, int Nf, const int *keep, float *xv, float *Pv, float *w, float *lm_mean,
                        float *lm_cov);

void so_estimate(long P, const float *xv, const float *w, double *x, double *y, double *t);

#ifdef __cplusplus
}
#endif
#endif
