// TEST INFRASTRUCTURE ONLY — not part of the product path.
//
// Flat-array C shim over the UNMODIFIED reference sources, which are compiled
// in place from /root/reference by oracle/Makefile into oracle/_ref/libslamref.so
// (git-ignored; no reference source is copied into this repo).  Only tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
// may load it.
//
// What it gives the tests:
//   * an interposed rand()/srand() (linked -Bsymbolic) that replays an injected
//     integer stream, so the reference and the CUDA path see identical noise
//     (all reference randomness goes through std::rand: core.cpp:392,428,776);
//   * vector<Particle>  <->  the flat SoA layout of include/slamgpu.h;
//   * the reference entry points FastSLAM1/2::predict, ::update
//     (algorithms/fastslam1.cpp:18-66, fastslam2.cpp:21-60), their protected
//     members through a Probe subclass, and the free functions of core.cpp.
//
// Flat layout (identical to slamgpu_upload_particles):
//   xv[c*P+p] c<3 | Pv[(r*3+c)*P+p] | w[p] | lm_mean[(l*P+p)*2+j] |
//   lm_cov[(l*P+p)*4+j], j = Eigen linear (column-major) index of the 2x2.
#include <cstdlib>
#include <cstring>
#include <vector>

#include <Eigen/Dense>

#include "src/backend/core.h"
#include "src/backend/Particle.h"
#include "src/backend/algorithms/fastslam1.h"
#include "src/backend/algorithms/fastslam2.h"

using namespace Eigen;
using std::vector;

// ---------------------------------------------------------------------------
// rand() interposition
// ---------------------------------------------------------------------------
static std::vector<int> g_stream;
static long g_pos = 0;
static long g_underflow = 0;
static unsigned long g_lcg = 1;

extern "C" int rand(void) {
    if (g_pos < (long) g_stream.size()) return g_stream[g_pos++];
    g_underflow++;
    // deterministic fallback so a mis-sized stream never crashes the test process
    g_lcg = g_lcg * 6364136223846793005ULL + 1442695040888963407ULL;
    return (int) ((g_lcg >> 33) & 0x7fffffff);
}
extern "C" void srand(unsigned s) { g_lcg = s; }

extern "C" void slamref_rand_set_stream(const int *v, long n) {
    g_stream.assign(v, v + n);
    g_pos = 0;
    g_underflow = 0;
}
extern "C" long slamref_rand_pos() { return g_pos; }
extern "C" void slamref_rand_seek(long p) { g_pos = p; }
extern "C" long slamref_rand_underflow() { return g_underflow; }
extern "C" int slamref_rand_max() { return RAND_MAX; }

// ---------------------------------------------------------------------------
// helpers
// ---------------------------------------------------------------------------
namespace {

struct Probe2 : public FastSLAM2 {
    using FastSLAM2::sampleProposal;
    using FastSLAM2::likelihoodGivenXv;
    using FastSLAM2::gaussEvaluate;
    using FastSLAM2::predictState;
    using FastSLAM2::observeHeading;
};
struct Probe1 : public FastSLAM1 {
    using FastSLAM1::computeWeight;
    using FastSLAM1::predictState;
    using FastSLAM1::observeHeading;
};

struct RefState {
    vector<Particle> particles;
};

MatrixXf mat2(const float *m) {  // row-major 2x2 in
    MatrixXf M(2, 2);
    M << m[0], m[1], m[2], m[3];
    return M;
}

void obs_vectors(const float *z, int K, vector<VectorXf> &out) {
    out.clear();
    for (int i = 0; i < K; i++) {
        VectorXf v(2);
        v << z[2 * i], z[2 * i + 1];
        out.push_back(v);
    }
}

template <class A>
void configure(A &a, float wheelBase, float sigmaPhi, int useHeading, int addPredictNoise, int nEffective,
               int resample) {
    a.wheelBase = wheelBase;
    a.sigmaPhi = sigmaPhi;
    a.useHeading = useHeading != 0;
    a.addPredictNoise = addPredictNoise != 0;
    a.nEffective = nEffective;
    a.resample = resample != 0;
}

}  // namespace

extern "C" {

// ---------------------------------------------------------------------------
// primitives of core.cpp
// ---------------------------------------------------------------------------
void slamref_randn(int m, int n, float *out) {  // core.cpp:383-419, row-major out
    MatrixXf x = nRandMat::randn(m, n);
    for (int i = 0; i < m; i++)
        for (int j = 0; j < n; j++) out[i * n + j] = x(i, j);
}

float slamref_trig_offset(float a) { return trigonometricOffset(a); }  // core.cpp:460-477

double slamref_unif_rand() { return unifRand(); }  // core.cpp:775-777

void slamref_multivariate_gauss(int n, const float *x, const float *P, float *out) {  // core.cpp:452-458
    VectorXf xv(n);
    MatrixXf Pm(n, n);
    for (int i = 0; i < n; i++) {
        xv(i) = x[i];
        for (int j = 0; j < n; j++) Pm(i, j) = P[i * n + j];
    }
    VectorXf r = multivariateGauss(xv, Pm, 1);
    for (int i = 0; i < n; i++) out[i] = r(i);
}

float slamref_sum(const float *w, int n) {  // Eigen redux as used at core.cpp:726,782,788
    VectorXf v(n);
    for (int i = 0; i < n; i++) v(i) = w[i];
    return v.sum();
}

void slamref_cumsum(float *w, int n) {  // core.cpp:813-824 (O(n^2))
    VectorXf v(n);
    for (int i = 0; i < n; i++) v(i) = w[i];
    cumulativeSum(v);
    for (int i = 0; i < n; i++) w[i] = v(i);
}

// Number of strata the reference's own loop would emit (core.cpp:751-759), so the caller can
// avoid the process-killing assert at core.cpp:762.  Restated here, not called.
long slamref_stratified_count(long N) {
    float k = 1.0 / (float) N;
    float temp = k / 2;
    long c = 0;
    while (temp < (1 - k / 2)) {
        c++;
        temp = temp + k;
        if (c > N + 8) break;
    }
    return c;
}

// core.cpp:751-769. Consumes N rand(). Caller must ensure slamref_stratified_count(N)==N.
void slamref_stratified_random(long N, float *di) {
    vector<float> v;
    stratifiedRandom((unsigned long) N, v);
    for (long i = 0; i < N; i++) di[i] = v[i];
}

// core.cpp:780-807. Consumes n rand().
void slamref_stratified_resample(const float *w, int n, int *keep, float *neff) {
    VectorXf v(n);
    for (int i = 0; i < n; i++) v(i) = w[i];
    vector<int> k;
    float ne = 0;
    stratifiedResample(v, k, ne);
    for (int i = 0; i < n; i++) keep[i] = k[i];
    *neff = ne;
}

// The accelerator record of core.cpp:586-664:
//   in : xv[3], R(0..3) linear (column-major), then n x { xf[2], Pf(0..3) linear }
//   out: n x { zp[2], Hf[4] row-major, Hv[6] row-major, Sf[4] row-major }
// computed by the CPU path core.cpp:666-713.
void slamref_compute_jacobians(const float *in, int n, float *out) {
    Particle p;
    VectorXf xv(3);
    xv << in[0], in[1], in[2];
    p.setXv(xv);
    MatrixXf R(2, 2);
    for (int j = 0; j < 4; j++) R(j) = in[3 + j];
    vector<int> idf;
    for (int i = 0; i < n; i++) {
        VectorXf xf(2);
        xf << in[7 + 6 * i], in[7 + 6 * i + 1];
        MatrixXf Pf(2, 2);
        for (int j = 0; j < 4; j++) Pf(j) = in[7 + 6 * i + 2 + j];
        p.setLandmarkX(i, xf);
        p.setLandmarkP(i, Pf);
        idf.push_back(i);
    }
    vector<VectorXf> zp;
    vector<MatrixXf> Hv, Hf, Sf;
    computeJacobians(p, idf, R, &zp, &Hv, &Hf, &Sf);
    for (int i = 0; i < n; i++) {
        float *o = out + 16 * i;
        o[0] = zp[i](0);
        o[1] = zp[i](1);
        o[2] = Hf[i](0, 0); o[3] = Hf[i](0, 1); o[4] = Hf[i](1, 0); o[5] = Hf[i](1, 1);
        o[6] = Hv[i](0, 0); o[7] = Hv[i](0, 1); o[8] = Hv[i](0, 2);
        o[9] = Hv[i](1, 0); o[10] = Hv[i](1, 1); o[11] = Hv[i](1, 2);
        o[12] = Sf[i](0, 0); o[13] = Sf[i](0, 1); o[14] = Sf[i](1, 0); o[15] = Sf[i](1, 1);
    }
}

// fastslam2.cpp:127-163 (logflag=0). v[D], S row-major DxD.
float slamref_gauss_evaluate(const float *v, const float *S, int D) {
    Probe2 a;
    VectorXf vv(D);
    MatrixXf Sm(D, D);
    for (int i = 0; i < D; i++) {
        vv(i) = v[i];
        for (int j = 0; j < D; j++) Sm(i, j) = S[i * D + j];
    }
    return a.gaussEvaluate(vv, Sm, 0);
}

// core.cpp:275-291. x[2], P linear(col-major)[4], v[2], R row-major, H row-major; in place.
void slamref_cholesky_update(float *x, float *P, const float *v, const float *R, const float *H) {
    VectorXf xv(2), vv(2);
    xv << x[0], x[1];
    vv << v[0], v[1];
    MatrixXf Pm(2, 2);
    for (int j = 0; j < 4; j++) Pm(j) = P[j];
    MatrixXf Rm = mat2(R), Hm = mat2(H);
    choleskyUpdate(xv, Pm, vv, Rm, Hm);
    x[0] = xv(0);
    x[1] = xv(1);
    for (int j = 0; j < 4; j++) P[j] = Pm(j);
}

// core.cpp:294-317 with H=[0 0 1]. x[3], P row-major 3x3; in place.
void slamref_joseph_update(float *x, float *P, float v, float R) {
    VectorXf xv(3);
    MatrixXf Pm(3, 3), H(1, 3);
    H << 0, 0, 1;
    for (int i = 0; i < 3; i++) {
        xv(i) = x[i];
        for (int j = 0; j < 3; j++) Pm(i, j) = P[i * 3 + j];
    }
    josephUpdate(xv, Pm, v, R, H);
    for (int i = 0; i < 3; i++) {
        x[i] = xv(i);
        for (int j = 0; j < 3; j++) P[i * 3 + j] = Pm(i, j);
    }
}

// ---------------------------------------------------------------------------
// particle-set handle
// ---------------------------------------------------------------------------
void *slamref_create(int P) {
    RefState *s = new RefState();
    s->particles = vector<Particle>(P);
    float u = 1.0 / (float) P;  // ParticleSLAMWrapper.cpp:14-25
    for (int p = 0; p < P; p++) {
        s->particles[p] = Particle();
        s->particles[p].setW(u);
    }
    return s;
}

void slamref_destroy(void *h) { delete (RefState *) h; }

int slamref_num_particles(void *h) { return (int) ((RefState *) h)->particles.size(); }

int slamref_num_landmarks(void *h) {
    RefState *s = (RefState *) h;
    return s->particles.empty() ? 0 : (int) s->particles[0].landmarkXs().size();
}

void slamref_upload(void *h, int Nf, const float *xv, const float *Pv, const float *w, const float *lm_mean,
                    const float *lm_cov) {
    RefState *s = (RefState *) h;
    long P = (long) s->particles.size();
    for (long p = 0; p < P; p++) {
        VectorXf x(3);
        MatrixXf Pm(3, 3);
        for (int c = 0; c < 3; c++) x(c) = xv[c * P + p];
        for (int r = 0; r < 3; r++)
            for (int c = 0; c < 3; c++) Pm(r, c) = Pv[(r * 3 + c) * P + p];
        vector<VectorXf> xs;
        vector<MatrixXf> ps;
        for (long l = 0; l < Nf; l++) {
            VectorXf m(2);
            MatrixXf C(2, 2);
            m << lm_mean[(l * P + p) * 2], lm_mean[(l * P + p) * 2 + 1];
            for (int j = 0; j < 4; j++) C(j) = lm_cov[(l * P + p) * 4 + j];
            xs.push_back(m);
            ps.push_back(C);
        }
        float ww = w[p];
        s->particles[p] = Particle(ww, x, Pm, xs, ps);
    }
}

void slamref_download(void *h, float *xv, float *Pv, float *w, float *lm_mean, float *lm_cov) {
    RefState *s = (RefState *) h;
    long P = (long) s->particles.size();
    for (long p = 0; p < P; p++) {
        Particle &q = s->particles[p];
        if (w) w[p] = q.w();
        for (int c = 0; c < 3; c++)
            if (xv) xv[c * P + p] = q.xv()(c);
        for (int r = 0; r < 3; r++)
            for (int c = 0; c < 3; c++)
                if (Pv) Pv[(r * 3 + c) * P + p] = q.Pv()(r, c);
        long Nf = (long) q.landmarkXs().size();
        for (long l = 0; l < Nf; l++) {
            if (lm_mean) {
                lm_mean[(l * P + p) * 2] = q.landmarkXs()[l](0);
                lm_mean[(l * P + p) * 2 + 1] = q.landmarkXs()[l](1);
            }
            if (lm_cov)
                for (int j = 0; j < 4; j++) lm_cov[(l * P + p) * 4 + j] = q.landmarkPs()[l](j);
        }
    }
}

// ---------------------------------------------------------------------------
// FastSLAM2 (algorithms/fastslam2.cpp)
// ---------------------------------------------------------------------------
void slamref_fs2_predict(void *h, float V, float G, const float *Q, float dt, float phiTrue, float wheelBase,
                         float sigmaPhi, int useHeading, int addPredictNoise) {
    RefState *s = (RefState *) h;
    FastSLAM2 a;
    configure(a, wheelBase, sigmaPhi, useHeading, addPredictNoise, 0, 0);
    MatrixXf Qm = mat2(Q);
    VectorXf xTrue(3);
    xTrue << 0, 0, phiTrue;
    a.predict(s->particles, xTrue, V, G, Qm, dt);  // fastslam2.cpp:51-60
}

// The whole of FastSLAM2::update (fastslam2.cpp:21-48), resample included.
void slamref_fs2_update(void *h, const float *zf, const int *idf, int K, const float *zn, int Kn, const float *R,
                        int nEffective, int resample) {
    RefState *s = (RefState *) h;
    FastSLAM2 a;
    configure(a, 0, 0, 0, 0, nEffective, resample);
    vector<VectorXf> zfv, znv, z;
    obs_vectors(zf, K, zfv);
    obs_vectors(zn, Kn, znv);
    vector<int> idfv(idf, idf + K);
    VectorXf table(1);
    MatrixXf Rm = mat2(R);
    a.update(s->particles, zfv, znv, idfv, z, table, Rm);
}

// Per-particle part of FastSLAM2::update only (fastslam2.cpp:26-45), no resample.
void slamref_fs2_update_particles(void *h, const float *zf, const int *idf, int K, const float *zn, int Kn,
                                  const float *R) {
    RefState *s = (RefState *) h;
    Probe2 a;
    vector<VectorXf> zfv, znv;
    obs_vectors(zf, K, zfv);
    obs_vectors(zn, Kn, znv);
    vector<int> idfv(idf, idf + K);
    MatrixXf Rm = mat2(R);
    for (size_t i = 0; i < s->particles.size(); i++) {
        if (!zfv.empty()) {
            a.sampleProposal(s->particles[i], zfv, idfv, Rm);
            featureUpdate(s->particles[i], zfv, idfv, Rm);
        }
        if (!znv.empty()) {
            if (zfv.empty()) {
                VectorXf xv = multivariateGauss(s->particles[i].xv(), s->particles[i].Pv(), 1);
                s->particles[i].setXv(xv);
                MatrixXf pv(3, 3);
                pv.setZero();
                s->particles[i].setPv(pv);
            }
            addFeature(s->particles[i], znv, Rm);
        }
    }
}

// Probe: sampleProposal on one particle (fastslam2.cpp:290-368).
void slamref_fs2_sample_proposal(void *h, int p, const float *zf, const int *idf, int K, const float *R) {
    RefState *s = (RefState *) h;
    Probe2 a;
    vector<VectorXf> zfv;
    obs_vectors(zf, K, zfv);
    vector<int> idfv(idf, idf + K);
    MatrixXf Rm = mat2(R);
    a.sampleProposal(s->particles[p], zfv, idfv, Rm);
}

// ---------------------------------------------------------------------------
// FastSLAM1 (algorithms/fastslam1.cpp)
// ---------------------------------------------------------------------------
void slamref_fs1_predict(void *h, float V, float G, const float *Q, float dt, float phiTrue, float wheelBase,
                         float sigmaPhi, int useHeading, int addPredictNoise) {
    RefState *s = (RefState *) h;
    FastSLAM1 a;
    configure(a, wheelBase, sigmaPhi, useHeading, addPredictNoise, 0, 0);
    MatrixXf Qm = mat2(Q);
    VectorXf xTrue(3);
    xTrue << 0, 0, phiTrue;
    a.predict(s->particles, xTrue, V, G, Qm, dt);  // fastslam1.cpp:57-66
}

void slamref_fs1_update(void *h, const float *zf, const int *idf, int K, const float *zn, int Kn, const float *R,
                        int nEffective, int resample) {
    RefState *s = (RefState *) h;
    FastSLAM1 a;
    configure(a, 0, 0, 0, 0, nEffective, resample);
    vector<VectorXf> zfv, znv;
    obs_vectors(zf, K, zfv);
    obs_vectors(zn, Kn, znv);
    vector<int> idfv(idf, idf + K), vis;
    VectorXf table(1);
    MatrixXf Rm = mat2(R);
    a.update(s->particles, zfv, znv, idfv, vis, table, Rm);  // fastslam1.cpp:18-35
}

void slamref_fs1_update_particles(void *h, const float *zf, const int *idf, int K, const float *zn, int Kn,
                                  const float *R) {
    RefState *s = (RefState *) h;
    Probe1 a;
    vector<VectorXf> zfv, znv;
    obs_vectors(zf, K, zfv);
    obs_vectors(zn, Kn, znv);
    vector<int> idfv(idf, idf + K);
    MatrixXf Rm = mat2(R);
    for (size_t i = 0; i < s->particles.size(); i++) {
        if (!zfv.empty()) {
            float w = a.computeWeight(s->particles[i], zfv, idfv, Rm);
            w = s->particles[i].w() * w;
            s->particles[i].setW(w);
            featureUpdate(s->particles[i], zfv, idfv, Rm);
        }
        if (!znv.empty()) addFeature(s->particles[i], znv, Rm);
    }
}

// resampleParticles verbatim (core.cpp:718-749).  Caller must check slamref_stratified_count.
void slamref_resample(void *h, int nMin, int doResample) {
    RefState *s = (RefState *) h;
    resampleParticles(s->particles, nMin, doResample != 0);
}

}  // extern "C"
