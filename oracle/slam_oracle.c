/* TEST INFRASTRUCTURE ONLY — CPU restatement of the reference's FastSLAM hot path.
 *
 * See slam_oracle.h for the role of this file and who may load it.  Every function
 * cites the reference lines (relative to /root/reference/) it follows.  The
 * arithmetic mirrors the reference's float/double promotions and, where cheap, the
 * operation order of the vendored Eigen 3.1.3 kernels the reference calls, so that
 * most functions agree with oracle/_ref/libslamref.so bit for bit
 * (tests/test_oracle_vs_ref.py states which do and the bound for those that do not).
 *
 * Build: gcc -std=c99 -O2 -msse2 -mfpmath=sse -ffp-contract=off (oracle/Makefile).
 */
#include "slam_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

/* ------------------------------------------------------------------------- */
/* small dense helpers (row-major, k-sequential accumulation like Eigen's gebp/gemv) */
/* ------------------------------------------------------------------------- */
/* C[m x n] = A[m x k] * B[k x n] */
static void mm(int m, int k, int n, const float *A, const float *B, float *C) {
    for (int i = 0; i < m; i++)
        for (int j = 0; j < n; j++) {
            float acc = A[i * k] * B[j];
            for (int t = 1; t < k; t++) acc = acc + A[i * k + t] * B[t * n + j];
            C[i * n + j] = acc;
        }
}
/* C[m x n] = A[m x k] * B^T, B is [n x k] */
static void mmt(int m, int k, int n, const float *A, const float *B, float *C) {
    for (int i = 0; i < m; i++)
        for (int j = 0; j < n; j++) {
            float acc = A[i * k] * B[j * k];
            for (int t = 1; t < k; t++) acc = acc + A[i * k + t] * B[j * k + t];
            C[i * n + j] = acc;
        }
}

/* Eigen LLT, lower, unblocked (libs/eigen3/Eigen/src/Cholesky/LLT.h:263-287).  A row-major n x n,
 * factor written into the lower triangle in place.  Returns -1 on success, else the column at
 * which it stopped (the rest of the matrix is left as it was, exactly like Eigen). */
static int llt_lower(int n, float *A) {
    for (int k = 0; k < n; k++) {
        float x = A[k * n + k];
        if (k > 0) {
            float sq = A[k * n] * A[k * n];
            for (int t = 1; t < k; t++) sq = sq + A[k * n + t] * A[k * n + t];
            x = x - sq;
        }
        if (x <= 0.0f) return k;
        x = sqrtf(x);
        A[k * n + k] = x;
        int rs = n - k - 1;
        if (k > 0 && rs > 0) {
            for (int i = k + 1; i < n; i++) {
                /* A21 -= A20 * A10^T : column-major gemv, one column of A20 at a time */
                for (int t = 0; t < k; t++) A[i * n + k] = A[i * n + k] - A[i * n + t] * A[k * n + t];
            }
        }
        if (rs > 0) {
            float r = 1.0f / x;
            for (int i = k + 1; i < n; i++) A[i * n + k] = A[i * n + k] * r;
        }
    }
    return -1;
}

/* matrixL(): lower triangle of the (possibly partial) factor, zeros above. */
static void take_lower(int n, const float *A, float *L) {
    for (int i = 0; i < n; i++)
        for (int j = 0; j < n; j++) L[i * n + j] = (j <= i) ? A[i * n + j] : 0.0f;
}

/* A.llt().solve(Identity)  (LLT.h:448-454 + Core/products/TriangularSolverMatrix.h:104-134):
 * forward substitution in axpy form on the column-major lower factor, backward substitution in
 * dot form on its adjoint; both multiply by the reciprocal of the diagonal. */
static void llt_inverse(int n, const float *Ain, float *X) {
    float A[9];
    memcpy(A, Ain, sizeof(float) * n * n);
    llt_lower(n, A);
    for (int i = 0; i < n; i++)
        for (int j = 0; j < n; j++) X[i * n + j] = (i == j) ? 1.0f : 0.0f;
    for (int k = 0; k < n; k++) { /* L y = b */
        float a = 1.0f / A[k * n + k];
        for (int j = 0; j < n; j++) {
            float b = X[k * n + j] * a;
            X[k * n + j] = b;
            for (int i = k + 1; i < n; i++) X[i * n + j] = X[i * n + j] - b * A[i * n + k];
        }
    }
    for (int k = 0; k < n; k++) { /* L^T x = y */
        int i = n - k - 1;
        float a = 1.0f / A[i * n + i];
        for (int j = 0; j < n; j++) {
            float b = 0.0f;
            for (int t = 0; t < k; t++) b = b + A[(i + 1 + t) * n + i] * X[(i + 1 + t) * n + j];
            X[i * n + j] = (X[i * n + j] - b) * a;
        }
    }
}

/* Dynamic-size MatrixXf::inverse() for a 2x2 = partialPivLu().solve(I)
 * (libs/eigen3/Eigen/src/LU/Inverse.h, PartialPivLU.h lu_unblocked + triangular solves). */
static void inv2_lu(const float *A, float *X, float *det) {
    float a = A[0], b = A[1], c = A[2], d = A[3];
    int swap = fabsf(c) > fabsf(a);
    if (swap) {
        float t = a; a = c; c = t;
        t = b; b = d; d = t;
    }
    float l = c / a;
    float u11 = d - l * b;
    float B[4];
    if (swap) { B[0] = 0; B[1] = 1; B[2] = 1; B[3] = 0; }
    else      { B[0] = 1; B[1] = 0; B[2] = 0; B[3] = 1; }
    for (int j = 0; j < 2; j++) { /* unit lower */
        float bb = B[j] * 1.0f;
        B[j] = bb;
        B[2 + j] = B[2 + j] - bb * l;
    }
    for (int j = 0; j < 2; j++) { /* upper, axpy form */
        float a1 = 1.0f / u11;
        float bb = B[2 + j] * a1;
        B[2 + j] = bb;
        B[j] = B[j] - bb * b;
    }
    {
        float a0 = 1.0f / a;
        B[0] = B[0] * a0;
        B[1] = B[1] * a0;
    }
    memcpy(X, B, sizeof(B));
    if (det) *det = (swap ? -1.0f : 1.0f) * (a * u11);
}

/* ------------------------------------------------------------------------- */
/* core.cpp primitives */
/* ------------------------------------------------------------------------- */
float so_trig_offset(float ang) { /* core.cpp:460-477 */
    int n;
    if ((ang < -2 * M_PI) || (ang > 2 * M_PI)) {
        n = (int) floor(ang / (2 * M_PI));
        ang = (float) (ang - n * (2 * M_PI));
    }
    if (ang > M_PI) ang = (float) (ang - (2 * M_PI));
    if (ang < -M_PI) ang = (float) (ang + (2 * M_PI));
    return ang;
}

int so_randn(const int *rnd, int count, float *out) { /* core.cpp:383-419 */
    int urows = count + 1;
    float amp = 0, angle = 0, square;
    for (int k = 0; k < count; k++) {
        if (k % 2 == 0) {
            float uk = (float) (rnd[k] * 1.0 / SO_RAND_MAX);
            float uk1 = (float) (rnd[k + 1] * 1.0 / SO_RAND_MAX);
            square = (float) (-2. * logf(uk));
            if (square < 0.) square = 0.;
            amp = sqrtf(square);
            angle = (float) (2. * M_PI * uk1);
            out[k] = amp * sinf(angle);
        } else {
            out[k] = amp * cosf(angle);
        }
    }
    return urows;
}

float so_eigen_sum(const float *v, long n) { /* Redux.h:193-248 with the SSE2 predux of PacketMath.h:406-410 */
    const long ps = 4;
    long size2 = (n / (2 * ps)) * (2 * ps);
    long size1 = (n / ps) * ps;
    float res;
    if (size1) {
        float p0[4], p1[4];
        for (int j = 0; j < 4; j++) p0[j] = v[j];
        if (size1 > ps) {
            for (int j = 0; j < 4; j++) p1[j] = v[ps + j];
            for (long idx = 2 * ps; idx < size2; idx += 2 * ps)
                for (int j = 0; j < 4; j++) {
                    p0[j] = p0[j] + v[idx + j];
                    p1[j] = p1[j] + v[idx + ps + j];
                }
            for (int j = 0; j < 4; j++) p0[j] = p0[j] + p1[j];
            if (size1 > size2)
                for (int j = 0; j < 4; j++) p0[j] = p0[j] + v[size2 + j];
        }
        res = (p0[0] + p0[2]) + (p0[1] + p0[3]);
        for (long idx = size1; idx < n; idx++) res = res + v[idx];
    } else {
        res = v[0];
        for (long idx = 1; idx < n; idx++) res = res + v[idx];
    }
    return res;
}

/* core.cpp:666-713 on the accelerator record of core.cpp:586-664:
 *   in : xv[3], R(0..3) column-major linear, then n x { xf[2], Pf(0..3) column-major linear }
 *   out: n x { zp[2], Hf[4] row-major, Hv[6] row-major, Sf[4] row-major } */
static void jacobian_one(const float *xv, const float *R /* row-major */, const float *xf,
                         const float *Pf /* row-major */, float *zp, float *Hf, float *Hv, float *Sf) {
    float dx = xf[0] - xv[0];
    float dy = xf[1] - xv[1];
    float d2 = (float) ((double) dx * (double) dx + (double) dy * (double) dy);
    float d = sqrtf(d2);
    zp[0] = d;
    zp[1] = so_trig_offset(atan2f(dy, dx) - xv[2]);
    Hv[0] = -dx / d; Hv[1] = -dy / d; Hv[2] = 0;
    Hv[3] = dy / d2; Hv[4] = -dx / d2; Hv[5] = -1;
    Hf[0] = dx / d; Hf[1] = dy / d; Hf[2] = -dy / d2; Hf[3] = dx / d2;
    float T[4];
    mm(2, 2, 2, Hf, Pf, T);
    mmt(2, 2, 2, T, Hf, Sf);
    for (int j = 0; j < 4; j++) Sf[j] = Sf[j] + R[j];
}

void so_compute_jacobians(const float *in, int n, float *out) {
    const float *xv = in;
    float R[4] = {in[3], in[5], in[4], in[6]};
    for (int i = 0; i < n; i++) {
        const float *r = in + 7 + 6 * i;
        float Pf[4] = {r[2], r[4], r[3], r[5]};
        float *o = out + 16 * i;
        jacobian_one(xv, R, r, Pf, o, o + 2, o + 6, o + 12);
    }
}

/* fastslam2.cpp:127-163, logflag = 0.  The reference solves Sc*nin = v through a Jacobi SVD of the
 * triangular factor; restated as the forward substitution it equals mathematically (agreement with
 * the linked reference is ~1e-6 relative, not bitwise; see tests/test_oracle_vs_ref.py). */
float so_gauss_evaluate(const float *v, const float *S, int D) {
    float A[9], nin[3];
    memcpy(A, S, sizeof(float) * D * D);
    llt_lower(D, A);
    for (int i = 0; i < D; i++) {
        float b = v[i];
        for (int t = 0; t < i; t++) b = b - A[i * D + t] * nin[t];
        nin[i] = b / A[i * D + i];
    }
    float E = 0;
    for (int s = 0; s < D; s++) {
        float q = (float) ((double) nin[s] * (double) nin[s]);
        E += q;
    }
    E = (float) (-0.5 * E);
    float prod = 1;
    for (int i = 0; i < D; i++) prod = prod * A[i * D + i];
    float C = (float) (pow(2 * M_PI, (double) (D / 2)) * prod);
    return expf(E) / C;
}

/* core.cpp:275-291.  x[2]; P column-major linear [4] (in/out); v[2]; R, H row-major. */
void so_cholesky_update(float *x, float *Plin, const float *v, const float *R, const float *H) {
    float P[4] = {Plin[0], Plin[2], Plin[1], Plin[3]};
    float PHt[4], S[4], HP[4];
    mmt(2, 2, 2, P, H, PHt);
    mm(2, 2, 2, H, PHt, HP);
    for (int j = 0; j < 4; j++) S[j] = HP[j] + R[j];
    float Ss[4] = {(S[0] + S[0]) * 0.5f, (S[1] + S[2]) * 0.5f, (S[2] + S[1]) * 0.5f, (S[3] + S[3]) * 0.5f};
    llt_lower(2, Ss);
    float U[4] = {Ss[0], Ss[2], 0.0f, Ss[3]}; /* matrixU() = adjoint of the lower factor */
    float Ui[4];
    inv2_lu(U, Ui, 0);
    float W1[4], W[4];
    mm(2, 2, 2, PHt, Ui, W1);
    mmt(2, 2, 2, W1, Ui, W);
    float Wv[2] = {W[0] * v[0] + W[1] * v[1], W[2] * v[0] + W[3] * v[1]};
    x[0] = x[0] + Wv[0];
    x[1] = x[1] + Wv[1];
    float WW[4];
    mmt(2, 2, 2, W1, W1, WW);
    for (int j = 0; j < 4; j++) P[j] = P[j] - WW[j];
    Plin[0] = P[0]; Plin[1] = P[2]; Plin[2] = P[1]; Plin[3] = P[3];
}

/* core.cpp:294-317 with H = [0 0 1].  x[3], P row-major 3x3; in place. */
void so_joseph_update(float *x, float *P, float v, float R) {
    const float H[3] = {0, 0, 1};
    float PHt[3];
    for (int i = 0; i < 3; i++) PHt[i] = (P[i * 3] * H[0] + P[i * 3 + 1] * H[1]) + P[i * 3 + 2] * H[2];
    float S = (H[0] * PHt[0] + H[1] * PHt[1]) + H[2] * PHt[2];
    S = S + 1.0f * R;
    float Si = 1.0f * (1.0f / S);
    float W[3];
    for (int i = 0; i < 3; i++) W[i] = PHt[i] * Si;
    for (int i = 0; i < 3; i++) x[i] = x[i] + W[i] * v;
    float C[9], CP[9], CPC[9];
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) C[i * 3 + j] = ((i == j) ? 1.0f : 0.0f) - W[i] * H[j];
    mm(3, 3, 3, C, P, CP);
    mmt(3, 3, 3, CP, C, CPC);
    float eps = (float) (2.2204 * pow(10.0, -16));
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
            float wrw = (W[i] * R) * W[j];
            float eye = (i == j) ? 1.0f : 0.0f;
            P[i * 3 + j] = (CPC[i * 3 + j] + wrw) + eye * eps;
        }
}

/* core.cpp:452-458 with the n Gaussians already drawn (g = nRandMat::randn(n,1)). P row-major. */
void so_multivariate_gauss(int n, const float *x, const float *P, const float *g, float *out) {
    float A[9], L[9];
    memcpy(A, P, sizeof(float) * n * n);
    llt_lower(n, A);
    take_lower(n, A, L);
    for (int i = 0; i < n; i++) {
        float acc = L[i * n] * g[0];
        for (int t = 1; t < n; t++) acc = acc + L[i * n + t] * g[t];
        out[i] = acc + x[i];
    }
}

/* ------------------------------------------------------------------------- */
/* predict */
/* ------------------------------------------------------------------------- */
static void load_pose(long P, long p, const float *xv, const float *Pv, float *x, float *C) {
    for (int c = 0; c < 3; c++) x[c] = xv[c * P + p];
    if (C)
        for (int j = 0; j < 9; j++) C[j] = Pv[j * P + p];
}
static void store_pose(long P, long p, float *xv, float *Pv, const float *x, const float *C) {
    for (int c = 0; c < 3; c++) xv[c * P + p] = x[c];
    if (C)
        for (int j = 0; j < 9; j++) Pv[j * P + p] = C[j];
}

static void observe_heading(float *x, float *C, float phi, float sigma_phi) {
    /* fastslam2.cpp:113-125 == fastslam1.cpp:74-86 */
    float v = so_trig_offset(phi - x[2]);
    float R = (float) pow((double) sigma_phi, 2.0);
    so_joseph_update(x, C, v, R);
}

static void control_noise(float *V, float *G, const float *Q, const float *g2) {
    /* multivariateGauss(A, Q, 1), core.cpp:452-458 */
    float A[2] = {*V, *G}, o[2];
    so_multivariate_gauss(2, A, Q, g2, o);
    *V = o[0];
    *G = o[1];
}

static void motion(float *x, float V, float G, float dt, float wb) {
    /* fastslam2.cpp:101-104 == fastslam1.cpp:50-53 (note sin(G / wheelBase)) */
    float nx = x[0] + V * dt * cosf(G + x[2]);
    float ny = x[1] + V * dt * sinf(G + x[2]);
    float nt = so_trig_offset(x[2] + V * dt * sinf(G / wb));
    x[0] = nx; x[1] = ny; x[2] = nt;
}

void so_fs2_predict(long P, float *xv, float *Pv, float V0, float G0, const float *Q, float dt, float phi_true,
                    float wheel_base, float sigma_phi, int use_heading, const float *gauss) {
    for (long p = 0; p < P; p++) { /* fastslam2.cpp:51-60 */
        float x[3], C[9];
        float V = V0, G = G0;
        load_pose(P, p, xv, Pv, x, C);
        /* predictState, fastslam2.cpp:70-105 */
        float phi = x[2];
        float Gv[9] = {1, 0, -V * dt * sinf(G + phi), 0, 1, V * dt * cosf(G + phi), 0, 0, 1};
        float Gu[6] = {dt * cosf(G + phi), -V * dt * sinf(G + phi), dt * sinf(G + phi), V * dt * cosf(G + phi),
                       dt * sinf(G) / wheel_base, V * dt * cosf(G) / wheel_base};
        float T1[9], T2[9], T3[6], T4[9];
        mm(3, 3, 3, Gv, C, T1);
        mmt(3, 3, 3, T1, Gv, T2);
        mm(3, 2, 2, Gu, Q, T3);
        mmt(3, 2, 3, T3, Gu, T4);
        for (int j = 0; j < 9; j++) C[j] = T2[j] + T4[j];
        if (gauss) {
            float g2[2] = {gauss[p], gauss[P + p]};
            control_noise(&V, &G, Q, g2);
        }
        motion(x, V, G, dt, wheel_base);
        if (use_heading) observe_heading(x, C, phi_true, sigma_phi);
        store_pose(P, p, xv, Pv, x, C);
    }
}

void so_fs1_predict(long P, float *xv, float *Pv, float V0, float G0, const float *Q, float dt, float phi_true,
                    float wheel_base, float sigma_phi, int use_heading, const float *gauss) {
    for (long p = 0; p < P; p++) { /* fastslam1.cpp:57-66, 37-54 */
        float x[3], C[9];
        float V = V0, G = G0;
        load_pose(P, p, xv, Pv, x, C);
        if (gauss) {
            float g2[2] = {gauss[p], gauss[P + p]};
            control_noise(&V, &G, Q, g2);
        }
        motion(x, V, G, dt, wheel_base);
        if (use_heading) observe_heading(x, C, phi_true, sigma_phi);
        store_pose(P, p, xv, Pv, x, C);
    }
}

/* ------------------------------------------------------------------------- */
/* update, per particle */
/* ------------------------------------------------------------------------- */
static void lm_load(long P, long p, long l, const float *lm_mean, const float *lm_cov, float *xf, float *Pf) {
    const float *m = lm_mean + (l * P + p) * 2;
    const float *c = lm_cov + (l * P + p) * 4;
    xf[0] = m[0]; xf[1] = m[1];
    Pf[0] = c[0]; Pf[1] = c[2]; Pf[2] = c[1]; Pf[3] = c[3]; /* linear (col-major) -> row-major */
}

static void add_feature(long P, long p, int Nf, const float *x, float *lm_mean, float *lm_cov, const float *zn,
                        int Kn, const float *R) {
    for (int i = 0; i < Kn; i++) { /* core.cpp:479-509 */
        float r = zn[2 * i], b = zn[2 * i + 1];
        float s = sinf(x[2] + b), c = cosf(x[2] + b);
        float *m = lm_mean + ((long) (Nf + i) * P + p) * 2;
        float *cv = lm_cov + ((long) (Nf + i) * P + p) * 4;
        m[0] = x[0] + r * c;
        m[1] = x[1] + r * s;
        float Gz[4] = {c, -r * s, s, r * c};
        float T[4], Pf[4];
        mm(2, 2, 2, Gz, R, T);
        mmt(2, 2, 2, T, Gz, Pf);
        cv[0] = Pf[0]; cv[1] = Pf[2]; cv[2] = Pf[1]; cv[3] = Pf[3];
    }
}

static void feature_update(long P, long p, const float *x, float *lm_mean, float *lm_cov, const float *zf,
                           const int *idf, int K, const float *R) {
    for (int i = 0; i < K; i++) { /* core.cpp:132-175 */
        float xf[2], Pf[4], zp[2], Hf[4], Hv[6], Sf[4];
        lm_load(P, p, idf[i], lm_mean, lm_cov, xf, Pf);
        jacobian_one(x, R, xf, Pf, zp, Hf, Hv, Sf);
        float v[2] = {zf[2 * i] - zp[0], zf[2 * i + 1] - zp[1]};
        v[1] = so_trig_offset(v[1]);
        float *m = lm_mean + ((long) idf[i] * P + p) * 2;
        float *cv = lm_cov + ((long) idf[i] * P + p) * 4;
        float xx[2] = {m[0], m[1]};
        float Pl[4] = {cv[0], cv[1], cv[2], cv[3]};
        so_cholesky_update(xx, Pl, v, R, Hf);
        m[0] = xx[0]; m[1] = xx[1];
        for (int j = 0; j < 4; j++) cv[j] = Pl[j];
    }
}

static void fs2_sample_proposal(long P, long p, float *x, float *C, float *w, const float *lm_mean,
                                const float *lm_cov, const float *zf, const int *idf, int K, const float *R,
                                const float *g3) {
    /* fastslam2.cpp:290-368 */
    float x0[3] = {x[0], x[1], x[2]}, C0[9];
    memcpy(C0, C, sizeof(C0));
    for (int i = 0; i < K; i++) {
        float xf[2], Pf[4], zp[2], Hf[4], Hv[6], Sf[4], Sfi[4];
        lm_load(P, p, idf[i], lm_mean, lm_cov, xf, Pf);
        jacobian_one(x, R, xf, Pf, zp, Hf, Hv, Sf);
        inv2_lu(Sf, Sfi, 0);
        float vi[2] = {zf[2 * i] - zp[0], zf[2 * i + 1] - zp[1]};
        vi[1] = so_trig_offset(vi[1]);
        float Pinv[9];
        llt_inverse(3, C, Pinv);
        /* Pv = Hv^T * Sfi * Hv + Pv_inv */
        float HtS[6], HSH[9], A[9];
        for (int a = 0; a < 3; a++)
            for (int b = 0; b < 2; b++) HtS[a * 2 + b] = Hv[a] * Sfi[b] + Hv[3 + a] * Sfi[2 + b];
        mm(3, 2, 3, HtS, Hv, HSH);
        for (int j = 0; j < 9; j++) A[j] = HSH[j] + Pinv[j];
        llt_inverse(3, A, C);
        /* xv = xv + Pv * Hv^T * Sfi * vi */
        float PH[6], PHS[6];
        mmt(3, 3, 2, C, Hv, PH);
        mm(3, 2, 2, PH, Sfi, PHS);
        for (int a = 0; a < 3; a++) x[a] = x[a] + (PHS[a * 2] * vi[0] + PHS[a * 2 + 1] * vi[1]);
    }
    float xs[3];
    so_multivariate_gauss(3, x, C, g3, xs);
    float v1[3] = {x0[0] - xs[0], x0[1] - xs[1], so_trig_offset(x0[2] - xs[2])};
    float v2[3] = {x[0] - xs[0], x[1] - xs[1], so_trig_offset(x[2] - xs[2])};
    /* likelihoodGivenXv at the sampled pose, landmarks not yet updated: fastslam2.cpp:370-400 */
    float like = 1;
    for (int i = 0; i < K; i++) {
        float xf[2], Pf[4], zp[2], Hf[4], Hv[6], Sf[4];
        lm_load(P, p, idf[i], lm_mean, lm_cov, xf, Pf);
        jacobian_one(xs, R, xf, Pf, zp, Hf, Hv, Sf);
        float v[2] = {zf[2 * i] - zp[0], zf[2 * i + 1] - zp[1]};
        v[1] = so_trig_offset(v[1]);
        like = like * so_gauss_evaluate(v, Sf, 2);
    }
    float prior = so_gauss_evaluate(v1, C0, 3);
    float prop = so_gauss_evaluate(v2, C, 3);
    *w = *w * like * prior / prop;
    x[0] = xs[0]; x[1] = xs[1]; x[2] = xs[2];
    for (int j = 0; j < 9; j++) C[j] = 0;
}

int so_fs2_update_particles(long P, int Nf, float *xv, float *Pv, float *w, float *lm_mean, float *lm_cov,
                            const float *zf, const int *idf, int K, const float *zn, int Kn, const float *R,
                            const float *gauss) {
    for (long p = 0; p < P; p++) { /* fastslam2.cpp:26-45 */
        float x[3], C[9];
        load_pose(P, p, xv, Pv, x, C);
        float g3[3] = {gauss[p], gauss[P + p], gauss[2 * P + p]};
        if (K > 0) {
            fs2_sample_proposal(P, p, x, C, &w[p], lm_mean, lm_cov, zf, idf, K, R, g3);
            feature_update(P, p, x, lm_mean, lm_cov, zf, idf, K, R);
        }
        if (Kn > 0) {
            if (K == 0) {
                float xs[3];
                so_multivariate_gauss(3, x, C, g3, xs);
                x[0] = xs[0]; x[1] = xs[1]; x[2] = xs[2];
                for (int j = 0; j < 9; j++) C[j] = 0;
            }
            add_feature(P, p, Nf, x, lm_mean, lm_cov, zn, Kn, R);
        }
        store_pose(P, p, xv, Pv, x, C);
    }
    return Nf + Kn;
}

int so_fs1_update_particles(long P, int Nf, float *xv, float *Pv, float *w, float *lm_mean, float *lm_cov,
                            const float *zf, const int *idf, int K, const float *zn, int Kn, const float *R) {
    (void) Pv;
    for (long p = 0; p < P; p++) { /* fastslam1.cpp:18-35 */
        float x[3];
        load_pose(P, p, xv, 0, x, 0);
        if (K > 0) {
            float wt = 1.0f; /* computeWeight, fastslam1.cpp:91-118 */
            for (int i = 0; i < K; i++) {
                float xf[2], Pf[4], zp[2], Hf[4], Hv[6], S[4], Si[4], det;
                lm_load(P, p, idf[i], lm_mean, lm_cov, xf, Pf);
                jacobian_one(x, R, xf, Pf, zp, Hf, Hv, S);
                float v[2] = {zf[2 * i] - zp[0], zf[2 * i + 1] - zp[1]};
                v[1] = so_trig_offset(v[1]);
                inv2_lu(S, Si, &det);
                float den = (float) (2 * M_PI * sqrtf(det));
                float hv[2] = {-0.5f * v[0], -0.5f * v[1]};
                float t[2] = {hv[0] * Si[0] + hv[1] * Si[2], hv[0] * Si[1] + hv[1] * Si[3]};
                float e = t[0] * v[0] + t[1] * v[1];
                float num = expf(e);
                wt = wt * num / den;
            }
            w[p] = w[p] * wt;
            feature_update(P, p, x, lm_mean, lm_cov, zf, idf, K, R);
        }
        if (Kn > 0) add_feature(P, p, Nf, x, lm_mean, lm_cov, zn, Kn, R);
    }
    return Nf + Kn;
}

/* ------------------------------------------------------------------------- */
/* resampling */
/* ------------------------------------------------------------------------- */
long so_stratified_count(long N) { /* core.cpp:751-759: how many strata the reference's own loop emits */
    float k = (float) (1.0 / (float) N);
    float temp = k / 2;
    long c = 0;
    while (temp < (1 - k / 2)) {
        c++;
        temp = temp + k;
        if (c > N + 8) break;
    }
    return c;
}

void so_stratified_di(long N, float *di) { /* core.cpp:751-759 with the loop run exactly N times */
    float k = (float) (1.0 / (float) N);
    float temp = k / 2;
    for (long i = 0; i < N; i++) {
        di[i] = temp;
        temp = temp + k;
    }
}

/* Fixed-shape blocked inclusive scan ("scan_mode 1"): the order the CUDA path uses at scale.
 *   level 0: blocks of SO_SCAN_BLOCK consecutive elements; inside a block, 32-element groups are
 *            scanned by the Hillis-Steele ladder (d = 1,2,4,8,16: x[i] += x[i-d] for i >= d), the
 *            group totals are scanned by the same ladder, and element = group_exclusive + local;
 *   level 1: the block totals are scanned by this same function (recursively); element =
 *            block_exclusive + level-0 value.
 * Every addition is a single fp32 add, so any implementation with this shape gives the same bits. */
static void ladder32(float *x, int n) {
    for (int d = 1; d < 32; d <<= 1) {
        float t[32];
        for (int i = 0; i < n; i++) t[i] = (i >= d) ? x[i] + x[i - d] : x[i];
        memcpy(x, t, sizeof(float) * n);
    }
}

static void blocked_scan(float *x, long n) {
    if (n <= 0) return;
    long nb = (n + SO_SCAN_BLOCK - 1) / SO_SCAN_BLOCK;
    float *tot = (float *) malloc(sizeof(float) * nb);
    for (long b = 0; b < nb; b++) {
        float *xb = x + b * SO_SCAN_BLOCK;
        long len = n - b * SO_SCAN_BLOCK;
        if (len > SO_SCAN_BLOCK) len = SO_SCAN_BLOCK;
        int ng = (int) ((len + 31) / 32);
        float gt[32];
        for (int g = 0; g < 32; g++) gt[g] = 0.0f;
        for (int g = 0; g < ng; g++) {
            int gl = (int) ((len - g * 32 < 32) ? len - g * 32 : 32);
            /* lanes past the end hold 0 in the CUDA ladder; they never feed a live lane */
            ladder32(xb + g * 32, gl);
            gt[g] = xb[g * 32 + gl - 1];
        }
        ladder32(gt, 32);
        for (int g = 1; g < ng; g++) {
            int gl = (int) ((len - g * 32 < 32) ? len - g * 32 : 32);
            for (int i = 0; i < gl; i++) xb[g * 32 + i] = gt[g - 1] + xb[g * 32 + i];
        }
        tot[b] = xb[len - 1];
    }
    if (nb > 1) {
        blocked_scan(tot, nb);
        for (long b = 1; b < nb; b++) {
            float *xb = x + b * SO_SCAN_BLOCK;
            long len = n - b * SO_SCAN_BLOCK;
            if (len > SO_SCAN_BLOCK) len = SO_SCAN_BLOCK;
            for (long i = 0; i < len; i++) xb[i] = tot[b - 1] + xb[i];
        }
    }
    free(tot);
}

static float blocked_sum(const float *x, long n) { /* == last element of blocked_scan(x) */
    float *t = (float *) malloc(sizeof(float) * n);
    memcpy(t, x, sizeof(float) * n);
    blocked_scan(t, n);
    float r = t[n - 1];
    free(t);
    return r;
}

void so_resample_indices(long P, float *w, const double *unif, int scan_mode, int *keep, float *neff,
                         float *cdf_out) {
    /* resampleParticles core.cpp:718-729 then stratifiedResample core.cpp:780-807 */
    float ws = scan_mode ? blocked_sum(w, P) : so_eigen_sum(w, P);
    float *wn = (float *) malloc(sizeof(float) * P);
    float *sq = (float *) malloc(sizeof(float) * P);
    /* particles[i].w = w(i)/ws (core.cpp:727-729); stratifiedResample re-derives the same quotient from
     * the unnormalised copy with the same sum (core.cpp:782-786), so one division serves both. */
    for (long i = 0; i < P; i++) {
        wn[i] = w[i] / ws;
        sq[i] = (float) ((double) wn[i] * (double) wn[i]); /* pow(w,2) in double, stored float */
    }
    float s2 = scan_mode ? blocked_sum(sq, P) : so_eigen_sum(sq, P);
    *neff = 1 / s2;
    for (long i = 0; i < P; i++) keep[i] = -1;
    /* stratifiedRandom, core.cpp:751-769 */
    float k = (float) (1.0 / (float) P);
    float *sel = (float *) malloc(sizeof(float) * P);
    so_stratified_di(P, sel);
    for (long i = 0; i < P; i++) sel[i] = (float) ((double) sel[i] + unif[i] * (double) k - (double) (k / 2));
    /* cumulativeSum, core.cpp:813-824 (each prefix re-summed from zero == running fp32 sum) */
    float *cdf = (float *) malloc(sizeof(float) * P);
    memcpy(cdf, wn, sizeof(float) * P);
    if (scan_mode) {
        blocked_scan(cdf, P);
    } else {
        float s = 0;
        for (long i = 0; i < P; i++) {
            s += wn[i];
            cdf[i] = s;
        }
    }
    long ctr = 0; /* core.cpp:800-806 */
    for (long i = 0; i < P; i++)
        while ((ctr < P) && (sel[ctr] < cdf[i])) {
            keep[ctr] = (int) i;
            ctr++;
        }
    memcpy(w, wn, sizeof(float) * P);
    if (cdf_out) memcpy(cdf_out, cdf, sizeof(float) * P);
    free(wn); free(sq); free(sel); free(cdf);
}

void so_resample_gather(long P, int Nf, const int *keep, float *xv, float *Pv, float *w, float *lm_mean,
                        float *lm_cov) {
    /* core.cpp:736-748: particles[i] = old[keep[i]]; w = 1/N.  keep<0 is UB in the reference; clamp. */
    float *tmp = (float *) malloc(sizeof(float) * P * 4);
    for (int c = 0; c < 3; c++) {
        for (long p = 0; p < P; p++) tmp[p] = xv[c * P + ((keep[p] < 0) ? P - 1 : keep[p])];
        memcpy(xv + c * P, tmp, sizeof(float) * P);
    }
    for (int c = 0; c < 9; c++) {
        for (long p = 0; p < P; p++) tmp[p] = Pv[c * P + ((keep[p] < 0) ? P - 1 : keep[p])];
        memcpy(Pv + c * P, tmp, sizeof(float) * P);
    }
    for (long l = 0; l < Nf; l++) {
        float *m = lm_mean + l * P * 2, *cv = lm_cov + l * P * 4;
        for (long p = 0; p < P; p++) {
            long s = (keep[p] < 0) ? P - 1 : keep[p];
            tmp[2 * p] = m[2 * s]; tmp[2 * p + 1] = m[2 * s + 1];
        }
        memcpy(m, tmp, sizeof(float) * P * 2);
        for (long p = 0; p < P; p++) {
            long s = (keep[p] < 0) ? P - 1 : keep[p];
            for (int j = 0; j < 4; j++) tmp[4 * p + j] = cv[4 * s + j];
        }
        memcpy(cv, tmp, sizeof(float) * P * 4);
    }
    float nw = 1.0f / (float) P;
    for (long p = 0; p < P; p++) w[p] = nw;
    free(tmp);
}

void so_estimate(long P, const float *xv, const float *w, double *x, double *y, double *t) {
    /* ParticleSLAMWrapper.cpp:56-77: double means of x,y; heading of the first max-weight particle */
    double sx = 0, sy = 0;
    float wmax = -1;
    long imax = 0;
    for (long p = 0; p < P; p++) {
        sx += xv[p];
        sy += xv[P + p];
        if (w[p] > wmax) {
            wmax = w[p];
            imax = p;
        }
    }
    *x = sx / P;
    *y = sy / P;
    *t = xv[2 * P + imax];
}
