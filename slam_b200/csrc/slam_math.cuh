// slam_math.cuh — per-particle FastSLAM arithmetic as inline device functions.
//
// Every function cites the reference lines (relative to /root/reference/) whose result it reproduces.  All
// state is fp32 like the reference's VectorXf/MatrixXf.  The reference's few double promotions are replaced by
// fp32 forms that round identically or within one ulp (noted per function); the 1e-4 norm-relative contract of
// BASELINE.json is checked against the CPU oracle in tests/.
//
// Matrices are row-major in registers.  The functions are __host__ __device__ only so that tests/host_emul can
// compile this same arithmetic on the build container (no GPU there) and diff it against the oracle before a
// GPU run is spent; the product library exports no host compute path.
#pragma once

#include <math.h>

#if defined(__CUDACC__)
#define SLAM_HD __host__ __device__ __forceinline__
#else
#define SLAM_HD inline
#endif
// single fp32 operations that must not be contracted into an FMA
#if defined(__CUDA_ARCH__)
#define SLAM_FADD(a, b) __fadd_rn((a), (b))
#define SLAM_FMUL(a, b) __fmul_rn((a), (b))
#else
#define SLAM_FADD(a, b) ((a) + (b))
#define SLAM_FMUL(a, b) ((a) * (b))
#endif

namespace slam {

#define SLAM_PI_F 3.14159274f          // float(M_PI); the largest float below it is < M_PI
#define SLAM_TWO_PI_HI 6.28318548f     // float(2*M_PI)
#define SLAM_TWO_PI_LO 1.74845553e-7f  // float(2*M_PI) - 2*M_PI
#define SLAM_TWO_PI_D 6.283185307179586

// trigonometricOffset, core.cpp:460-477.  The reference compares a float against the double M_PI and subtracts the
// double 2*M_PI; "ang > M_PI" is "ang >= float(M_PI)" for a float ang, and (ang - hi) + lo is the correctly
// rounded double result whenever ang is within [pi, 2pi] (Sterbenz), which the first branch guarantees.
SLAM_HD float trig_offset(float ang) {
    if (ang <= -SLAM_TWO_PI_HI || ang >= SLAM_TWO_PI_HI) {  // rare: keep the reference's double arithmetic
        int n = (int) floor((double) ang / SLAM_TWO_PI_D);
        ang = (float) ((double) ang - n * SLAM_TWO_PI_D);
    }
    if (ang >= SLAM_PI_F) ang = (ang - SLAM_TWO_PI_HI) + SLAM_TWO_PI_LO;
    if (ang <= -SLAM_PI_F) ang = (ang + SLAM_TWO_PI_HI) - SLAM_TWO_PI_LO;
    return ang;
}

// C[m x n] = A[m x k] * B[k x n], k-sequential accumulation (Eigen's small-product order).
template <int M, int K, int N>
SLAM_HD void mm(const float *A, const float *B, float *C) {
#pragma unroll
    for (int i = 0; i < M; i++)
#pragma unroll
        for (int j = 0; j < N; j++) {
            float acc = A[i * K] * B[j];
#pragma unroll
            for (int t = 1; t < K; t++) acc = acc + A[i * K + t] * B[t * N + j];
            C[i * N + j] = acc;
        }
}
// C[m x n] = A[m x k] * B^T, B is [n x k].
template <int M, int K, int N>
SLAM_HD void mmt(const float *A, const float *B, float *C) {
#pragma unroll
    for (int i = 0; i < M; i++)
#pragma unroll
        for (int j = 0; j < N; j++) {
            float acc = A[i * K] * B[j * K];
#pragma unroll
            for (int t = 1; t < K; t++) acc = acc + A[i * K + t] * B[j * K + t];
            C[i * N + j] = acc;
        }
}

// Eigen's unblocked lower LLT (libs/eigen3/Eigen/src/Cholesky/LLT.h:263-287) in place on the lower triangle.
// Returns -1 on success, else the column where a non-positive pivot stopped it; like Eigen, the rest of the
// matrix is then left untouched and callers go on using it (the reference never checks info()).
template <int N>
SLAM_HD int llt_lower(float *A) {
#pragma unroll
    for (int k = 0; k < N; k++) {
        float x = A[k * N + k];
        if (k > 0) {
            float sq = A[k * N] * A[k * N];
#pragma unroll
            for (int t = 1; t < k; t++) sq = sq + A[k * N + t] * A[k * N + t];
            x = x - sq;
        }
        if (x <= 0.0f) return k;
        x = sqrtf(x);
        A[k * N + k] = x;
        if (k > 0) {
#pragma unroll
            for (int i = k + 1; i < N; i++)
#pragma unroll
                for (int t = 0; t < k; t++) A[i * N + k] = A[i * N + k] - A[i * N + t] * A[k * N + t];
        }
        if (k + 1 < N) {
            float r = 1.0f / x;
#pragma unroll
            for (int i = k + 1; i < N; i++) A[i * N + k] = A[i * N + k] * r;
        }
    }
    return -1;
}

// A.llt().solve(Identity) (LLT.h:448-454): forward substitution in axpy form, backward in dot form, both
// multiplying by the reciprocal of the diagonal (Core/products/TriangularSolverMatrix.h:104-134).
template <int N>
SLAM_HD void llt_inverse(const float *Ain, float *X) {
    float A[N * N];
#pragma unroll
    for (int j = 0; j < N * N; j++) A[j] = Ain[j];
    llt_lower<N>(A);
#pragma unroll
    for (int i = 0; i < N; i++)
#pragma unroll
        for (int j = 0; j < N; j++) X[i * N + j] = (i == j) ? 1.0f : 0.0f;
#pragma unroll
    for (int k = 0; k < N; k++) {
        float a = 1.0f / A[k * N + k];
#pragma unroll
        for (int j = 0; j < N; j++) {
            float b = X[k * N + j] * a;
            X[k * N + j] = b;
#pragma unroll
            for (int i = k + 1; i < N; i++) X[i * N + j] = X[i * N + j] - b * A[i * N + k];
        }
    }
#pragma unroll
    for (int k = 0; k < N; k++) {
        const int i = N - k - 1;
        float a = 1.0f / A[i * N + i];
#pragma unroll
        for (int j = 0; j < N; j++) {
            float b = 0.0f;
#pragma unroll
            for (int t = 0; t < k; t++) b = b + A[(i + 1 + t) * N + i] * X[(i + 1 + t) * N + j];
            X[i * N + j] = (X[i * N + j] - b) * a;
        }
    }
}

// MatrixXf::inverse() of a dynamic 2x2 = partialPivLu().solve(I) (LU/Inverse.h, PartialPivLU.h): row pivot on
// |a| vs |c|, unit-lower then upper solve in axpy form.  det (nullable) = determinant() of the same LU.
SLAM_HD void inv2_lu(const float *A, float *X, float *det) {
    float a = A[0], b = A[1], c = A[2], d = A[3];
    const bool swap = fabsf(c) > fabsf(a);
    if (swap) {
        float t = a; a = c; c = t;
        t = b; b = d; d = t;
    }
    const float l = c / a;
    const float u11 = d - l * b;
    float B0 = swap ? 0.0f : 1.0f, B1 = swap ? 1.0f : 0.0f, B2 = B1, B3 = B0;
    B2 = B2 - B0 * l;
    B3 = B3 - B1 * l;
    const float a1 = 1.0f / u11;
    B2 = B2 * a1;
    B0 = B0 - B2 * b;
    B3 = B3 * a1;
    B1 = B1 - B3 * b;
    const float a0 = 1.0f / a;
    X[0] = B0 * a0;
    X[1] = B1 * a0;
    X[2] = B2;
    X[3] = B3;
    if (det) *det = (swap ? -1.0f : 1.0f) * (a * u11);
}

// computeJacobians for one landmark, core.cpp:666-713.  R, Pf, Hf, Sf row-major.  Hv is implied:
// Hv = [-Hf0 -Hf1 0; -Hf2 -Hf3 -1].  d2 is formed in double like the reference's pow(dx,2)+pow(dy,2) (core.cpp:685):
// the innovation z - zp cancels ~3 digits, so an ulp of zp is worth ~1e-5 of likelihood and is not given away.
SLAM_HD void jacobian_one(const float *xv, const float *R, const float *xf, const float *Pf, float *zp, float *Hf,
                          float *Sf) {
    const float dx = xf[0] - xv[0];
    const float dy = xf[1] - xv[1];
    const float d2 = (float) ((double) dx * (double) dx + (double) dy * (double) dy);
    const float d = sqrtf(d2);
    zp[0] = d;
    zp[1] = trig_offset(atan2f(dy, dx) - xv[2]);
    Hf[0] = dx / d;
    Hf[1] = dy / d;
    Hf[2] = -dy / d2;
    Hf[3] = dx / d2;
    float T[4];
    mm<2, 2, 2>(Hf, Pf, T);
    mmt<2, 2, 2>(T, Hf, Sf);
#pragma unroll
    for (int j = 0; j < 4; j++) Sf[j] = Sf[j] + R[j];
}

SLAM_HD void hv_from_hf(const float *Hf, float *Hv) {
    Hv[0] = -Hf[0]; Hv[1] = -Hf[1]; Hv[2] = 0.0f;
    Hv[3] = -Hf[2]; Hv[4] = -Hf[3]; Hv[5] = -1.0f;
}

// FastSLAM2::gaussEvaluate with logflag = 0, fastslam2.cpp:127-163: exp(-0.5*|L^-1 v|^2) / ((2pi)^(D/2) * prod diag L)
// with the reference's INTEGER D/2 (so (2pi)^1 for D = 2 and for D = 3).  The reference solves L*nin = v through
// a Jacobi SVD; this is the forward substitution it equals.
template <int D>
SLAM_HD float gauss_evaluate(const float *v, const float *S) {
    float A[D * D], nin[D];
#pragma unroll
    for (int j = 0; j < D * D; j++) A[j] = S[j];
    llt_lower<D>(A);
#pragma unroll
    for (int i = 0; i < D; i++) {
        float b = v[i];
#pragma unroll
        for (int t = 0; t < i; t++) b = b - A[i * D + t] * nin[t];
        nin[i] = b / A[i * D + i];
    }
    float E = 0.0f;  // each square rounded to float before it is added (pow() in double, stored float)
#pragma unroll
    for (int s = 0; s < D; s++) E = SLAM_FADD(E, SLAM_FMUL(nin[s], nin[s]));
    E = -0.5f * E;
    float prod = 1.0f;
#pragma unroll
    for (int i = 0; i < D; i++) prod = prod * A[i * D + i];
    const float C = SLAM_TWO_PI_HI * prod;  // pow(2*M_PI, D/2) with D/2 == 1, rounded to float after the product
    return expf(E) / C;
}

// choleskyUpdate, core.cpp:275-291 (2x2 landmark EKF).  x[2]; P row-major 2x2 in/out (may be asymmetric: the
// reference symmetrises S only); v innovation; R, H row-major.
SLAM_HD void cholesky_update(float *x, float *P, const float *v, const float *R, const float *H) {
    float PHt[4], S[4];
    mmt<2, 2, 2>(P, H, PHt);
    mm<2, 2, 2>(H, PHt, S);
#pragma unroll
    for (int j = 0; j < 4; j++) S[j] = S[j] + R[j];
    float Ss[4] = {(S[0] + S[0]) * 0.5f, (S[1] + S[2]) * 0.5f, (S[2] + S[1]) * 0.5f, (S[3] + S[3]) * 0.5f};
    llt_lower<2>(Ss);
    const float U[4] = {Ss[0], Ss[2], 0.0f, Ss[3]};  // matrixU() = adjoint of the lower factor
    float Ui[4];
    inv2_lu(U, Ui, nullptr);
    float W1[4], W[4], WW[4];
    mm<2, 2, 2>(PHt, Ui, W1);
    mmt<2, 2, 2>(W1, Ui, W);
    x[0] = x[0] + (W[0] * v[0] + W[1] * v[1]);
    x[1] = x[1] + (W[2] * v[0] + W[3] * v[1]);
    mmt<2, 2, 2>(W1, W1, WW);
#pragma unroll
    for (int j = 0; j < 4; j++) P[j] = P[j] - WW[j];
}

// josephUpdate with H = [0 0 1], core.cpp:294-317 (heading observation).  The reference's 2.2204e-16*I term is
// kept (it is what makes Pv positive definite after the first tick).
SLAM_HD void joseph_update_heading(float *x, float *P, float v, float R) {
    const float PHt[3] = {P[2], P[5], P[8]};
    const float S = PHt[2] + R;
    const float Si = 1.0f / S;
    float W[3];
#pragma unroll
    for (int i = 0; i < 3; i++) W[i] = PHt[i] * Si;
#pragma unroll
    for (int i = 0; i < 3; i++) x[i] = x[i] + W[i] * v;
    // C = I - W*H differs from I only in its last column
    float C[9] = {1.0f, 0.0f, 0.0f - W[0], 0.0f, 1.0f, 0.0f - W[1], 0.0f, 0.0f, 1.0f - W[2]};
    float CP[9], CPC[9];
    mm<3, 3, 3>(C, P, CP);
    mmt<3, 3, 3>(CP, C, CPC);
    const float eps = 2.2204e-16f;
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 3; j++) {
            const float wrw = (W[i] * R) * W[j];
            P[i * 3 + j] = (CPC[i * 3 + j] + wrw) + ((i == j) ? eps : 0.0f);
        }
}

// multivariateGauss with the Gaussians already drawn, core.cpp:452-458: chol_L(P)*g + x.
template <int N>
SLAM_HD void multivariate_gauss(const float *x, const float *P, const float *g, float *out) {
    float A[N * N];
#pragma unroll
    for (int j = 0; j < N * N; j++) A[j] = P[j];
    llt_lower<N>(A);
#pragma unroll
    for (int i = 0; i < N; i++) {
        // matrixL(): entries above the diagonal read as zero
        float acc = A[i * N] * g[0];
#pragma unroll
        for (int t = 1; t < N; t++) acc = acc + ((t <= i) ? A[i * N + t] : 0.0f) * g[t];
        out[i] = acc + x[i];
    }
}

// Motion model shared by FastSLAM1/2::predictState (fastslam2.cpp:101-104, fastslam1.cpp:50-53); note the
// reference's sin(G / wheelBase).
SLAM_HD void motion(float *x, float V, float G, float dt, float wb) {
    const float nx = x[0] + V * dt * cosf(G + x[2]);
    const float ny = x[1] + V * dt * sinf(G + x[2]);
    const float nt = trig_offset(x[2] + V * dt * sinf(G / wb));
    x[0] = nx; x[1] = ny; x[2] = nt;
}

// FastSLAM2::predictState covariance part, fastslam2.cpp:70-99: Pv = Gv*Pv*Gv' + Gu*Q*Gu'.
SLAM_HD void predict_cov(float *C, float phi, float V, float G, float dt, float wb, const float *Q) {
    const float s = sinf(G + phi), c = cosf(G + phi);
    const float Gv[9] = {1.0f, 0.0f, -V * dt * s, 0.0f, 1.0f, V * dt * c, 0.0f, 0.0f, 1.0f};
    const float Gu[6] = {dt * c, -V * dt * s, dt * s, V * dt * c, dt * sinf(G) / wb, V * dt * cosf(G) / wb};
    float T1[9], T2[9], T3[6], T4[9];
    mm<3, 3, 3>(Gv, C, T1);
    mmt<3, 3, 3>(T1, Gv, T2);
    mm<3, 2, 2>(Gu, Q, T3);
    mmt<3, 2, 3>(T3, Gu, T4);
#pragma unroll
    for (int j = 0; j < 9; j++) C[j] = T2[j] + T4[j];
}

// observeHeading, fastslam2.cpp:113-125 == fastslam1.cpp:74-86.
SLAM_HD void observe_heading(float *x, float *C, float phi_true, float sigma_phi) {
    const float v = trig_offset(phi_true - x[2]);
    joseph_update_heading(x, C, v, sigma_phi * sigma_phi);
}

// One full predict for one particle.  g2 = nullptr (no control noise) or two standard normals.
SLAM_HD void predict_particle(int method, float *x, float *C, float V, float G, const float *Q, float dt,
                              float phi_true, float wb, float sigma_phi, bool use_heading, const float *g2) {
    if (method == 2) predict_cov(C, x[2], V, G, dt, wb, Q);
    if (g2) {  // multivariateGauss([V;G], Q, 1), fastslam2.cpp:88-97 / fastslam1.cpp:39-48
        const float A[2] = {V, G};
        float o[2];
        multivariate_gauss<2>(A, Q, g2, o);
        V = o[0];
        G = o[1];
    }
    motion(x, V, G, dt, wb);
    if (use_heading) observe_heading(x, C, phi_true, sigma_phi);
}

// addFeature for one new observation, core.cpp:479-509.  Pf out row-major.
SLAM_HD void add_feature_one(const float *x, float r, float b, const float *R, float *xf, float *Pf) {
    const float s = sinf(x[2] + b), c = cosf(x[2] + b);
    xf[0] = x[0] + r * c;
    xf[1] = x[1] + r * s;
    const float Gz[4] = {c, -r * s, s, r * c};
    float T[4];
    mm<2, 2, 2>(Gz, R, T);
    mmt<2, 2, 2>(T, Gz, Pf);
}

// One iteration of FastSLAM2::sampleProposal's loop, fastslam2.cpp:314-350: refine (x, C) with one observation.
SLAM_HD void proposal_step(float *x, float *C, const float *R, const float *xf, const float *Pf, const float *z) {
    float zp[2], Hf[4], Hv[6], Sf[4], Sfi[4];
    jacobian_one(x, R, xf, Pf, zp, Hf, Sf);
    hv_from_hf(Hf, Hv);
    inv2_lu(Sf, Sfi, nullptr);
    const float vi[2] = {z[0] - zp[0], trig_offset(z[1] - zp[1])};
    float Pinv[9];
    llt_inverse<3>(C, Pinv);
    float HtS[6], HSH[9], A[9];
#pragma unroll
    for (int a = 0; a < 3; a++)
#pragma unroll
        for (int b = 0; b < 2; b++) HtS[a * 2 + b] = Hv[a] * Sfi[b] + Hv[3 + a] * Sfi[2 + b];
    mm<3, 2, 3>(HtS, Hv, HSH);
#pragma unroll
    for (int j = 0; j < 9; j++) A[j] = HSH[j] + Pinv[j];
    llt_inverse<3>(A, C);
    float PH[6], PHS[6];
    mmt<3, 3, 2>(C, Hv, PH);
    mm<3, 2, 2>(PH, Sfi, PHS);
#pragma unroll
    for (int a = 0; a < 3; a++) x[a] = x[a] + (PHS[a * 2] * vi[0] + PHS[a * 2 + 1] * vi[1]);
}

// Second half, per observation, at the sampled pose: likelihoodGivenXv's factor (fastslam2.cpp:370-400) and
// featureUpdate's landmark EKF (core.cpp:132-175).  Both evaluate computeJacobians at the same pose on the same
// not-yet-updated landmark, so one evaluation serves both (idf rows are distinct within an update).
SLAM_HD float likelihood_and_feature_update(const float *xs, const float *R, float *xf, float *Pf, const float *z) {
    float zp[2], Hf[4], Sf[4];
    jacobian_one(xs, R, xf, Pf, zp, Hf, Sf);
    const float v[2] = {z[0] - zp[0], trig_offset(z[1] - zp[1])};
    const float like = gauss_evaluate<2>(v, Sf);
    cholesky_update(xf, Pf, v, R, Hf);
    return like;
}

// FastSLAM1::computeWeight's factor for one observation (fastslam1.cpp:91-118), folded into *wt in the
// reference's order (wt*num/den), followed by the landmark EKF (core.cpp:132-175).
SLAM_HD void fs1_weight_and_feature_update(const float *x, const float *R, float *xf, float *Pf, const float *z,
                                           float *wt) {
    float zp[2], Hf[4], S[4], Si[4], det;
    jacobian_one(x, R, xf, Pf, zp, Hf, S);
    const float v[2] = {z[0] - zp[0], trig_offset(z[1] - zp[1])};
    inv2_lu(S, Si, &det);
    const float den = SLAM_TWO_PI_HI * sqrtf(det);  // 2*M_PI*sqrt(det) in double, cast to float
    const float hv0 = -0.5f * v[0], hv1 = -0.5f * v[1];
    const float t0 = hv0 * Si[0] + hv1 * Si[2], t1 = hv0 * Si[1] + hv1 * Si[3];
    const float e = t0 * v[0] + t1 * v[1];
    *wt = *wt * expf(e) / den;
    cholesky_update(xf, Pf, v, R, Hf);
}

}  // namespace slam
