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


// =============================================================================================================
// FAST formulation (SLAMGPU_MATH_FAST): the same estimator with the arithmetic reorganised for instruction count.
//
// The reference-order functions above cost ~1000 SASS instructions per particle-landmark (ncu, profiles/): IEEE
// divisions, a Jacobi-free but still explicit 3x3 inverse TWICE per observation, an LU inverse of S, separate
// Cholesky factorisations for the likelihood and for the landmark EKF.  The forms below are algebraically identical
// and keep every reference quirk that changes results (integer D/2 in the Gaussian normaliser, sin(G/wheelBase), Pv
// zeroed after sampling, S symmetrised but P not), but:
//   * sampleProposal runs in INFORMATION form: Lambda = Pv^-1 is carried across observations (the reference inverts Pv,
//     adds Hv' Sf^-1 Hv and inverts back: inv(inv(A)) == A), the pose increment is one Cholesky SOLVE, and Pv itself
//     is formed once per particle after the last observation (it is only needed for the pose sample and the proposal
//     density);
//   * Sf^-1 is the 2x2 adjugate / determinant; Cholesky factors use rsqrt (no division, no sqrt);
//   * likelihoodGivenXv's factor and choleskyUpdate share ONE factorisation of the symmetrised S, and the landmark
//     mean update reuses the whitened innovation (W*v == W1 * (L^-1 v)).
// Differences from the reference are rounding-level (a few ulp per operation); tests bound them against the oracle
// with the same tolerances as the reference-order path.
// =============================================================================================================
// MUFU approximations issued directly (rsqrt/rcp/ex2 .approx.ftz): rsqrtf(), __fdividef() and __expf() without
// -ftz wrap the same MUFU instruction in a denormal-range test and two rescaling multiplies (3-5 extra instructions
// each, 11 such calls per particle-landmark: ncu r01, profiles/).  Denormal inputs cannot occur for rsqrt / rcp here
// (ranges, variances and determinants of metre-scale quantities); ex2 is the exception, see mufu_ex2.
#if defined(__CUDA_ARCH__)
__device__ __forceinline__ float mufu_rsqrt(float x) {
    float r;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ float mufu_rcp(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
// NOT flushed to zero: a likelihood below 2^-126 stays a denormal like the reference's expf result.  That matters
// exactly when a filter is in trouble — every particle far from the observations: with a flush all weights become 0,
// their sum 0, the normalised weights NaN and nEff NaN, so nothing ever resamples again, whereas the reference keeps
// a denormal weight on the least bad particle, resamples onto it and recovers (seen as two lost tracks in 37 seeded
// runs of BASELINE.json config 0 with the flushing version, none in 41 without).  Costs two scaling instructions.
__device__ __forceinline__ float mufu_ex2(float x) {
    float r;
    asm("ex2.approx.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
#define SLAM_RSQRT(x) mufu_rsqrt(x)
#define SLAM_RCP(x) mufu_rcp(x)
#define SLAM_EXP2(x) mufu_ex2(x)
#define SLAM_SINCOS_APPROX(a, s, c) __sincosf((a), (s), (c))
#else
#define SLAM_SINCOS_APPROX(a, s, c) (*(s) = sinf(a), *(c) = cosf(a))
#define SLAM_RSQRT(x) (1.0f / sqrtf(x))
#define SLAM_RCP(x) (1.0f / (x))
#define SLAM_EXP2(x) exp2f(x)
#endif
#define SLAM_NEG_HALF_LOG2E (-0.72134752f)  // exp(-q/2) == exp2(q * -0.5*log2(e))

#define SLAM_INV_TWO_PI 0.159154943f

// atan2 for the bearing prediction: octant reduction to a = min/max in [0,1], degree-8 minimax polynomial in a^2
// (max |error| 1.1e-7 rad in fp32, about what atan2f's 2 ulp is worth at |angle| ~ pi/2..pi), ~24 instructions
// instead of ~60 for atan2f + the reference's two-branch wrap.
SLAM_HD float atan2_fast(float y, float x) {
    const float ax = fabsf(x), ay = fabsf(y);
    const float mx = fmaxf(ax, ay), mn = fminf(ax, ay);
    const float a = mn * SLAM_RCP(mx);
    const float q = a * a;
    float r = 0.0028340641874819994f;
    r = fmaf(r, q, -0.016005029901862144f);
    r = fmaf(r, q, 0.042587608098983765f);
    r = fmaf(r, q, -0.07495445758104324f);
    r = fmaf(r, q, 0.10636754333972931f);
    r = fmaf(r, q, -0.14202570915222168f);
    r = fmaf(r, q, 0.19992484152317047f);
    r = fmaf(r, q, -0.3333306610584259f);
    r = fmaf(r * q, a, a);
    if (ay > ax) r = 1.57079637f - r;
    if (x < 0.0f) r = 3.14159274f - r;
    return copysignf(r, y);
}

// Wrap an angle difference into [-pi, pi] (trigonometricOffset's job, core.cpp:460-477) by one rounding to the
// nearest multiple of 2*pi; identical to the reference whenever no wrap is needed.
SLAM_HD float wrap_fast(float t) {
    const float k = rintf(t * SLAM_INV_TWO_PI);
    return fmaf(k, SLAM_TWO_PI_LO, fmaf(-k, SLAM_TWO_PI_HI, t));
}

// The bearing innovation v1 = wrap(z_b - (atan2(dy, dx) - theta)) (core.cpp:686 + trigonometricOffset) WITHOUT an
// arctangent per landmark.  With phi = z_b + theta (the observed ray in the world frame) and u = (dx, dy) / d the unit
// vector to the landmark — which the Jacobian needs anyway — the innovation is the angle from u to the ray:
//     sin v1 = sin(phi) ux - cos(phi) uy,    cos v1 = cos(phi) ux + sin(phi) uy.
// An innovation is small when the filter works, so v1 = asin(sin v1) by an odd polynomial for |sin v1| <= 0.5 (30
// degrees: max error 6e-8 rad; wide enough that a warp almost never has a lane outside it, which would make the whole
// warp execute both branches); anything larger — and the far side, cos v1 <= 0, where the reference's wrap matters —
// takes the arctangent, which returns the wrapped angle directly.  ~16 instructions instead of the ~29 of atan2 +
// subtract + wrap, twice per particle-landmark.
SLAM_HD float bearing_innovation(float sphi, float cphi, float ux, float uy) {
    const float sn = fmaf(sphi, ux, -cphi * uy);
    const float cs = fmaf(cphi, ux, sphi * uy);
    if (fabsf(sn) <= 0.5f && cs > 0.0f) {
        // asin(s) = s + s^3 g(s^2), g a cubic fitted (minimax) on s^2 in [0, 0.25]: max |error| 6e-8 rad in fp32
        const float q = sn * sn;
        float r = 0.05159852f;
        r = fmaf(r, q, 0.03918817f);
        r = fmaf(r, q, 0.07554101f);
        r = fmaf(r, q, 0.16664924f);
        return fmaf(r * q, sn, sn);
    }
    return atan2_fast(sn, cs);
}

// computeJacobians (core.cpp:666-713) for one landmark: predicted range *dist, the unit vector to the landmark in
// Hf[0..1] (first row of Hf), Hf, T = Hf * sym(Pf) (row-major; its transpose is the PHt of choleskyUpdate) and the
// symmetrised innovation covariance S = Hf Pf Hf' + R as (s00, s10, s11).  The predicted bearing itself is never
// formed: see bearing_innovation().
// (S + S')/2 == Hf sym(Pf) Hf' + sym(R) exactly, so the stored Pf (row-major, off-diagonals possibly an ulp apart —
// the reference never symmetrises P, core.cpp:290) enters through pm = (Pf01 + Pf10)/2 and S needs 3 entries, not 4.
SLAM_HD void jacobian_fast(const float *xv, const float *R, const float *xf, const float *Pf, float *dist, float *Hf,
                           float *T, float *S3) {
    const float dx = xf[0] - xv[0];
    const float dy = xf[1] - xv[1];
    const float d2 = fmaf(dx, dx, dy * dy);
    const float rd = SLAM_RSQRT(d2);
    float d = d2 * rd;
    d = fmaf(0.5f * rd, fmaf(-d, d, d2), d);  // one Newton step: the range innovation cancels ~3 digits
    const float rd2 = rd * rd;
    *dist = d;
    Hf[0] = dx * rd;
    Hf[1] = dy * rd;
    Hf[2] = -dy * rd2;
    Hf[3] = dx * rd2;
    const float pm = 0.5f * (Pf[1] + Pf[2]), rm = 0.5f * (R[1] + R[2]);
    T[0] = fmaf(Hf[0], Pf[0], Hf[1] * pm);
    T[1] = fmaf(Hf[0], pm, Hf[1] * Pf[3]);
    T[2] = fmaf(Hf[2], Pf[0], Hf[3] * pm);
    T[3] = fmaf(Hf[2], pm, Hf[3] * Pf[3]);
    S3[0] = fmaf(T[0], Hf[0], fmaf(T[1], Hf[1], R[0]));
    S3[1] = fmaf(T[0], Hf[2], fmaf(T[1], Hf[3], rm));
    S3[2] = fmaf(T[2], Hf[2], fmaf(T[3], Hf[3], R[3]));
}

// Lower Cholesky factor of a symmetric 3x3 given as (a00, a10, a11, a20, a21, a22), in reciprocal-diagonal form:
// L = {l00i, l10, l11i, l20, l21, l22i} with lXXi = 1 / L[X][X].
SLAM_HD void chol3_rs(const float *a, float *L) {
    L[0] = SLAM_RSQRT(a[0]);
    L[1] = a[1] * L[0];
    L[3] = a[3] * L[0];
    L[2] = SLAM_RSQRT(fmaf(-L[1], L[1], a[2]));
    L[4] = fmaf(-L[3], L[1], a[4]) * L[2];
    L[5] = SLAM_RSQRT(fmaf(-L[4], L[4], fmaf(-L[3], L[3], a[5])));
}
// y = L^-1 b
SLAM_HD void chol3_fwd(const float *L, const float *b, float *y) {
    y[0] = b[0] * L[0];
    y[1] = fmaf(-L[1], y[0], b[1]) * L[2];
    y[2] = fmaf(-L[4], y[1], fmaf(-L[3], y[0], b[2])) * L[5];
}
// x = L^-T y
SLAM_HD void chol3_bwd(const float *L, const float *y, float *x) {
    x[2] = y[2] * L[5];
    x[1] = fmaf(-L[4], x[2], y[1]) * L[2];
    x[0] = fmaf(-L[3], x[2], fmaf(-L[1], x[1], y[0])) * L[0];
}
// (L L')^-1 as packed symmetric (c00, c10, c11, c20, c21, c22)
SLAM_HD void chol3_inverse(const float *L, float *c) {
    // M = L^-1 (lower): m00 = l00i, m10 = -l10 m00 l11i, m11 = l11i, m20 = -(l20 m00 + l21 m10) l22i,
    // m21 = -l21 m11 l22i, m22 = l22i;  inverse = M' M
    const float m00 = L[0], m11 = L[2], m22 = L[5];
    const float m10 = -L[1] * m00 * m11;
    const float m21 = -L[4] * m11 * m22;
    const float m20 = -fmaf(L[3], m00, L[4] * m10) * m22;
    c[0] = fmaf(m00, m00, fmaf(m10, m10, m20 * m20));
    c[1] = fmaf(m10, m11, m20 * m21);
    c[2] = fmaf(m11, m11, m21 * m21);
    c[3] = m20 * m22;
    c[4] = m21 * m22;
    c[5] = m22 * m22;
}

// ---- predict, fast formulation -------------------------------------------------------------------------------------
// One noise-free FastSLAM2::predict tick (fastslam2.cpp:51-60,70-125) with the structure of its matrices used instead
// of dense 3x3 products: Gv = I + [0 0 a; 0 0 b; 0 0 0], and josephUpdate's C = I - W*[0 0 1] (core.cpp:294-317), so
// Gv*Pv*Gv' and C*Pv*C' are two rank-one row/column updates each.  Skipping the exact zeros and ones of the dense
// products changes no rounding by itself; what differs from the reference-order path is FMA contraction, sincosf
// for the pair (sin, cos)(G + phi), the reciprocal and the single-rounding angle wrap.  Everything that depends only
// on the controls (sin(G/wb): note the reference's G/wheelBase INSIDE the sine, fastslam2.cpp:103; sin G, cos G in
// Gu) is evaluated once on the host (PredictTick).  Pv stays a general (not symmetrised) 3x3, like the reference's.
// ~120 instructions per tick instead of ~430 (ncu r01: k_predict_batch was fp32-issue bound at 122 us for 8 ticks).
struct PredictTick {
    float vdt;      // V*dt
    float dt;
    float dtheta;   // V*dt*sin(G/wheelBase)
    float G;
    float gu20, gu21;  // third row of Gu: dt*sin(G)/wb, V*dt*cos(G)/wb
    float Q[4];
    float phi_true, R_phi;  // heading observation and its variance sigmaT^2
};

SLAM_HD void predict_tick_fast(float *x, float *C, const PredictTick &t, bool with_cov, bool use_heading) {
    // sin / cos of the heading from the special-function unit (abs. error ~5e-7): they only enter multiplied by
    // V*dt (centimetres per tick), i.e. ~4e-8 m per tick on the pose and as much, relatively, on the covariance
    // increments — and the accurate sincosf was a quarter of this kernel's instructions (it is fp32-issue bound)
    float s, c;
    SLAM_SINCOS_APPROX(wrap_fast(t.G + x[2]), &s, &c);
    const float a = -t.vdt * s, b = t.vdt * c;
    if (with_cov) {
        // T1 = Gv*C: rows 0 and 1 pick up a, b times row 2;  T2 = T1*Gv': columns 0 and 1 pick up a, b times column 2
        float T[9];
#pragma unroll
        for (int j = 0; j < 3; j++) {
            T[j] = fmaf(a, C[6 + j], C[j]);
            T[3 + j] = fmaf(b, C[6 + j], C[3 + j]);
            T[6 + j] = C[6 + j];
        }
#pragma unroll
        for (int i = 0; i < 3; i++) {
            T[i * 3 + 0] = fmaf(a, T[i * 3 + 2], T[i * 3 + 0]);
            T[i * 3 + 1] = fmaf(b, T[i * 3 + 2], T[i * 3 + 1]);
        }
        // Gu*Q*Gu', Gu = [dt*c a; dt*s b; gu20 gu21]
        const float Gu[6] = {t.dt * c, a, t.dt * s, b, t.gu20, t.gu21};
        float M[6];
#pragma unroll
        for (int i = 0; i < 3; i++) {
            M[i * 2 + 0] = fmaf(Gu[i * 2], t.Q[0], Gu[i * 2 + 1] * t.Q[2]);
            M[i * 2 + 1] = fmaf(Gu[i * 2], t.Q[1], Gu[i * 2 + 1] * t.Q[3]);
        }
#pragma unroll
        for (int i = 0; i < 3; i++)
#pragma unroll
            for (int j = 0; j < 3; j++)
                C[i * 3 + j] = fmaf(M[i * 2], Gu[j * 2], fmaf(M[i * 2 + 1], Gu[j * 2 + 1], T[i * 3 + j]));
    }
    x[0] += b;   // x + V*dt*cos(G + phi)
    x[1] -= a;   // y + V*dt*sin(G + phi)
    x[2] = wrap_fast(x[2] + t.dtheta);
    if (use_heading) {
        const float v = wrap_fast(t.phi_true - x[2]);
        const float Si = SLAM_RCP(C[8] + t.R_phi);
        const float W[3] = {C[2] * Si, C[5] * Si, C[8] * Si};
        x[0] = fmaf(W[0], v, x[0]);
        x[1] = fmaf(W[1], v, x[1]);
        x[2] = fmaf(W[2], v, x[2]);
        // CP = C - W * C[2,:];  CPC = CP - CP[:,2] * W';  + W*R*W' + eps*I
        float CP[9];
#pragma unroll
        for (int i = 0; i < 3; i++)
#pragma unroll
            for (int j = 0; j < 3; j++) CP[i * 3 + j] = fmaf(-W[i], C[6 + j], C[i * 3 + j]);
#pragma unroll
        for (int i = 0; i < 3; i++) {
            const float wr = W[i] * t.R_phi;
#pragma unroll
            for (int j = 0; j < 3; j++) {
                const float v2 = fmaf(wr, W[j], fmaf(-CP[i * 3 + 2], W[j], CP[i * 3 + j]));
                C[i * 3 + j] = (i == j) ? v2 + 2.2204e-16f : v2;
            }
        }
    }
}

// Pose-side state of the information-form proposal.
struct ProposalInfo {
    float lam[6];  // Lambda = Pv^-1, packed lower (00, 10, 11, 20, 21, 22)
};

// One observation of FastSLAM2::sampleProposal (fastslam2.cpp:314-350) in information form.
SLAM_HD void proposal_step_fast(float *x, ProposalInfo &pi, const float *R, const float *xf, const float *Pf,
                                const float *z) {
    float d, Hf[4], T[4], S[3];
    jacobian_fast(x, R, xf, Pf, &d, Hf, T, S);
    // the pose moves with every observation here, so sin / cos of the ray come from the special-function unit (abs.
    // error ~5e-7: what reaches the proposal is that times the Kalman gain of one bearing on the pose, < 1e-8)
    float sphi, cphi;
    SLAM_SINCOS_APPROX(z[1] + x[2], &sphi, &cphi);
    const float v0 = z[0] - d, v1 = bearing_innovation(sphi, cphi, Hf[0], Hf[1]);
    // G = S^-1 (symmetric): adjugate / det
    const float rdet = SLAM_RCP(fmaf(S[0], S[2], -S[1] * S[1]));
    const float g00 = S[2] * rdet, g01 = -S[1] * rdet, g11 = S[0] * rdet;
    // rows of Hv are -(Hf0, Hf1, 0) and -(Hf2, Hf3, 1); the signs cancel in Hv' G Hv
    const float t10 = fmaf(g00, Hf[0], g01 * Hf[2]), t11 = fmaf(g00, Hf[1], g01 * Hf[3]), t12 = g01;
    const float t20 = fmaf(g01, Hf[0], g11 * Hf[2]), t21 = fmaf(g01, Hf[1], g11 * Hf[3]), t22 = g11;
    pi.lam[0] += fmaf(Hf[0], t10, Hf[2] * t20);
    pi.lam[1] += fmaf(Hf[1], t10, Hf[3] * t20);
    pi.lam[2] += fmaf(Hf[1], t11, Hf[3] * t21);
    pi.lam[3] += t20;
    pi.lam[4] += t21;
    pi.lam[5] += t22;
    (void) t12;
    // b = Hv' G v = -(h1 u0 + h2 u1), u = G v
    const float u0 = fmaf(g00, v0, g01 * v1), u1 = fmaf(g01, v0, g11 * v1);
    const float b[3] = {-fmaf(Hf[0], u0, Hf[2] * u1), -fmaf(Hf[1], u0, Hf[3] * u1), -u1};
    float L[6], y[3], dxv[3];
    chol3_rs(pi.lam, L);
    chol3_fwd(L, b, y);
    chol3_bwd(L, y, dxv);
    x[0] += dxv[0];
    x[1] += dxv[1];
    x[2] += dxv[2];
}

// exp(-0.5*|L^-1 v|^2) / (2*pi * prod diag L) for a packed reciprocal-diagonal factor (gaussEvaluate with D = 3,
// fastslam2.cpp:127-163: the normaliser is (2*pi)^(3/2 as int) = 2*pi).
SLAM_HD float gauss3_from_factor(const float *L, const float *v) {
    float nin[3];
    chol3_fwd(L, v, nin);
    const float q = fmaf(nin[0], nin[0], fmaf(nin[1], nin[1], nin[2] * nin[2]));
    return SLAM_EXP2(q * SLAM_NEG_HALF_LOG2E) * (SLAM_INV_TWO_PI * L[0] * L[2] * L[5]);
}

// Per-particle prologue: factor Pv0 once; gives Lambda0 = Pv0^-1 and the factor for the prior density.
// Returns false when Pv0 is not numerically positive definite (a pivot <= 0 or NaN): the caller then takes the
// reference-order path, which reproduces what Eigen's LLT does with such a matrix.
SLAM_HD bool proposal_begin_fast(const float *C0, float *L0, ProposalInfo &pi) {
    const float a[6] = {C0[0], C0[3], C0[4], C0[6], C0[7], C0[8]};  // lower triangle, like Eigen's LLT
    chol3_rs(a, L0);
    chol3_inverse(L0, pi.lam);
    const float t = L0[0] * L0[2] * L0[5];
    return (L0[0] > 0.0f) && (L0[2] > 0.0f) && (L0[5] > 0.0f) && (t < 3.0e38f);  // false for NaN / inf
}

// Per-particle epilogue (fastslam2.cpp:352-367): Pv = Lambda^-1, pose sample xs = x + chol(Pv)*g, and the weight
// factor prior/proposal.  Returns prior / proposal; xs out.
SLAM_HD float proposal_finish_fast(const float *x0, const float *L0, const float *x, const ProposalInfo &pi,
                                   const float *g, float *xs) {
    float Ll[6], c[6], Lc[6];
    chol3_rs(pi.lam, Ll);
    chol3_inverse(Ll, c);  // Pv, packed lower
    chol3_rs(c, Lc);       // its lower Cholesky factor (reciprocal diagonals)
    const float l00 = c[0] * Lc[0], l11 = fmaf(-Lc[1], Lc[1], c[2]) * Lc[2],
                l22 = fmaf(-Lc[4], Lc[4], fmaf(-Lc[3], Lc[3], c[5])) * Lc[5];  // diag = a * rsqrt(a)
    xs[0] = l00 * g[0] + x[0];
    xs[1] = fmaf(l11, g[1], Lc[1] * g[0]) + x[1];
    xs[2] = fmaf(l22, g[2], fmaf(Lc[4], g[1], Lc[3] * g[0])) + x[2];
    const float v1[3] = {x0[0] - xs[0], x0[1] - xs[1], trig_offset(x0[2] - xs[2])};
    const float v2[3] = {x[0] - xs[0], x[1] - xs[1], trig_offset(x[2] - xs[2])};
    const float prior = gauss3_from_factor(L0, v1);
    const float prop = gauss3_from_factor(Lc, v2);
    return prior * SLAM_RCP(prop);
}

// Pass 2 for one observation at the (sampled) pose: likelihood factor (likelihoodGivenXv, fastslam2.cpp:370-400, or
// FastSLAM1::computeWeight's term, fastslam1.cpp:91-118 — the same density) and the landmark EKF
// (featureUpdate -> choleskyUpdate, core.cpp:132-175,275-291), sharing one factorisation of S.
// The pose is fixed over the K observations of this pass: its (sin, cos)(theta) are evaluated once per particle,
// (sin, cos)(z_b) once per observation (staged with the observation set), and the ray's are their angle sum.
SLAM_HD float likelihood_and_feature_update_fast(const float *xs, float sth, float cth, const float *R, float *xf,
                                                 float *Pf, float z_range, float sz, float cz) {
    float d, Hf[4], T[4], S[3];
    jacobian_fast(xs, R, xf, Pf, &d, Hf, T, S);
    const float sphi = fmaf(sz, cth, cz * sth), cphi = fmaf(cz, cth, -sz * sth);
    const float v0 = z_range - d, v1 = bearing_innovation(sphi, cphi, Hf[0], Hf[1]);
    const float l00i = SLAM_RSQRT(S[0]);
    const float l10 = S[1] * l00i;
    const float l11i = SLAM_RSQRT(fmaf(-l10, l10, S[2]));
    const float n0 = v0 * l00i;
    const float n1 = fmaf(-l10, n0, v1) * l11i;
    const float q = fmaf(n0, n0, n1 * n1);
    const float like = SLAM_EXP2(q * SLAM_NEG_HALF_LOG2E) * (SLAM_INV_TWO_PI * l00i * l11i);
    // PHt = Pf Hf' = (Hf Pf)' = T' (Pf through its symmetric part, see jacobian_fast)
    const float p00 = T[0], p01 = T[2], p10 = T[1], p11 = T[3];
    // W1 = PHt * U^-1, U = L' = [l00 l10; 0 l11]
    const float w00 = p00 * l00i, w10 = p10 * l00i;
    const float w01 = fmaf(-w00, l10, p01) * l11i, w11 = fmaf(-w10, l10, p11) * l11i;
    xf[0] += fmaf(w00, n0, w01 * n1);
    xf[1] += fmaf(w10, n0, w11 * n1);
    const float c01 = fmaf(w00, w10, w01 * w11);
    Pf[0] -= fmaf(w00, w00, w01 * w01);
    Pf[1] -= c01;
    Pf[2] -= c01;
    Pf[3] -= fmaf(w10, w10, w11 * w11);
    return like;
}
// The same with everything evaluated on the spot (one call per particle-landmark: tests, host emulation).
SLAM_HD float likelihood_and_feature_update_fast(const float *xs, const float *R, float *xf, float *Pf,
                                                 const float *z) {
    return likelihood_and_feature_update_fast(xs, sinf(xs[2]), cosf(xs[2]), R, xf, Pf, z[0], sinf(z[1]), cosf(z[1]));
}

}  // namespace slam
