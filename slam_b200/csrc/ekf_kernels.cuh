// ekf_kernels.cuh — K9: the EKF-SLAM entry point as sm_100a kernels.
//
// One joint state on the device: x[n], P[n][n] row-major with leading dimension ld (n = 3 + 2*Nf grows by augment).
// The reference (algorithms/ekfslam.cpp) forms dense n x n products for everything; here the structure is used:
//   * predict touches the 3x3 pose block and the 3 x (n-3) cross covariance only (ekfslam.cpp:46-77);
//   * observeHeading's Joseph form C*P*C' + W*R*W' + eps*I with C = I - W*e3' is two rank-one corrections, one
//     element-wise pass over P (core.cpp:294-317);
//   * the observation Jacobian H has 5 non-zero columns per landmark (ekfslam.cpp:97-131), so the gated
//     nearest-neighbour association (ekfslam.cpp:133-189) costs 25 entries of P per (observation, landmark) pair
//     instead of a dense H*P*H', and P*H' in batchUpdate is a 5-term sum per entry;
//   * batchUpdate's P -= W1*W1' (core.cpp:275-291) is the one rank-2m pass over P: HBM bound at large n.
// Arithmetic follows the order of the dense products (sequential over the inner index, no FMA contraction: skipping
// an exact zero changes nothing), so the association sees the same nis / nd as the reference up to libm differences.
// Grids are 2-D tiles over P; nothing here is GEMM-shaped enough for tensor cores (inner dimensions 1, 5 and 2m).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "slam_math.cuh"

namespace slam {

#define EM(a, b) __fmul_rn((a), (b))
#define EA(a, b) __fadd_rn((a), (b))
#define ES(a, b) __fsub_rn((a), (b))

struct EkfState {
    float *x;   // [n_cap]
    float *P;   // [n_cap][ld]
    int ld;
};

// ---- predict (ekfslam.cpp:46-77): one CTA; the pose block by thread 0, the cross covariance by column -------------
__global__ void __launch_bounds__(256) k_ekf_predict(EkfState s, int n, float V, float G, float4 Q, float wb, float dt) {
    __shared__ float Gv[9];
    float *P = s.P;
    const int ld = s.ld;
    if (threadIdx.x == 0) {
        const float th = s.x[2];
        const float sn = sinf(EA(G, th)), cs = cosf(EA(G, th));
        const float vts = EM(EM(V, dt), sn), vtc = EM(EM(V, dt), cs);
        const float gv[9] = {1, 0, -vts, 0, 1, vtc, 0, 0, 1};
        const float Gu[6] = {EM(dt, cs), -vts, EM(dt, sn), vtc, __fdiv_rn(EM(dt, sinf(G)), wb),
                             __fdiv_rn(EM(EM(V, dt), cosf(G)), wb)};
        const float q[4] = {Q.x, Q.y, Q.z, Q.w};
        float Pvv[9], T1[9], T2[9], T3[6], T4[9];
        for (int i = 0; i < 3; i++)
            for (int j = 0; j < 3; j++) Pvv[i * 3 + j] = P[i * ld + j];
        for (int i = 0; i < 3; i++)
            for (int j = 0; j < 3; j++) {
                float a = 0;
                for (int k = 0; k < 3; k++) a = EA(a, EM(gv[i * 3 + k], Pvv[k * 3 + j]));
                T1[i * 3 + j] = a;
            }
        for (int i = 0; i < 3; i++)
            for (int j = 0; j < 3; j++) {
                float a = 0;
                for (int k = 0; k < 3; k++) a = EA(a, EM(T1[i * 3 + k], gv[j * 3 + k]));
                T2[i * 3 + j] = a;
            }
        for (int i = 0; i < 3; i++)
            for (int j = 0; j < 2; j++) {
                float a = 0;
                for (int k = 0; k < 2; k++) a = EA(a, EM(Gu[i * 2 + k], q[k * 2 + j]));
                T3[i * 2 + j] = a;
            }
        for (int i = 0; i < 3; i++)
            for (int j = 0; j < 3; j++) {
                float a = 0;
                for (int k = 0; k < 2; k++) a = EA(a, EM(T3[i * 2 + k], Gu[j * 2 + k]));
                T4[i * 3 + j] = a;
            }
        for (int i = 0; i < 3; i++)
            for (int j = 0; j < 3; j++) P[i * ld + j] = EA(T2[i * 3 + j], T4[i * 3 + j]);
        for (int j = 0; j < 9; j++) Gv[j] = gv[j];
        s.x[0] = EA(s.x[0], vtc);
        s.x[1] = EA(s.x[1], vts);
        s.x[2] = trig_offset(EA(th, __fdiv_rn(EM(EM(V, dt), sinf(G)), wb)));
    }
    __syncthreads();
    for (int j = 3 + threadIdx.x; j < n; j += blockDim.x) {
        const float c0 = P[0 * ld + j], c1 = P[1 * ld + j], c2 = P[2 * ld + j];
        for (int i = 0; i < 3; i++) {
            float a = EM(Gv[i * 3], c0);
            a = EA(a, EM(Gv[i * 3 + 1], c1));
            a = EA(a, EM(Gv[i * 3 + 2], c2));
            P[i * ld + j] = a;
            P[(size_t) j * ld + i] = a;
        }
    }
}

// ---- observeHeading -> josephUpdate with H = e3' (ekfslam.cpp:86-95, core.cpp:294-317) -------------------------------
// prep (one CTA): W = P[:,2] / (P22 + R); x += W*v; row 2 and column 2 of the OLD P are saved for the element pass.
__global__ void __launch_bounds__(256) k_ekf_heading_prep(EkfState s, int n, float phi, float R, float *W, float *row2,
                                                          float *col2) {
    const float *P = s.P;
    const int ld = s.ld;
    const float v = trig_offset(ES(phi, s.x[2]));
    const float Si = __fdiv_rn(1.0f, EA(P[2 * ld + 2], R));
    __syncthreads();  // every thread has read x[2] before anyone updates it
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const float pi2 = P[(size_t) i * ld + 2];
        const float w = EM(pi2, Si);
        W[i] = w;
        col2[i] = pi2;
        row2[i] = P[2 * ld + i];
        s.x[i] = EA(s.x[i], EM(w, v));
    }
}
// apply: P[i][j] = (C*P*C')[i][j] + (W_i*R)*W_j + eps*[i==j], in the order the dense products give each entry.
__global__ void __launch_bounds__(256) k_ekf_heading_apply(EkfState s, int n, float R, const float *__restrict__ W,
                                                           const float *__restrict__ row2,
                                                           const float *__restrict__ col2) {
    const int j = blockIdx.x * 32 + (threadIdx.x & 31);
    const int i = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (i >= n || j >= n) return;
    const float wi = W[i], wj = W[j], w2 = W[2];
    const float one_m_w2 = ES(1.0f, w2);
    const float p22 = row2[2];
    const float pij = s.P[(size_t) i * s.ld + j];
    // CP[i][j] and CP[i][2]
    const float cp_ij = (i == 2) ? EM(one_m_w2, pij) : EA(pij, EM(-wi, row2[j]));
    const float cp_i2 = (i == 2) ? EM(one_m_w2, p22) : EA(col2[i], EM(-wi, p22));
    const float cpc = (j == 2) ? EM(cp_i2, one_m_w2) : EA(cp_ij, EM(cp_i2, -wj));
    const float wrw = EM(EM(wi, R), wj);
    float out = EA(cpc, wrw);
    if (i == j) out = EA(out, 2.2204e-16f);
    s.P[(size_t) i * s.ld + j] = out;
}

// ---- observation model (ekfslam.cpp:97-131): 5 non-zero columns 0,1,2,fpos,fpos+1 -------------------------------------
__device__ __forceinline__ void ekf_observe_model(const float *x, int idf, float *zp, float *Hs) {
    const int fpos = 3 + idf * 2;
    const float dx = ES(x[fpos], x[0]), dy = ES(x[fpos + 1], x[1]);
    const float d2 = EA(EM(dx, dx), EM(dy, dy));
    const float d = __fsqrt_rn(d2);
    const float xd = __fdiv_rn(dx, d), yd = __fdiv_rn(dy, d), xd2 = __fdiv_rn(dx, d2), yd2 = __fdiv_rn(dy, d2);
    zp[0] = d;
    zp[1] = ES(atan2f(dy, dx), x[2]);
    Hs[0] = -xd; Hs[1] = -yd; Hs[2] = 0; Hs[3] = xd; Hs[4] = yd;
    Hs[5] = yd2; Hs[6] = -xd2; Hs[7] = -1; Hs[8] = -yd2; Hs[9] = xd2;
}

// ---- gated nearest neighbour (ekfslam.cpp:133-189): one thread per (observation, landmark) pair ----------------------
__global__ void __launch_bounds__(128) k_ekf_assoc_pairs(EkfState s, int Nf, int m, const float2 *__restrict__ z,
                                                         float4 R, float *__restrict__ nis_out,
                                                         float *__restrict__ nd_out) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    const int i = blockIdx.y;
    if (j >= Nf || i >= m) return;
    const float *P = s.P;
    const int ld = s.ld;
    float zp[2], Hs[10];
    ekf_observe_model(s.x, j, zp, Hs);
    const int cols[5] = {0, 1, 2, 3 + 2 * j, 4 + 2 * j};
    const float v0 = ES(z[i].x, zp[0]), v1 = trig_offset(ES(z[i].y, zp[1]));
    float HP[2][5];
#pragma unroll
    for (int c = 0; c < 5; c++) {
        float pc[5];
#pragma unroll
        for (int k = 0; k < 5; k++) pc[k] = P[(size_t) cols[k] * ld + cols[c]];
#pragma unroll
        for (int r = 0; r < 2; r++) {
            float a = 0;
#pragma unroll
            for (int k = 0; k < 5; k++) a = EA(a, EM(Hs[r * 5 + k], pc[k]));
            HP[r][c] = a;
        }
    }
    const float Rm[4] = {R.x, R.y, R.z, R.w};
    float S[4];
#pragma unroll
    for (int r = 0; r < 2; r++)
#pragma unroll
        for (int c = 0; c < 2; c++) {
            float a = 0;
#pragma unroll
            for (int k = 0; k < 5; k++) a = EA(a, EM(HP[r][k], Hs[c * 5 + k]));
            S[r * 2 + c] = EA(a, Rm[r * 2 + c]);
        }
    float Si[4], det;
    inv2_lu(S, Si, &det);  // Eigen's dynamic 2x2 inverse()/determinant(): PartialPivLU (slam_math.cuh)
    const float t0 = EA(EM(v0, Si[0]), EM(v1, Si[2])), t1 = EA(EM(v0, Si[1]), EM(v1, Si[3]));
    const float nis = EA(EM(t0, v0), EM(t1, v1));
    nis_out[(size_t) i * Nf + j] = nis;
    nd_out[(size_t) i * Nf + j] = EA(nis, logf(det));
}

// One CTA per observation: the reference's sequential scan over the landmarks (ekfslam.cpp:165-178) as a reduction.
// jbest = first landmark with the smallest nd among those inside gate1 (strict '<' keeps the first on ties);
// when there is none, every landmark went through the else branch and outer = min nis.
__global__ void __launch_bounds__(256) k_ekf_assoc_reduce(int Nf, const float *__restrict__ nis,
                                                          const float *__restrict__ nd, float gate1, int *jbest_out,
                                                          float *outer_out) {
    __shared__ float s_nd[256], s_out[256];
    __shared__ int s_j[256];
    const int i = blockIdx.x;
    float bnd = INFINITY, bout = INFINITY;  // the reference starts both at float(1e60) = +inf
    int bj = -1;
    for (int j = threadIdx.x; j < Nf; j += blockDim.x) {
        const float a = nis[(size_t) i * Nf + j], b = nd[(size_t) i * Nf + j];
        if (a < gate1 && b < bnd) {
            bnd = b;
            bj = j;
        }
        if (a < bout) bout = a;
    }
    s_nd[threadIdx.x] = bnd;
    s_j[threadIdx.x] = bj;
    s_out[threadIdx.x] = bout;
    __syncthreads();
    for (int st = 128; st > 0; st >>= 1) {
        if (threadIdx.x < st) {
            const float ond = s_nd[threadIdx.x + st];
            const int oj = s_j[threadIdx.x + st];
            if (oj >= 0 && (s_j[threadIdx.x] < 0 || ond < s_nd[threadIdx.x] ||
                            (ond == s_nd[threadIdx.x] && oj < s_j[threadIdx.x]))) {
                s_nd[threadIdx.x] = ond;
                s_j[threadIdx.x] = oj;
            }
            s_out[threadIdx.x] = fminf(s_out[threadIdx.x], s_out[threadIdx.x + st]);
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        jbest_out[i] = s_j[0];
        outer_out[i] = s_out[0];
    }
}

// ---- batchUpdate (ekfslam.cpp:238-267) -> choleskyUpdate (core.cpp:275-291) ---------------------------------------------
struct EkfBatch {        // scratch of one batch update, q = 2*m rows
    float *Hs;           // [m][10] sparse rows of H
    float *v;            // [q] innovation
    float *PHt;          // [n][q]
    float *U, *Ui;       // [q][q] chol(S) upper, its inverse
    float *W1;           // [n][q]
    const int *idf;      // [m]
    const float2 *zf;    // [m]
};

__global__ void k_ekf_innovation(EkfState s, EkfBatch b, int m) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    float zp[2];
    ekf_observe_model(s.x, b.idf[i], zp, b.Hs + 10 * i);
    b.v[2 * i] = ES(b.zf[i].x, zp[0]);
    b.v[2 * i + 1] = trig_offset(ES(b.zf[i].y, zp[1]));
}

// PHt = P*H': entry (i, 2a+r) is the 5-term sum over the non-zero columns of row r of observation a.
__global__ void __launch_bounds__(256) k_ekf_pht(EkfState s, EkfBatch b, int n, int m) {
    const int c = blockIdx.x * 32 + (threadIdx.x & 31);
    const int i = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (i >= n || c >= 2 * m) return;
    const int a = c >> 1, r = c & 1;
    const int f = 3 + 2 * b.idf[a];
    const int cols[5] = {0, 1, 2, f, f + 1};
    const float *Hs = b.Hs + 10 * a + 5 * r;
    float acc = 0;
#pragma unroll
    for (int k = 0; k < 5; k++) acc = EA(acc, EM(s.P[(size_t) i * s.ld + cols[k]], Hs[k]));
    b.PHt[(size_t) i * (2 * m) + c] = acc;
}

// One CTA: S = H*PHt + RR, symmetrised; S = U'U (llt().matrixU()); Ui = U^-1 by back substitution.
__global__ void __launch_bounds__(256) k_ekf_factor(EkfBatch b, int m, float4 R) {
    const int q = 2 * m;
    float *U = b.U, *Ui = b.Ui;
    const float Rm[4] = {R.x, R.y, R.z, R.w};
    // raw S into Ui (scratch), then the symmetrised upper triangle into U
    for (int e = threadIdx.x; e < q * q; e += blockDim.x) {
        const int i = e / q, j = e % q;
        const int a = i >> 1, r = i & 1;
        const int f = 3 + 2 * b.idf[a];
        const int cols[5] = {0, 1, 2, f, f + 1};
        const float *Hs = b.Hs + 10 * a + 5 * r;
        float acc = 0;
        for (int k = 0; k < 5; k++) acc = EA(acc, EM(Hs[k], b.PHt[(size_t) cols[k] * q + j]));
        const float rr = ((j >> 1) == a) ? Rm[r * 2 + (j & 1)] : 0.0f;
        Ui[e] = EA(acc, rr);
    }
    __syncthreads();
    for (int e = threadIdx.x; e < q * q; e += blockDim.x) {
        const int i = e / q, j = e % q;
        U[e] = (j >= i) ? EM(EA(Ui[i * q + j], Ui[j * q + i]), 0.5f) : 0.0f;
    }
    __syncthreads();
    for (int k = 0; k < q; k++) {
        if (threadIdx.x == 0) {
            float d = U[k * q + k];
            for (int t = 0; t < k; t++) d = ES(d, EM(U[t * q + k], U[t * q + k]));
            U[k * q + k] = __fsqrt_rn(d);
        }
        __syncthreads();
        const float d = U[k * q + k];
        for (int j = k + 1 + threadIdx.x; j < q; j += blockDim.x) {
            float a = U[k * q + j];
            for (int t = 0; t < k; t++) a = ES(a, EM(U[t * q + k], U[t * q + j]));
            U[k * q + j] = __fdiv_rn(a, d);
        }
        __syncthreads();
    }
    for (int c = threadIdx.x; c < q; c += blockDim.x)
        for (int i = q - 1; i >= 0; i--) {
            float a = (i == c) ? 1.0f : 0.0f;
            for (int k = i + 1; k < q; k++) a = ES(a, EM(U[i * q + k], Ui[k * q + c]));
            Ui[i * q + c] = __fdiv_rn(a, U[i * q + i]);
        }
}

// One thread per state row: W1 = PHt*Ui, W = W1*Ui', x += W*v.
__global__ void __launch_bounds__(128) k_ekf_gain(EkfState s, EkfBatch b, int n, int m) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int q = 2 * m;
    const float *ph = b.PHt + (size_t) i * q;
    float *w1 = b.W1 + (size_t) i * q;
    for (int j = 0; j < q; j++) {
        float a = 0;
        for (int k = 0; k < q; k++) a = EA(a, EM(ph[k], b.Ui[k * q + j]));
        w1[j] = a;
    }
    float dx = 0;
    for (int j = 0; j < q; j++) {
        float a = 0;
        for (int k = 0; k < q; k++) a = EA(a, EM(w1[k], b.Ui[j * q + k]));
        dx = EA(dx, EM(a, b.v[j]));
    }
    s.x[i] = EA(s.x[i], dx);
}

// P -= W1*W1': the rank-2m pass over P (read + write of n^2 floats; W1 rows come from L2).
__global__ void __launch_bounds__(256) k_ekf_downdate(EkfState s, EkfBatch b, int n, int m) {
    const int j = blockIdx.x * 32 + (threadIdx.x & 31);
    const int i = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (i >= n || j >= n) return;
    const int q = 2 * m;
    const float *wi = b.W1 + (size_t) i * q, *wj = b.W1 + (size_t) j * q;
    float a = 0;
    for (int k = 0; k < q; k++) a = EA(a, EM(wi[k], wj[k]));
    s.P[(size_t) i * s.ld + j] = ES(s.P[(size_t) i * s.ld + j], a);
}

// ---- augment (ekfslam.cpp:269-323): one new landmark appended at state index n; one CTA -------------------------------
__global__ void __launch_bounds__(256) k_ekf_add_one(EkfState s, int n, float r, float bg, float4 Re) {
    __shared__ float Gv[6];
    float *P = s.P;
    const int ld = s.ld;
    if (threadIdx.x == 0) {
        const float sn = sinf(EA(s.x[2], bg)), cs = cosf(EA(s.x[2], bg));
        s.x[n] = EA(s.x[0], EM(r, cs));
        s.x[n + 1] = EA(s.x[1], EM(r, sn));
        const float gv[6] = {1, 0, EM(-r, sn), 0, 1, EM(r, cs)};
        const float Gz[4] = {cs, EM(-r, sn), sn, EM(r, cs)};
        const float re[4] = {Re.x, Re.y, Re.z, Re.w};
        float gp[6], A[4], T[4], B[4];
        for (int i = 0; i < 2; i++)
            for (int j = 0; j < 3; j++) {
                float a = 0;
                for (int k = 0; k < 3; k++) a = EA(a, EM(gv[i * 3 + k], P[k * ld + j]));
                gp[i * 3 + j] = a;
            }
        for (int i = 0; i < 2; i++)
            for (int j = 0; j < 2; j++) {
                float a = 0;
                for (int k = 0; k < 3; k++) a = EA(a, EM(gp[i * 3 + k], gv[j * 3 + k]));
                A[i * 2 + j] = a;
                float t = 0;
                for (int k = 0; k < 2; k++) t = EA(t, EM(Gz[i * 2 + k], re[k * 2 + j]));
                T[i * 2 + j] = t;
            }
        for (int i = 0; i < 2; i++)
            for (int j = 0; j < 2; j++) {
                float a = 0;
                for (int k = 0; k < 2; k++) a = EA(a, EM(T[i * 2 + k], Gz[j * 2 + k]));
                B[i * 2 + j] = a;
            }
        for (int i = 0; i < 2; i++)
            for (int j = 0; j < 2; j++) P[(size_t) (n + i) * ld + n + j] = EA(A[i * 2 + j], B[i * 2 + j]);
        for (int i = 0; i < 2; i++)
            for (int j = 0; j < 3; j++) {
                P[(size_t) (n + i) * ld + j] = gp[i * 3 + j];
                P[(size_t) j * ld + n + i] = gp[i * 3 + j];
            }
        for (int j = 0; j < 6; j++) Gv[j] = gv[j];
    }
    __syncthreads();
    for (int j = 3 + threadIdx.x; j < n; j += blockDim.x) {
        const float c0 = P[0 * ld + j], c1 = P[1 * ld + j], c2 = P[2 * ld + j];
        for (int i = 0; i < 2; i++) {
            float a = EM(Gv[i * 3], c0);
            a = EA(a, EM(Gv[i * 3 + 1], c1));
            a = EA(a, EM(Gv[i * 3 + 2], c2));
            P[(size_t) (n + i) * ld + j] = a;
            P[(size_t) j * ld + n + i] = a;
        }
    }
}

#undef EM
#undef EA
#undef ES

}  // namespace slam
