/* TEST INFRASTRUCTURE ONLY — CPU restatement of the reference's EKF-SLAM.
 *
 * Plain C on flat row-major arrays, following src/backend/algorithms/ekfslam.cpp and the dense helpers of
 * src/backend/core.cpp (choleskyUpdate :275-291, josephUpdate :294-317) line by line.  Dense products accumulate
 * sequentially over the inner index in fp32 without contraction, like Eigen 3.1.3's SSE kernels do per output element
 * (the zero entries of H and C then contribute exact zeros).  The state is x[n] and P[n][n] stored with leading
 * dimension ld (>= n) so that augment can grow it in place.
 *
 * Parity pinning: tests/test_ekf_oracle_vs_ref.py drives this file and the UNMODIFIED reference
 * (oracle/_ref/libslamref.so through oracle/ref_shim_ekf.cpp) with the same inputs; tests/golden/ekf_drive.npz holds a
 * drive produced by that reference.  Only tests/ may load this; the product never does.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "ekf_oracle.h"
#include "slam_oracle.h"

#define PX(i, j) P[(size_t) (i) * ld + (j)]

/* ekfslam.cpp:46-77 */
void so_ekf_predict(int n, int ld, float *x, float *P, float V, float G, const float *Q, float wheel_base, float dt) {
    const float s = sinf(G + x[2]), c = cosf(G + x[2]);
    const float vts = V * dt * s, vtc = V * dt * c;
    const float Gv[9] = {1, 0, -vts, 0, 1, vtc, 0, 0, 1};
    const float Gu[6] = {dt * c, -vts, dt * s, vtc, dt * sinf(G) / wheel_base, V * dt * cosf(G) / wheel_base};
    float Pvv[9], T1[9], T2[9], T3[6], T4[9];
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) Pvv[i * 3 + j] = PX(i, j);
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
            float a = 0;
            for (int k = 0; k < 3; k++) a = a + Gv[i * 3 + k] * Pvv[k * 3 + j];
            T1[i * 3 + j] = a;
        }
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
            float a = 0;
            for (int k = 0; k < 3; k++) a = a + T1[i * 3 + k] * Gv[j * 3 + k];
            T2[i * 3 + j] = a;
        }
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 2; j++) {
            float a = 0;
            for (int k = 0; k < 2; k++) a = a + Gu[i * 2 + k] * Q[k * 2 + j];
            T3[i * 2 + j] = a;
        }
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
            float a = 0;
            for (int k = 0; k < 2; k++) a = a + T3[i * 2 + k] * Gu[j * 2 + k];
            T4[i * 3 + j] = a;
        }
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) PX(i, j) = T2[i * 3 + j] + T4[i * 3 + j];
    for (int j = 3; j < n; j++) { /* P(0:3, 3:) = Gv * P(0:3, 3:), then mirrored */
        float col[3] = {PX(0, j), PX(1, j), PX(2, j)}, out[3];
        for (int i = 0; i < 3; i++) {
            float a = 0;
            for (int k = 0; k < 3; k++) a = a + Gv[i * 3 + k] * col[k];
            out[i] = a;
        }
        for (int i = 0; i < 3; i++) {
            PX(i, j) = out[i];
            PX(j, i) = out[i];
        }
    }
    x[0] = x[0] + vtc;
    x[1] = x[1] + vts;
    x[2] = so_trig_offset(x[2] + V * dt * sinf(G) / wheel_base);
}

/* observeHeading ekfslam.cpp:86-95 -> josephUpdate core.cpp:294-317 with H = e3' (dense n x n Joseph form). */
void so_ekf_observe_heading(int n, int ld, float *x, float *P, float phi, float sigma_phi) {
    const float v = so_trig_offset(phi - x[2]);
    const float R = (float) pow((double) sigma_phi, 2.0); /* pow(sigmaPhi, 2) is double, passed as float */
    float *PHt = (float *) malloc(sizeof(float) * n), *W = (float *) malloc(sizeof(float) * n);
    float *CP = (float *) malloc(sizeof(float) * (size_t) n * n);
    for (int i = 0; i < n; i++) PHt[i] = PX(i, 2);
    const float S = PHt[2] + R;
    const float Si = 1.0f / S;
    for (int i = 0; i < n; i++) W[i] = PHt[i] * Si;
    for (int i = 0; i < n; i++) x[i] = x[i] + W[i] * v;
    /* C = I - W*H: identity except column 2.  CP = C*P, then CP*C', sequential over the inner index. */
    for (int i = 0; i < n; i++)
        for (int j = 0; j < n; j++) {
            float a = 0;
            for (int k = 0; k < n; k++) {
                const float cik = ((i == k) ? 1.0f : 0.0f) - ((k == 2) ? W[i] : 0.0f);
                a = a + cik * PX(k, j);
            }
            CP[(size_t) i * n + j] = a;
        }
    const float eps = (float) (2.2204 * pow(10.0, -16));
    for (int i = 0; i < n; i++)
        for (int j = 0; j < n; j++) {
            float a = 0;
            for (int k = 0; k < n; k++) {
                const float cjk = ((j == k) ? 1.0f : 0.0f) - ((k == 2) ? W[j] : 0.0f);
                a = a + CP[(size_t) i * n + k] * cjk;
            }
            const float wrw = (W[i] * R) * W[j];
            PX(i, j) = (a + wrw) + ((i == j) ? eps : 0.0f);
        }
    free(PHt);
    free(W);
    free(CP);
}

/* ekfObserveModel ekfslam.cpp:97-131.  H is returned sparse: its 5 non-zero columns are 0, 1, 2, fpos, fpos+1 and
 * Hs[r*5 + c] holds row r at those columns. */
void so_ekf_observe_model(const float *x, int idf, float *z, float *Hs) {
    const int fpos = 3 + idf * 2;
    const float dx = x[fpos] - x[0], dy = x[fpos + 1] - x[1];
    const float d2 = dx * dx + dy * dy;
    const float d = sqrtf(d2);
    const float xd = dx / d, yd = dy / d, xd2 = dx / d2, yd2 = dy / d2;
    z[0] = d;
    z[1] = atan2f(dy, dx) - x[2];
    Hs[0] = -xd; Hs[1] = -yd; Hs[2] = 0; Hs[3] = xd; Hs[4] = yd;
    Hs[5] = yd2; Hs[6] = -xd2; Hs[7] = -1; Hs[8] = -yd2; Hs[9] = xd2;
}

static void sparse_cols(int idf, int *cols) {
    cols[0] = 0; cols[1] = 1; cols[2] = 2; cols[3] = 3 + idf * 2; cols[4] = 4 + idf * 2;
}

/* Eigen's inverse()/determinant() of a dynamic 2x2: PartialPivLU (same steps as slam_oracle.c inv2_lu). */
static void inv2_lu_ekf(const float *A, float *X, float *det) {
    float a00 = A[0], a01 = A[1], a10 = A[2], a11 = A[3];
    int swap = fabsf(a10) > fabsf(a00);
    if (swap) { float t; t = a00; a00 = a10; a10 = t; t = a01; a01 = a11; a11 = t; }
    const float l = a10 / a00;
    const float u11 = a11 - l * a01;
    *det = swap ? -(a00 * u11) : (a00 * u11);
    for (int c = 0; c < 2; c++) { /* solve L U X = P I, column c */
        float b0 = (c == 0) ? 1.0f : 0.0f, b1 = (c == 1) ? 1.0f : 0.0f;
        if (swap) { float t = b0; b0 = b1; b1 = t; }
        const float y1 = b1 - l * b0;
        const float x1 = y1 / u11;
        const float x0 = (b0 - a01 * x1) / a00;
        X[0 * 2 + c] = x0;
        X[1 * 2 + c] = x1;
    }
}

/* ekfComputeAssociation ekfslam.cpp:133-149: S = H*P*H' + R over the 5 non-zero columns, in ascending column order. */
void so_ekf_compute_association(int ld, const float *x, const float *P, const float *z, const float *R, int idf,
                                float *nis, float *nd) {
    float zp[2], Hs[10], v[2];
    int cols[5];
    so_ekf_observe_model(x, idf, zp, Hs);
    sparse_cols(idf, cols);
    v[0] = z[0] - zp[0];
    v[1] = so_trig_offset(z[1] - zp[1]);
    float HP[2][5]; /* (H*P) at the 5 columns that H' then picks */
    for (int r = 0; r < 2; r++)
        for (int c = 0; c < 5; c++) {
            float a = 0;
            for (int k = 0; k < 5; k++) a = a + Hs[r * 5 + k] * PX(cols[k], cols[c]);
            HP[r][c] = a;
        }
    float S[4], Si[4], det;
    for (int r = 0; r < 2; r++)
        for (int c = 0; c < 2; c++) {
            float a = 0;
            for (int k = 0; k < 5; k++) a = a + HP[r][k] * Hs[c * 5 + k];
            S[r * 2 + c] = a + R[r * 2 + c];
        }
    inv2_lu_ekf(S, Si, &det);
    /* v' * Si * v: (v' Si) first, then the dot with v */
    const float t0 = v[0] * Si[0] + v[1] * Si[2], t1 = v[0] * Si[1] + v[1] * Si[3];
    *nis = t0 * v[0] + t1 * v[1];
    *nd = *nis + logf(det);
}

/* dataAssociate ekfslam.cpp:151-189: gated nearest neighbour.  zf_idx / zn_idx receive the indices (into z) of the
 * associated / new observations in the order the reference pushes them. */
void so_ekf_data_associate(int n, int ld, const float *x, const float *P, const float *z, int m, const float *R,
                           float gate1, float gate2, int *idf, int *zf_idx, int *n_zf, int *zn_idx, int *n_zn) {
    const int Nf = (n - 3) / 2;
    *n_zf = *n_zn = 0;
    for (int i = 0; i < m; i++) {
        long jbest = -1;
        float nbest = (float) 1e60, outer = (float) 1e60; /* 1e60 -> +inf as float, like the reference's initialisers */
        for (int j = 0; j < Nf; j++) {
            float nis, nd;
            so_ekf_compute_association(ld, x, P, z + 2 * i, R, j, &nis, &nd);
            if (nis < gate1 && nd < nbest) {
                nbest = nd;
                jbest = j;
            } else if (nis < outer) {
                outer = nis;
            }
        }
        if (jbest > -1) {
            zf_idx[*n_zf] = i;
            idf[*n_zf] = (int) jbest;
            (*n_zf)++;
        } else if (outer > gate2) {
            zn_idx[*n_zn] = i;
            (*n_zn)++;
        }
    }
}

/* dataAssociateKnown ekfslam.cpp:201-236; the table is updated in place (the wrapper's by-value call discards it). */
void so_ekf_data_associate_known(int n, const int *idz, int m, int *table, int *idf, int *zf_idx, int *n_zf,
                                 int *zn_idx, int *n_zn) {
    const int Nf = (n - 3) / 2;
    *n_zf = *n_zn = 0;
    for (int i = 0; i < m; i++) {
        const int ii = idz[i];
        if (table[ii] == -1) {
            zn_idx[(*n_zn)++] = i;
        } else {
            zf_idx[*n_zf] = i;
            idf[(*n_zf)++] = table[ii];
        }
    }
    for (int i = 0; i < *n_zn; i++) table[idz[zn_idx[i]]] = Nf + i;
}

/* batchUpdate ekfslam.cpp:238-267 -> choleskyUpdate core.cpp:275-291, dense (H is 2m x n). */
void so_ekf_batch_update(int n, int ld, float *x, float *P, const float *zf, const int *idf, int m, const float *R) {
    const int q = 2 * m;
    if (m == 0) return; /* empty H: every product is empty, x and P are unchanged */
    float *H = (float *) calloc((size_t) q * n, sizeof(float)), *v = (float *) calloc(q, sizeof(float));
    float *RR = (float *) calloc((size_t) q * q, sizeof(float));
    float *PHt = (float *) malloc(sizeof(float) * (size_t) n * q), *S = (float *) malloc(sizeof(float) * q * q);
    float *U = (float *) calloc((size_t) q * q, sizeof(float)), *Ui = (float *) calloc((size_t) q * q, sizeof(float));
    float *W1 = (float *) malloc(sizeof(float) * (size_t) n * q), *W = (float *) malloc(sizeof(float) * (size_t) n * q);
    for (int i = 0; i < m; i++) {
        float zp[2], Hs[10];
        int cols[5];
        so_ekf_observe_model(x, idf[i], zp, Hs);
        sparse_cols(idf[i], cols);
        for (int r = 0; r < 2; r++)
            for (int c = 0; c < 5; c++) H[(size_t) (2 * i + r) * n + cols[c]] = Hs[r * 5 + c];
        v[2 * i] = zf[2 * i] - zp[0];
        v[2 * i + 1] = so_trig_offset(zf[2 * i + 1] - zp[1]);
        for (int r = 0; r < 2; r++)
            for (int c = 0; c < 2; c++) RR[(size_t) (2 * i + r) * q + 2 * i + c] = R[r * 2 + c];
    }
    for (int i = 0; i < n; i++)
        for (int j = 0; j < q; j++) {
            float a = 0;
            for (int k = 0; k < n; k++) a = a + PX(i, k) * H[(size_t) j * n + k];
            PHt[(size_t) i * q + j] = a;
        }
    for (int i = 0; i < q; i++)
        for (int j = 0; j < q; j++) {
            float a = 0;
            for (int k = 0; k < n; k++) a = a + H[(size_t) i * n + k] * PHt[(size_t) k * q + j];
            S[i * q + j] = a + RR[i * q + j];
        }
    for (int i = 0; i < q; i++)
        for (int j = i; j < q; j++) {
            const float sym = (S[i * q + j] + S[j * q + i]) * 0.5f;
            U[i * q + j] = sym; /* upper triangle of the symmetrised S, factored in place below */
        }
    /* S = U'U (llt().matrixU()): row-by-row Cholesky on the upper triangle */
    for (int k = 0; k < q; k++) {
        float d = U[k * q + k];
        for (int t = 0; t < k; t++) d = d - U[t * q + k] * U[t * q + k];
        d = sqrtf(d);
        U[k * q + k] = d;
        for (int j = k + 1; j < q; j++) {
            float a = U[k * q + j];
            for (int t = 0; t < k; t++) a = a - U[t * q + k] * U[t * q + j];
            U[k * q + j] = a / d;
        }
    }
    /* SChol.inverse(): U * X = I by back substitution, column by column */
    for (int c = 0; c < q; c++)
        for (int i = q - 1; i >= 0; i--) {
            float a = (i == c) ? 1.0f : 0.0f;
            for (int k = i + 1; k < q; k++) a = a - U[i * q + k] * Ui[k * q + c];
            Ui[i * q + c] = a / U[i * q + i];
        }
    for (int i = 0; i < n; i++)
        for (int j = 0; j < q; j++) {
            float a = 0;
            for (int k = 0; k < q; k++) a = a + PHt[(size_t) i * q + k] * Ui[k * q + j];
            W1[(size_t) i * q + j] = a;
        }
    for (int i = 0; i < n; i++)
        for (int j = 0; j < q; j++) {
            float a = 0;
            for (int k = 0; k < q; k++) a = a + W1[(size_t) i * q + k] * Ui[j * q + k];
            W[(size_t) i * q + j] = a;
        }
    for (int i = 0; i < n; i++) {
        float a = 0;
        for (int k = 0; k < q; k++) a = a + W[(size_t) i * q + k] * v[k];
        x[i] = x[i] + a;
    }
    for (int i = 0; i < n; i++)
        for (int j = 0; j < n; j++) {
            float a = 0;
            for (int k = 0; k < q; k++) a = a + W1[(size_t) i * q + k] * W1[(size_t) j * q + k];
            PX(i, j) = PX(i, j) - a;
        }
    free(H); free(v); free(RR); free(PHt); free(S); free(U); free(Ui); free(W1); free(W);
}

/* ekfAddOneZ ekfslam.cpp:269-316; returns the new state size (n + 2).  Needs ld >= n + 2. */
int so_ekf_add_one_z(int n, int ld, float *x, float *P, const float *z, const float *Re) {
    const float r = z[0], b = z[1];
    const float s = sinf(x[2] + b), c = cosf(x[2] + b);
    x[n] = x[0] + r * c;
    x[n + 1] = x[1] + r * s;
    const float Gv[6] = {1, 0, -r * s, 0, 1, r * c};
    const float Gz[4] = {c, -r * s, s, r * c};
    float GP[6]; /* Gv * P(0:3,0:3) */
    for (int i = 0; i < 2; i++)
        for (int j = 0; j < 3; j++) {
            float a = 0;
            for (int k = 0; k < 3; k++) a = a + Gv[i * 3 + k] * PX(k, j);
            GP[i * 3 + j] = a;
        }
    float A[4], T[4], B[4];
    for (int i = 0; i < 2; i++)
        for (int j = 0; j < 2; j++) {
            float a = 0;
            for (int k = 0; k < 3; k++) a = a + GP[i * 3 + k] * Gv[j * 3 + k];
            A[i * 2 + j] = a;
        }
    for (int i = 0; i < 2; i++)
        for (int j = 0; j < 2; j++) {
            float a = 0;
            for (int k = 0; k < 2; k++) a = a + Gz[i * 2 + k] * Re[k * 2 + j];
            T[i * 2 + j] = a;
        }
    for (int i = 0; i < 2; i++)
        for (int j = 0; j < 2; j++) {
            float a = 0;
            for (int k = 0; k < 2; k++) a = a + T[i * 2 + k] * Gz[j * 2 + k];
            B[i * 2 + j] = a;
        }
    for (int i = 0; i < 2; i++)
        for (int j = 0; j < 2; j++) PX(n + i, n + j) = A[i * 2 + j] + B[i * 2 + j];
    for (int i = 0; i < 2; i++)
        for (int j = 0; j < 3; j++) {
            PX(n + i, j) = GP[i * 3 + j];
            PX(j, n + i) = GP[i * 3 + j];
        }
    for (int j = 3; j < n; j++) /* map to feature cross correlation: Gv * P(0:3, j) */
        for (int i = 0; i < 2; i++) {
            float a = 0;
            for (int k = 0; k < 3; k++) a = a + Gv[i * 3 + k] * PX(k, j);
            PX(n + i, j) = a;
            PX(j, n + i) = a;
        }
    return n + 2;
}

/* EKFSLAM::sim ekfslam.cpp:18-43.  Returns the new state size.  table may be NULL unless association_known. */
int so_ekf_sim(const so_ekf_config *cfg, int n, int ld, float *x, float *P, float V, float G, const float *Qe, float dt,
               float phi, int observe, const float *z, const int *idz, int m, int *table, const float *Re,
               const float *R, int *idf_out, int *n_zf_out, int *n_zn_out) {
    so_ekf_predict(n, ld, x, P, V, G, Qe, cfg->wheel_base, dt);
    if (cfg->use_heading) so_ekf_observe_heading(n, ld, x, P, phi, cfg->sigma_phi);
    int n_zf = 0, n_zn = 0;
    if (observe) {
        int *zf_idx = (int *) malloc(sizeof(int) * (m + 1)), *zn_idx = (int *) malloc(sizeof(int) * (m + 1));
        float *zf = (float *) malloc(sizeof(float) * 2 * (m + 1));
        if (cfg->association_known)
            so_ekf_data_associate_known(n, idz, m, table, idf_out, zf_idx, &n_zf, zn_idx, &n_zn);
        else
            so_ekf_data_associate(n, ld, x, P, z, m, Re, cfg->gate_reject, cfg->gate_augment, idf_out, zf_idx, &n_zf,
                                  zn_idx, &n_zn);
        for (int i = 0; i < n_zf; i++) {
            zf[2 * i] = z[2 * zf_idx[i]];
            zf[2 * i + 1] = z[2 * zf_idx[i] + 1];
        }
        if (cfg->batch_update) so_ekf_batch_update(n, ld, x, P, zf, idf_out, n_zf, R);
        for (int i = 0; i < n_zn; i++) n = so_ekf_add_one_z(n, ld, x, P, z + 2 * zn_idx[i], Re);
        free(zf_idx); free(zn_idx); free(zf);
    }
    if (n_zf_out) *n_zf_out = n_zf;
    if (n_zn_out) *n_zn_out = n_zn;
    return n;
}
