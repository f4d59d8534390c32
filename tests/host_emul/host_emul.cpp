// TEST INFRASTRUCTURE ONLY — compiles the product's per-particle arithmetic (slam_b200/csrc/slam_math.cuh) with the
// HOST compiler so that the build container (no GPU) can diff it against the oracle before GPU time is spent.
// The loops below mirror what the kernels in kernels.cuh do per thread.  Nothing in the product links this file.
#include "../../slam_b200/csrc/slam_math.cuh"

using namespace slam;

static void lm_load(long P, long p, long l, const float *lm_mean, const float *lm_cov, float *xf, float *Pf) {
    const float *m = lm_mean + (l * P + p) * 2;
    const float *c = lm_cov + (l * P + p) * 4;
    xf[0] = m[0]; xf[1] = m[1];
    Pf[0] = c[0]; Pf[1] = c[2]; Pf[2] = c[1]; Pf[3] = c[3];
}
static void lm_store(long P, long p, long l, float *lm_mean, float *lm_cov, const float *xf, const float *Pf) {
    float *m = lm_mean + (l * P + p) * 2;
    float *c = lm_cov + (l * P + p) * 4;
    m[0] = xf[0]; m[1] = xf[1];
    c[0] = Pf[0]; c[1] = Pf[2]; c[2] = Pf[1]; c[3] = Pf[3];
}

extern "C" {

float he_trig_offset(float a) { return trig_offset(a); }

void he_compute_jacobians(const float *in, int n, float *out) {
    const float x[3] = {in[0], in[1], in[2]};
    const float R[4] = {in[3], in[5], in[4], in[6]};
    for (int i = 0; i < n; i++) {
        const float *r = in + 7 + 6 * i;
        const float xf[2] = {r[0], r[1]};
        const float Pf[4] = {r[2], r[4], r[3], r[5]};
        float *o = out + 16 * i;
        jacobian_one(x, R, xf, Pf, o, o + 2, o + 12);
        hv_from_hf(o + 2, o + 6);
    }
}

void he_predict(int method, long P, float *xv, float *Pv, float V, float G, const float *Q, float dt, float phi_true,
                float wheel_base, float sigma_phi, int use_heading, const float *gauss) {
    for (long p = 0; p < P; p++) {
        float x[3], C[9];
        for (int c = 0; c < 3; c++) x[c] = xv[c * P + p];
        for (int j = 0; j < 9; j++) C[j] = Pv[j * P + p];
        float g2[2];
        if (gauss) { g2[0] = gauss[p]; g2[1] = gauss[P + p]; }
        predict_particle(method, x, C, V, G, Q, dt, phi_true, wheel_base, sigma_phi, use_heading != 0, gauss ? g2 : 0);
        for (int c = 0; c < 3; c++) xv[c * P + p] = x[c];
        if (method == 2 || use_heading)
            for (int j = 0; j < 9; j++) Pv[j * P + p] = C[j];
    }
}

// what capi.cu:flush_predicts + k_predict_batch_fast do for one noise-free tick
void he_predict_fast(int method, long P, float *xv, float *Pv, float V, float G, const float *Q, float dt,
                     float phi_true, float wheel_base, float sigma_phi, int use_heading) {
    PredictTick k;
    k.vdt = V * dt;
    k.dt = dt;
    k.dtheta = V * dt * sinf(G / wheel_base);
    k.G = G;
    k.gu20 = dt * sinf(G) / wheel_base;
    k.gu21 = V * dt * cosf(G) / wheel_base;
    for (int j = 0; j < 4; j++) k.Q[j] = Q[j];
    k.phi_true = phi_true;
    k.R_phi = sigma_phi * sigma_phi;
    for (long p = 0; p < P; p++) {
        float x[3], C[9];
        for (int c = 0; c < 3; c++) x[c] = xv[c * P + p];
        for (int j = 0; j < 9; j++) C[j] = Pv[j * P + p];
        predict_tick_fast(x, C, k, method == 2, use_heading != 0);
        for (int c = 0; c < 3; c++) xv[c * P + p] = x[c];
        if (method == 2 || use_heading)
            for (int j = 0; j < 9; j++) Pv[j * P + p] = C[j];
    }
}

int he_fs2_update_particles(long P, int Nf, float *xv, float *Pv, float *w, float *lm_mean, float *lm_cov,
                            const float *zf, const int *idf, int K, const float *zn, int Kn, const float *R,
                            const float *gauss) {
    for (long p = 0; p < P; p++) {
        float x[3], C[9];
        for (int c = 0; c < 3; c++) x[c] = xv[c * P + p];
        for (int j = 0; j < 9; j++) C[j] = Pv[j * P + p];
        const float g3[3] = {gauss[p], gauss[P + p], gauss[2 * P + p]};
        if (K > 0) {
            float x0[3] = {x[0], x[1], x[2]}, C0[9];
            for (int j = 0; j < 9; j++) C0[j] = C[j];
            for (int i = 0; i < K; i++) {
                float xf[2], Pf[4];
                lm_load(P, p, idf[i], lm_mean, lm_cov, xf, Pf);
                proposal_step(x, C, R, xf, Pf, zf + 2 * i);
            }
            float xs[3];
            multivariate_gauss<3>(x, C, g3, xs);
            const float v1[3] = {x0[0] - xs[0], x0[1] - xs[1], trig_offset(x0[2] - xs[2])};
            const float v2[3] = {x[0] - xs[0], x[1] - xs[1], trig_offset(x[2] - xs[2])};
            const float prior = gauss_evaluate<3>(v1, C0);
            const float prop = gauss_evaluate<3>(v2, C);
            float like = 1.0f;
            for (int i = 0; i < K; i++) {
                float xf[2], Pf[4];
                lm_load(P, p, idf[i], lm_mean, lm_cov, xf, Pf);
                like = like * likelihood_and_feature_update(xs, R, xf, Pf, zf + 2 * i);
                lm_store(P, p, idf[i], lm_mean, lm_cov, xf, Pf);
            }
            w[p] = w[p] * like * prior / prop;
            for (int c = 0; c < 3; c++) x[c] = xs[c];
            for (int j = 0; j < 9; j++) C[j] = 0;
        }
        if (Kn > 0) {
            if (K == 0) {
                float xs[3];
                multivariate_gauss<3>(x, C, g3, xs);
                for (int c = 0; c < 3; c++) x[c] = xs[c];
                for (int j = 0; j < 9; j++) C[j] = 0;
            }
            for (int i = 0; i < Kn; i++) {
                float xf[2], Pf[4];
                add_feature_one(x, zn[2 * i], zn[2 * i + 1], R, xf, Pf);
                lm_store(P, p, Nf + i, lm_mean, lm_cov, xf, Pf);
            }
        }
        for (int c = 0; c < 3; c++) xv[c * P + p] = x[c];
        for (int j = 0; j < 9; j++) Pv[j * P + p] = C[j];
    }
    return Nf + Kn;
}

int he_fs1_update_particles(long P, int Nf, float *xv, float *Pv, float *w, float *lm_mean, float *lm_cov,
                            const float *zf, const int *idf, int K, const float *zn, int Kn, const float *R) {
    (void) Pv;
    for (long p = 0; p < P; p++) {
        float x[3];
        for (int c = 0; c < 3; c++) x[c] = xv[c * P + p];
        if (K > 0) {
            float wt = 1.0f;
            for (int i = 0; i < K; i++) {
                float xf[2], Pf[4];
                lm_load(P, p, idf[i], lm_mean, lm_cov, xf, Pf);
                fs1_weight_and_feature_update(x, R, xf, Pf, zf + 2 * i, &wt);
                lm_store(P, p, idf[i], lm_mean, lm_cov, xf, Pf);
            }
            w[p] = w[p] * wt;
        }
        for (int i = 0; i < Kn; i++) {
            float xf[2], Pf[4];
            add_feature_one(x, zn[2 * i], zn[2 * i + 1], R, xf, Pf);
            lm_store(P, p, Nf + i, lm_mean, lm_cov, xf, Pf);
        }
    }
    return Nf + Kn;
}

int he_fs2_update_particles_fast(long P, int Nf, float *xv, float *Pv, float *w, float *lm_mean, float *lm_cov,
                                 const float *zf, const int *idf, int K, const float *R, const float *gauss) {
    for (long p = 0; p < P; p++) {
        float x[3], C[9], x0[3];
        for (int c = 0; c < 3; c++) x0[c] = x[c] = xv[c * P + p];
        for (int j = 0; j < 9; j++) C[j] = Pv[j * P + p];
        const float g3[3] = {gauss[p], gauss[P + p], gauss[2 * P + p]};
        float L0[6];
        ProposalInfo pi;
        proposal_begin_fast(C, L0, pi);
        for (int i = 0; i < K; i++) {
            float xf[2], Pf[4];
            lm_load(P, p, idf[i], lm_mean, lm_cov, xf, Pf);
            proposal_step_fast(x, pi, R, xf, Pf, zf + 2 * i);
        }
        float xs[3];
        const float ratio = proposal_finish_fast(x0, L0, x, pi, g3, xs);
        float like = 1.0f;
        for (int i = 0; i < K; i++) {
            float xf[2], Pf[4];
            lm_load(P, p, idf[i], lm_mean, lm_cov, xf, Pf);
            like = like * likelihood_and_feature_update_fast(xs, R, xf, Pf, zf + 2 * i);
            lm_store(P, p, idf[i], lm_mean, lm_cov, xf, Pf);
        }
        w[p] = w[p] * like * ratio;
        for (int c = 0; c < 3; c++) xv[c * P + p] = xs[c];
        for (int j = 0; j < 9; j++) Pv[j * P + p] = 0;
    }
    return Nf;
}

int he_fs1_update_particles_fast(long P, int Nf, float *xv, float *w, float *lm_mean, float *lm_cov, const float *zf,
                                 const int *idf, int K, const float *R) {
    for (long p = 0; p < P; p++) {
        float x[3];
        for (int c = 0; c < 3; c++) x[c] = xv[c * P + p];
        float wt = 1.0f;
        for (int i = 0; i < K; i++) {
            float xf[2], Pf[4];
            lm_load(P, p, idf[i], lm_mean, lm_cov, xf, Pf);
            wt = wt * likelihood_and_feature_update_fast(x, R, xf, Pf, zf + 2 * i);
            lm_store(P, p, idf[i], lm_mean, lm_cov, xf, Pf);
        }
        w[p] = w[p] * wt;
    }
    return Nf;
}
}
