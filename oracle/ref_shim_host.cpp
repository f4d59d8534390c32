// TEST INFRASTRUCTURE ONLY — flat C windows onto the UNMODIFIED reference's host-side functions (Conf / CParamArray,
// readInputFile, updateSteering, predictTruePosition, addControlNoise, getObservations, addObservationNoise,
// dataAssociationKnown: core.cpp, utils.cpp) for tests/test_host_parity.py, which holds host/*.cpp against them.
// Same signatures as host/test_hooks.cpp so that one Python driver can call either side.
#include <cstdio>
#include <cstring>
#include <vector>

#include <Eigen/Dense>

#include "src/backend/core.h"

using namespace Eigen;
using std::vector;

extern "C" {

int slamref_conf_dump(const char *ini, const char **keys, const char **vals, int n, char *out, int cap) {
    Conf c;
    int rc = 0;
    if (ini && ini[0]) rc = c.load(ini);
    for (int i = 0; i < n; i++) c.set_s(keys[i], vals[i]);
    c.parse();
    snprintf(out, cap,
             "Vtrue=%.9g MAXG=%.9g RATEG=%.9g WHEELBASE=%.9g DT_CONTROLS=%.9g sigmaV=%.9g sigmaG=%.9g MAX_RANGE=%.9g "
             "DT_OBSERVE=%.9g sigmaR=%.9g sigmaB=%.9g sigmaT=%.9g GATE_REJECT=%.9g GATE_AUGMENT=%.9g AT_WAYPOINT=%.9g "
             "NUMBER_LOOPS=%d NPARTICLES=%d NEFFECTIVE=%d SWITCH_CONTROL_NOISE=%d SWITCH_SENSOR_NOISE=%d "
             "SWITCH_INFLATE_NOISE=%d SWITCH_PREDICT_NOISE=%d SWITCH_SAMPLE_PROPOSAL=%d SWITCH_HEADING_KNOWN=%d "
             "SWITCH_RESAMPLE=%d SWITCH_PROFILE=%d SWITCH_SEED_RANDOM=%d SWITCH_ASSOCIATION_KNOWN=%d "
             "SWITCH_BATCH_UPDATE=%d SWITCH_USE_IEKF=%d",
             c.V, c.MAXG, c.RATEG, c.WHEELBASE, c.DT_CONTROLS, c.sigmaV, c.sigmaG, c.MAX_RANGE, c.DT_OBSERVE, c.sigmaR,
             c.sigmaB, c.sigmaT, c.GATE_REJECT, c.GATE_AUGMENT, c.AT_WAYPOINT, c.NUMBER_LOOPS, c.NPARTICLES,
             c.NEFFECTIVE, c.SWITCH_CONTROL_NOISE, c.SWITCH_SENSOR_NOISE, c.SWITCH_INFLATE_NOISE,
             c.SWITCH_PREDICT_NOISE, c.SWITCH_SAMPLE_PROPOSAL, c.SWITCH_HEADING_KNOWN, c.SWITCH_RESAMPLE,
             c.SWITCH_PROFILE, c.SWITCH_SEED_RANDOM, c.SWITCH_ASSOCIATION_KNOWN, c.SWITCH_BATCH_UPDATE,
             c.SWITCH_USE_IEKF);
    return rc;
}

int slamref_read_map(const char *path, int *n_lm, int *n_wp, float *lm, float *wp, char *err, int cap) {
    MatrixXf L, W;
    readInputFile(path, &L, &W);  // exit()s on a malformed file: only call it on well-formed ones
    *n_lm = (int) L.cols();
    *n_wp = (int) W.cols();
    if (lm)
        for (int i = 0; i < L.cols(); i++) {
            lm[2 * i] = L(0, i);
            lm[2 * i + 1] = L(1, i);
        }
    if (wp)
        for (int i = 0; i < W.cols(); i++) {
            wp[2 * i] = W(0, i);
            wp[2 * i + 1] = W(1, i);
        }
    (void) err;
    (void) cap;
    return 0;
}

void slamref_add_control_noise(float V, float G, const float *Q, float *Vn, float *Gn) {
    MatrixXf Qm(2, 2);
    Qm << Q[0], Q[1], Q[2], Q[3];
    addControlNoise(V, G, Qm, *Vn, *Gn);
}
void slamref_predict_true_position(float *x, float V, float G, float wb, float dt) {
    VectorXf xv(3);
    xv << x[0], x[1], x[2];
    predictTruePosition(xv, V, G, wb, dt);
    for (int i = 0; i < 3; i++) x[i] = xv(i);
}

static MatrixXf g_lm, g_wp;
void slamref_set_map(const float *lm, int n_lm, const float *wp, int n_wp) {
    g_lm.resize(2, n_lm);
    g_wp.resize(2, n_wp);
    for (int i = 0; i < n_lm; i++) {
        g_lm(0, i) = lm[2 * i];
        g_lm(1, i) = lm[2 * i + 1];
    }
    for (int i = 0; i < n_wp; i++) {
        g_wp(0, i) = wp[2 * i];
        g_wp(1, i) = wp[2 * i + 1];
    }
}
void slamref_update_steering(const float *x, int *iwp, float min_d, float *G, float rate, float maxg, float dt) {
    VectorXf xv(3);
    xv << x[0], x[1], x[2];
    updateSteering(xv, g_wp, *iwp, min_d, *G, rate, maxg, dt);
}
int slamref_get_observations(const float *x, float max_range, int add_noise, const float *R, int *ids, float *z) {
    VectorXf xv(3);
    xv << x[0], x[1], x[2];
    vector<int> id;
    for (int i = 0; i < g_lm.cols(); i++) id.push_back(i);
    vector<VectorXf> zz = getObservations(g_lm, xv, id, max_range);
    if (add_noise) {
        MatrixXf Rm(2, 2);
        Rm << R[0], R[1], R[2], R[3];
        addObservationNoise(zz, Rm);
    }
    for (size_t i = 0; i < id.size(); i++) {
        ids[i] = id[i];
        z[2 * i] = zz[i](0);
        z[2 * i + 1] = zz[i](1);
    }
    return (int) id.size();
}
int slamref_data_association_known(const float *z, const int *idz, int m, int *table, int ntab, int Nf, float *zf,
                                   int *idf, float *zn, int *Kn) {
    vector<VectorXf> zz, zfv, znv;
    vector<int> iv(idz, idz + m), idfv;
    for (int i = 0; i < m; i++) {
        VectorXf v(2);
        v << z[2 * i], z[2 * i + 1];
        zz.push_back(v);
    }
    VectorXf tab(ntab);  // the reference keeps the table as floats (ParticleSLAMWrapper.cpp:27-32)
    for (int i = 0; i < ntab; i++) tab(i) = (float) table[i];
    dataAssociationKnown(zz, iv, tab, Nf, zfv, idfv, znv);
    for (int i = 0; i < ntab; i++) table[i] = (int) tab(i);
    for (size_t i = 0; i < zfv.size(); i++) {
        zf[2 * i] = zfv[i](0);
        zf[2 * i + 1] = zfv[i](1);
        idf[i] = idfv[i];
    }
    for (size_t i = 0; i < znv.size(); i++) {
        zn[2 * i] = znv[i](0);
        zn[2 * i + 1] = znv[i](1);
    }
    *Kn = (int) znv.size();
    return (int) idfv.size();
}

}  // extern "C"
