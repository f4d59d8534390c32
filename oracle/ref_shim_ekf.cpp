// TEST INFRASTRUCTURE ONLY — not part of the product path.
//
// Flat-array C shim over the UNMODIFIED reference EKF-SLAM (src/backend/algorithms/ekfslam.cpp, compiled in place
// from /root/reference by oracle/Makefile into oracle/_ref/libslamref.so).  It drives EKFSLAM::sim
// (ekfslam.cpp:18-43) exactly as wrappers/ekfslamwrapper.cpp:80-83 does — zf / idf / zn / dataAssociationTable are
// passed BY VALUE there (ekfslam.h:20-23), so to let the tests see the association the same predict -> observeHeading
// -> dataAssociate sequence is replayed on a copy of the state through a Probe subclass (the members are protected).
//
// Matrices cross this boundary row-major.
#include <vector>

#include <Eigen/Dense>

#include "src/backend/core.h"
#include "src/backend/algorithms/ekfslam.h"

using namespace Eigen;
using std::vector;

namespace {

struct ProbeEkf : public EKFSLAM {
    using EKFSLAM::augment;
    using EKFSLAM::batchUpdate;
    using EKFSLAM::dataAssociate;
    using EKFSLAM::dataAssociateKnown;
    using EKFSLAM::ekfComputeAssociation;
    using EKFSLAM::observeHeading;
    using EKFSLAM::predict;
};

struct RefEkf {
    ProbeEkf a;
    VectorXf x;
    MatrixXf P;
};

MatrixXf m2(const float *m) {
    MatrixXf M(2, 2);
    M << m[0], m[1], m[2], m[3];
    return M;
}

}  // namespace

extern "C" {

void *slamref_ekf_create(float wheelBase, float sigmaPhi, float gateReject, float gateAugment, int useHeading,
                         int batchUpdate, int associationKnown) {
    RefEkf *e = new RefEkf();
    e->a.wheelBase = wheelBase;
    e->a.sigmaPhi = sigmaPhi;
    e->a.gateReject = gateReject;
    e->a.gateAugment = gateAugment;
    e->a.useHeading = useHeading != 0;
    e->a.enableBatchUpdate = batchUpdate != 0;
    e->a.associationKnown = associationKnown;
    e->x = VectorXf::Zero(3);          // slamwrapper.cpp: xEstimated starts at zero, P = 0 (ekfslamwrapper.cpp:40-41)
    e->P = MatrixXf::Zero(3, 3);
    return e;
}

void slamref_ekf_destroy(void *h) { delete (RefEkf *) h; }

int slamref_ekf_size(void *h) { return (int) ((RefEkf *) h)->x.size(); }

void slamref_ekf_set(void *h, int n, const float *x, const float *P) {
    RefEkf *e = (RefEkf *) h;
    e->x.resize(n);
    e->P.resize(n, n);
    for (int i = 0; i < n; i++) {
        e->x(i) = x[i];
        for (int j = 0; j < n; j++) e->P(i, j) = P[i * n + j];
    }
}

void slamref_ekf_get(void *h, float *x, float *P) {
    RefEkf *e = (RefEkf *) h;
    const int n = (int) e->x.size();
    for (int i = 0; i < n; i++) {
        x[i] = e->x(i);
        for (int j = 0; j < n; j++) P[i * n + j] = e->P(i, j);
    }
}

// One EKFSLAM::sim call.  z [m][2] = landmarksRangeBearing of the visible landmarks, idz [m] their identifiers,
// table [ntab] the wrapper's dataAssociationTable (read-only: by value in the reference).
// Out: idf_out [<= m] and the zf / zn counts of this call's association (0 / 0 when !observe).
void slamref_ekf_sim(void *h, float V, float G, const float *Qe, float dt, float phi, int observe, const float *z,
                     const int *idz, int m, const int *table, int ntab, const float *Re, const float *R, int *idf_out,
                     int *n_zf, int *n_zn) {
    RefEkf *e = (RefEkf *) h;
    MatrixXf Q = m2(Qe), Rem = m2(Re), Rm = m2(R);
    vector<VectorXf> lrb;
    vector<int> ids(idz ? idz : nullptr, idz ? idz + m : nullptr);
    for (int i = 0; i < m; i++) {
        VectorXf v(2);
        v << z[2 * i], z[2 * i + 1];
        lrb.push_back(v);
    }
    vector<int> tab(table ? table : nullptr, table ? table + ntab : nullptr);
    MatrixXf lm, wp;  // unused by sim
    vector<VectorXf> zf, zn;
    vector<int> idf;
    *n_zf = *n_zn = 0;
    if (observe) {  // the association this call will make, replayed on a copy
        VectorXf xc = e->x;
        MatrixXf Pc = e->P;
        e->a.predict(xc, Pc, V, G, Q, e->a.wheelBase, dt);
        if (e->a.useHeading) e->a.observeHeading(xc, Pc, phi);
        vector<VectorXf> zf2, zn2;
        vector<int> idf2, tab2 = tab;
        if (e->a.associationKnown) e->a.dataAssociateKnown(xc, lrb, ids, zf2, idf2, zn2, tab2);
        else e->a.dataAssociate(xc, Pc, lrb, Rem, e->a.gateReject, e->a.gateAugment, zf2, idf2, zn2);
        *n_zf = (int) zf2.size();
        *n_zn = (int) zn2.size();
        for (size_t i = 0; i < idf2.size(); i++) idf_out[i] = idf2[i];
    }
    e->a.sim(lm, wp, e->x, e->P, V, G, Q, dt, phi, ids, lrb, Rem, observe != 0, zf, idf, zn, tab, Rm);
}

// nis / nd of one (observation, feature) pair: ekfComputeAssociation, ekfslam.cpp:133-149.
void slamref_ekf_association(void *h, const float *z, const float *R, int idf, float *nis, float *nd) {
    RefEkf *e = (RefEkf *) h;
    VectorXf zz(2);
    zz << z[0], z[1];
    MatrixXf Rm = m2(R);
    e->a.ekfComputeAssociation(e->x, e->P, zz, Rm, idf, *nis, *nd);
}

}  // extern "C"
