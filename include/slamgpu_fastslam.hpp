// slamgpu_fastslam.hpp — C++ host mirror of the reference's FastSLAM method objects over the C ABI (slamgpu.h).
//
// The reference declares (src/backend/algorithms/fastslam2.h:14-31, fastslam1.h:14-33):
//     class FastSLAM2 { void predict(vector<Particle>&, VectorXf& xTrue, float V, float G, MatrixXf& Q, float dt);
//                       void update(vector<Particle>&, vector<VectorXf>& zf, vector<VectorXf>& zn, vector<int>& idf,
//                                   vector<VectorXf>& z, VectorXf& dataAssociationTable, MatrixXf& R);
//                       bool addPredictNoise, useHeading; float wheelBase; int nEffective; bool resample; float sigmaPhi; };
// Here `vector<Particle>& particles` is the device-resident set owned by the object, and the noise the reference
// draws from rand() inside those calls is passed in (nullptr = noise staged earlier with stageNoise()).  Vector
// arguments are templates over anything with Eigen's element access `v(i)` / `m(r,c)`, so the reference's call sites
// (wrappers/fastslam2wrapper.cpp:64,88) compile against this class unchanged apart from the particle argument.
// Errors throw std::runtime_error like AcceleratorHandler's constructor does (AcceleratorHandler.cpp:20-28).
#pragma once

#include <stdexcept>
#include <string>
#include <vector>

#include "slamgpu.h"

namespace slamb200 {

struct Vec2 {  // minimal stand-in where Eigen is not at hand
    float d[2];
    float operator()(int i) const { return d[i]; }
};

class FastSLAM {
public:
    // public switches with the reference's names (fastslam2.h:26-31)
    bool addPredictNoise = false;
    bool useHeading = false;
    float wheelBase = 4.0f;
    int nEffective = 75;
    bool resample = true;
    float sigmaPhi = 0.0f;

    FastSLAM(int method, long nParticles, int maxLandmarks) : method_(method), n_(nParticles), cap_(maxLandmarks) {}
    ~FastSLAM() {
        if (h_) slamgpu_destroy(h_);
    }
    FastSLAM(const FastSLAM &) = delete;
    FastSLAM &operator=(const FastSLAM &) = delete;

    // Allocates the device state from the current switch values (call once, after setting them like
    // FastSLAM2Wrapper's constructor does, fastslam2wrapper.cpp:18-23).
    void initializeParticles(int device = 0, int scanMode = SLAMGPU_SCAN_BLOCKED) {
        slamgpu_config c{};
        c.n_particles = n_;
        c.nf_cap = cap_;
        c.method = method_;
        c.wheel_base = wheelBase;
        c.sigma_phi = sigmaPhi;
        c.n_effective = nEffective;
        c.use_heading = useHeading;
        c.do_resample = resample;
        c.scan_mode = scanMode;
        c.device = device;
        check(slamgpu_create(&c, &h_));
    }

    void stageNoise(const float *gauss, int rows, const double *unif) { check(slamgpu_stage_noise(h_, gauss, rows, unif)); }

    // predict(particles, xTrue, V, G, Q, dt): gauss = [2][N] standard normals when addPredictNoise, else nullptr.
    template <class Vec, class Mat>
    void predict(const Vec &xTrue, float V, float G, const Mat &Q, float dt, const float *gauss = nullptr) {
        const float q[4] = {Q(0, 0), Q(0, 1), Q(1, 0), Q(1, 1)};
        check(slamgpu_predict(h_, V, G, q, dt, xTrue(2), addPredictNoise ? gauss : nullptr));
    }

    // update(particles, zf, zn, idf, z, dataAssociationTable, R): z and the table are not used by the reference's
    // update either (fastslam2.cpp:21-48).  gauss = [3][N] (FastSLAM2), unif = [N] doubles; returns nEff.
    template <class Vec, class Mat>
    float update(const std::vector<Vec> &zf, const std::vector<Vec> &zn, const std::vector<int> &idf, const Mat &R,
                 const float *gauss = nullptr, const double *unif = nullptr, bool *didResample = nullptr) {
        std::vector<float> f(2 * zf.size()), n(2 * zn.size());
        for (size_t i = 0; i < zf.size(); i++) { f[2 * i] = zf[i](0); f[2 * i + 1] = zf[i](1); }
        for (size_t i = 0; i < zn.size(); i++) { n[2 * i] = zn[i](0); n[2 * i + 1] = zn[i](1); }
        const float r[4] = {R(0, 0), R(0, 1), R(1, 0), R(1, 1)};
        int did = 0;
        float neff = 0;
        check(slamgpu_update(h_, f.data(), idf.data(), (int) zf.size(), n.data(), (int) zn.size(), r, gauss, unif, &did,
                             &neff));
        if (didResample) *didResample = did != 0;
        return neff;
    }

    // ParticleSLAMWrapper::computeEstimatedPosition (ParticleSLAMWrapper.cpp:56-77)
    void computeEstimatedPosition(double &x, double &y, double &t) { check(slamgpu_estimate(h_, &x, &y, &t)); }
    int landmarkCount() const { return slamgpu_landmark_count(h_); }
    slamgpu_t *handle() { return h_; }

private:
    void check(int rc) {
        if (rc != SLAMGPU_OK) throw std::runtime_error(std::string("slamgpu: ") + slamgpu_last_error());
    }
    int method_;
    long n_;
    int cap_;
    slamgpu_t *h_ = nullptr;
};

struct FastSLAM1 : FastSLAM {
    FastSLAM1(long n, int cap) : FastSLAM(SLAMGPU_FASTSLAM1, n, cap) { addPredictNoise = true; }  // fastslam1wrapper.cpp:20
};
struct FastSLAM2 : FastSLAM {
    FastSLAM2(long n, int cap) : FastSLAM(SLAMGPU_FASTSLAM2, n, cap) {}
};

// computeJacobians through the accelerator record (core.cpp:586-664): the literal AcceleratorHandler replacement.
inline void computeJacobiansRecord(const float *in, int n, float *out) {
    if (slamgpu_compute_jacobians(nullptr, in, n, out) != SLAMGPU_OK)
        throw std::runtime_error(std::string("slamgpu: ") + slamgpu_last_error());
}

}  // namespace slamb200
