// world.cpp — the simulated vehicle, sensor and random numbers of the host loop (reference: core.cpp:22-120, 176-273,
// 383-477, 751-777).  Every random number is drawn from std::rand() in the reference's order so that a run with the
// same SWITCH_SEED_RANDOM consumes the same stream as the reference binary.
#include <cmath>
#include <cstdlib>

#include "slam_host.hpp"

namespace slamhost {

// trigonometricOffset, core.cpp:460-477
float trig_offset(float ang) {
    if ((ang < -2 * M_PI) || (ang > 2 * M_PI)) {
        const int n = (int) floor(ang / (2 * M_PI));
        ang = (float) (ang - n * (2 * M_PI));
    }
    if (ang > M_PI) ang = (float) (ang - (2 * M_PI));
    if (ang < -M_PI) ang = (float) (ang + (2 * M_PI));
    return ang;
}

double unif_rand() { return rand() / double(RAND_MAX); }

// nRandMat::randn, core.cpp:383-419: count + 1 uniforms stored as float, Box-Muller pairs (sin first, then cos).
void randn(int count, float *out) {
    std::vector<float> u(count + 1);
    for (int r = 0; r <= count; r++) u[r] = (float) (std::rand() * 1.0 / RAND_MAX);
    float amp = 0, angle = 0;
    for (int k = 0; k < count; k++) {
        if (k % 2 == 0) {
            float square = (float) (-2. * std::log(u[k]));
            if (square < 0.) square = 0.;
            amp = std::sqrt(square);
            angle = (float) (2. * M_PI * u[k + 1]);
            out[k] = amp * std::sin(angle);
        } else {
            out[k] = amp * std::cos(angle);
        }
    }
}

// addControlNoise (core.cpp:22-31) = multivariateGauss([V; G], Q, 1): chol_L(Q) * randn(2, 1) + [V; G].
void add_control_noise(float V, float G, const float Q[4], float *Vn, float *Gn) {
    // lower Cholesky factor of the 2x2 (Eigen LLT reads the lower triangle)
    const float l00 = std::sqrt(Q[0]);
    const float l10 = Q[2] / l00;
    const float l11 = std::sqrt(Q[3] - l10 * l10);
    float g[2];
    randn(2, g);
    *Vn = (l00 * g[0] + 0.0f * g[1]) + V;
    *Gn = (l10 * g[0] + l11 * g[1]) + G;
}

// predictTruePosition, core.cpp:34-38 (note sin(G)/wheelBase here, unlike the particles' sin(G/wheelBase))
void predict_true_position(float x[3], float V, float G, float wheel_base, float dt) {
    x[0] = x[0] + V * dt * std::cos(G + x[2]);
    x[1] = x[1] + V * dt * std::sin(G + x[2]);
    x[2] = trig_offset(x[2] + V * dt * std::sin(G) / wheel_base);
}

// updateSteering, core.cpp:40-78
void update_steering(const float x[3], const Map &map, int *iwp, float minimum_distance, float *G, float max_rate,
                     float max_angle, float dt) {
    double cw[2] = {map.wp_x[*iwp], map.wp_y[*iwp]};
    const float d2 = (float) (std::pow(cw[0] - x[0], 2) + std::pow(cw[1] - x[1], 2));
    if (d2 < minimum_distance * minimum_distance) {
        (*iwp)++;
        if (*iwp >= (int) map.wp_x.size()) {
            *iwp = -1;
            return;
        }
        cw[0] = map.wp_x[*iwp];
        cw[1] = map.wp_y[*iwp];
    }
    float deltaG = (float) (std::atan2(cw[1] - x[1], cw[0] - x[0]) - x[2] - *G);
    deltaG = trig_offset(deltaG);
    const float maxDelta = max_rate * dt;
    if (std::fabs(deltaG) > maxDelta) {
        const int sign = (deltaG > 0) ? 1 : ((deltaG < 0) ? -1 : 0);
        deltaG = sign * maxDelta;
    }
    *G = *G + deltaG;
    if (std::fabs(*G) > max_angle) {
        const int sign2 = (*G > 0) ? 1 : ((*G < 0) ? -1 : 0);
        *G = sign2 * max_angle;
    }
}

// getObservations (core.cpp:185-273): findVisibleLandmarks' four tests, then range / bearing of the survivors.
void get_observations(const Map &map, const float x[3], float max_range, std::vector<int> *ids, std::vector<float> *z) {
    ids->clear();
    z->clear();
    const float phi = x[2];
    for (size_t j = 0; j < map.lm_x.size(); j++) {
        const float dx = map.lm_x[j] - x[0], dy = map.lm_y[j] - x[1];
        if ((std::fabs(dx) < max_range) && (std::fabs(dy) < max_range) &&
            ((dx * std::cos(phi) + dy * std::sin(phi)) > 0.0) &&
            ((std::pow(dx, 2) + std::pow(dy, 2)) < std::pow(max_range, 2))) {
            ids->push_back((int) j);
            z->push_back((float) std::sqrt(std::pow(dx, 2) + std::pow(dy, 2)));
            z->push_back(std::atan2(dy, dx) - phi);
        }
    }
}

// addObservationNoise, core.cpp:438-449: two randn(1, len) draws, R assumed diagonal.
void add_observation_noise(std::vector<float> *z, const float R[4]) {
    const int len = (int) z->size() / 2;
    if (len == 0) return;
    std::vector<float> r1(len), r2(len);
    randn(len, r1.data());
    randn(len, r2.data());
    for (int c = 0; c < len; c++) {
        (*z)[2 * c] = (*z)[2 * c] + r1[c] * std::sqrt(R[0]);
        (*z)[2 * c + 1] = (*z)[2 * c + 1] + r2[c] * std::sqrt(R[3]);
    }
}

// dataAssociationKnown, core.cpp:91-120
void data_association_known(const std::vector<float> &z, const std::vector<int> &idz, std::vector<int> *table, int Nf,
                            std::vector<float> *zf, std::vector<int> *idf, std::vector<float> *zn) {
    std::vector<int> idn;
    zf->clear();
    zn->clear();
    idf->clear();
    for (size_t i = 0; i < idz.size(); i++) {
        const int ii = idz[i];
        if ((*table)[ii] == -1) {
            zn->push_back(z[2 * i]);
            zn->push_back(z[2 * i + 1]);
            idn.push_back(ii);
        } else {
            zf->push_back(z[2 * i]);
            zf->push_back(z[2 * i + 1]);
            idf->push_back((*table)[ii]);
        }
    }
    for (size_t i = 0; i < idn.size(); i++) (*table)[idn[i]] = Nf + (int) i;
}

}  // namespace slamhost
