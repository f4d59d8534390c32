// Builds against include/slamgpu_fastslam.hpp + libslamgpu.so like a C++ host would.  Without a CUDA device the
// constructor path must throw (no CPU fallback); with one it runs a tiny FastSLAM2 cycle.
#include <cmath>
#include <cstdio>
#include <vector>

#include "slamgpu_fastslam.hpp"

struct Mat2 {
    float m[4];
    float operator()(int r, int c) const { return m[r * 2 + c]; }
};
struct Vec3 {
    float d[3];
    float operator()(int i) const { return d[i]; }
};

int main() {
    try {
        slamb200::FastSLAM2 slam(256, 8);
        slam.useHeading = true;
        slam.nEffective = 192;
        slam.sigmaPhi = 0.0175f;
        slam.initializeParticles();
        const Mat2 Q{{0.09f, 0, 0, 0.00274f}}, R{{0.01f, 0, 0, 0.0003f}};
        const Vec3 xTrue{{0, 0, 0}};
        std::vector<float> gauss(3 * 256, 0.1f);
        std::vector<double> unif(256, 0.5);
        for (int t = 0; t < 8; t++) slam.predict(xTrue, 3.0f, 0.01f, Q, 0.025f);
        std::vector<slamb200::Vec2> zf, zn{{{10.0f, 0.3f}}, {{7.0f, -0.5f}}};
        std::vector<int> idf;
        slam.update(zf, zn, idf, R, gauss.data(), unif.data());
        for (int t = 0; t < 8; t++) slam.predict(xTrue, 3.0f, 0.01f, Q, 0.025f);
        zf = {{{9.4f, 0.31f}}, {{6.5f, -0.55f}}};
        zn.clear();
        idf = {0, 1};
        bool did = false;
        const float neff = slam.update(zf, zn, idf, R, gauss.data(), unif.data(), &did);
        double x, y, th;
        slam.computeEstimatedPosition(x, y, th);
        if (!(neff > 0) || !std::isfinite(x) || slam.landmarkCount() != 2) {
            std::printf("HOST_CPP_BAD neff=%f x=%f\n", neff, x);
            return 2;
        }
        std::printf("HOST_CPP_OK neff=%.3f x=%.4f y=%.4f landmarks=%d\n", neff, x, y, slam.landmarkCount());
        return 0;
    } catch (const std::exception &e) {
        std::printf("HOST_CPP_THROW %s\n", e.what());
        return 3;
    }
}
